set -x
timeout 200 python -m pytest tests/test_bn_head_gpu.py tests/test_engine_gpu.py tests/test_zz_no_alloc.py -q --timeout 200 2>&1 | tail -3
timeout 120 python bench.py --steps 200 --warmup 10 2>/dev/null | python -c "import sys,json; d=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print(d['ms_per_step'], d['e2e']['ms_per_step'], d['clocks'])"
