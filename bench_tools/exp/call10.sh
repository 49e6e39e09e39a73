python -m pytest tests/test_engine_gpu.py tests/test_bn_head_gpu.py tests/test_zz_no_alloc.py -x -q 2>&1 | tail -4
python bench.py --steps 200 --warmup 10 > gpurun_out/r2_bench6.json 2> gpurun_out/r2_bench6.err; echo "bench rc $?"; cat gpurun_out/r2_bench6.json | cut -c1-300; tail -3 gpurun_out/r2_bench6.err
