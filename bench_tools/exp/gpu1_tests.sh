# 1-GPU call: whole GPU test suite + smoke() + headline bench, bounded
set -x
timeout 420 python -m pytest tests -q -m gpu --timeout 300 > gpurun_out/gputests_final.log 2>&1; tail -5 gpurun_out/gputests_final.log
timeout 100 python -c "import __graft_entry__ as g; g.smoke(); print('SMOKE OK')" 2>&1 | tail -2
timeout 120 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_1gpu_sync.json 2> gpurun_out/bench_1gpu_sync.err; grep '^{' gpurun_out/bench_1gpu_sync.json | cut -c1-260
timeout 100 python bench.py --steps 200 --warmup 10 --mode async > gpurun_out/bench_1gpu_async.json 2>/dev/null; grep '^{' gpurun_out/bench_1gpu_async.json | cut -c1-200
