# final 1-GPU call of the round: whole GPU test suite, headline bench, launch list of one step, 224x224 configuration
set -x
timeout 400 python -m pytest tests -q -m gpu --timeout 300 > gpurun_out/gputests_final.log 2>&1; tail -4 gpurun_out/gputests_final.log
timeout 120 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_1gpu_sync.json 2> gpurun_out/bench_1gpu_sync.err; grep '^{' gpurun_out/bench_1gpu_sync.json | cut -c1-200
timeout 120 python bench.py --steps 200 --warmup 10 --mode async > gpurun_out/bench_1gpu_async.json 2>/dev/null; grep '^{' gpurun_out/bench_1gpu_async.json | cut -c1-200
timeout 150 python bench.py --steps 10 --warmup 3 --image-size 224 --no-e2e > gpurun_out/bench_1gpu_224.json 2> gpurun_out/bench_1gpu_224.err; grep '^{' gpurun_out/bench_1gpu_224.json | cut -c1-200; tail -2 gpurun_out/bench_1gpu_224.err
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-e2e"
timeout 250 ncu --metrics gpu__time_duration.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active,lts__throughput.avg.pct_of_peak_sustained_elapsed,dram__throughput.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:'conv|grad|bn_|head|stem|push|pull|wait|bump|gate' --launch-skip 485 --launch-count 125 --csv --log-file gpurun_out/launches_v6.csv $CMD > gpurun_out/ncu_list.log 2>&1
tail -2 gpurun_out/ncu_list.log
