# 1-GPU call: whole GPU test suite, headline bench, one full step under ncu --set full
set -x
timeout 600 python -m pytest tests -x -q -m gpu --timeout 300 -s > gpurun_out/gputests.log 2>&1; tail -25 gpurun_out/gputests.log
timeout 200 python bench.py --steps 200 --warmup 10 > gpurun_out/bench_1gpu_sync.json 2> gpurun_out/bench_1gpu_sync.err; cut -c1-260 gpurun_out/bench_1gpu_sync.json
timeout 120 python bench.py --steps 2 --warmup 3 --no-graph --no-e2e > gpurun_out/ncu_plain.log 2>&1 && \
timeout 500 ncu --set full --clock-control none --import-source on -k regex:'conv|bn_|head|stem|push|pull|wait|bump|gate' \
    --launch-skip 372 --launch-count 128 -f -o gpurun_out/prof_step_r2 \
    python bench.py --steps 2 --warmup 3 --no-graph --no-e2e > gpurun_out/ncu_step.log 2>&1
tail -3 gpurun_out/ncu_step.log; ls -la gpurun_out/prof_step_r2.ncu-rep
