python -m pytest tests/test_bn_head_gpu.py tests/test_engine_gpu.py -x -q 2>&1 | tail -8
python bench.py --steps 100 --warmup 10 > gpurun_out/r2_bench5.json 2> gpurun_out/r2_bench5.err; echo "bench rc $?"; cat gpurun_out/r2_bench5.json | cut -c1-300; tail -3 gpurun_out/r2_bench5.err
python bench_tools/step_timeline.py > gpurun_out/timeline_r2_4.log 2>&1; head -16 gpurun_out/timeline_r2_4.log
