python -m pytest tests/test_conv_gpu.py tests/test_engine_gpu.py -x -q 2>&1 | tail -5
python bench.py --steps 100 --warmup 10 > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; echo "bench rc $?"; cat gpurun_out/r2_bench1.json | cut -c1-400
python bench_tools/step_timeline.py > gpurun_out/timeline_r2_1.log 2>&1; head -16 gpurun_out/timeline_r2_1.log
