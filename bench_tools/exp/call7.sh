python -m pytest tests/test_ps_kernels.py tests/test_engine_gpu.py -x -q 2>&1 | tail -8
python bench.py --steps 100 --warmup 10 > gpurun_out/r2_bench4.json 2> gpurun_out/r2_bench4.err; echo "bench rc $?"; cat gpurun_out/r2_bench4.json | cut -c1-300; tail -3 gpurun_out/r2_bench4.err
