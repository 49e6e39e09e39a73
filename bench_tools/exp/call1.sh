set -x
nvidia-smi -L | head -2
./bench_tools/bin/ingest_probe > gpurun_out/ingest_probe.log 2>&1; echo "probe rc $?"
python bench_tools/exp/cublas_shapes.py > gpurun_out/cublas_shapes.log 2>&1; echo "cublas rc $?"
python bench.py --steps 100 --warmup 10 > gpurun_out/r2_bench0.json 2> gpurun_out/r2_bench0.err; echo "bench rc $?"
tail -3 gpurun_out/ingest_probe.log; cat gpurun_out/cublas_shapes.log; cat gpurun_out/r2_bench0.json
