set -x
mkdir -p gpurun_out/mgpu
timeout 200 python -m pytest tests/test_multi_gpu.py -q -k "ipc and 2" --timeout 150 > gpurun_out/mgpu/tests2b.log 2>&1; tail -5 gpurun_out/mgpu/tests2b.log
TR="timeout 150 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 2 --master-port 29802 bench.py --gpus 2 --steps 200 --warmup 10 --mode async > gpurun_out/mgpu/bench_2gpu_async.json 2> gpurun_out/mgpu/bench_2gpu_async.err; grep '^{' gpurun_out/mgpu/bench_2gpu_async.json | cut -c1-330
$TR --nproc-per-node 2 --master-port 29801 bench.py --gpus 2 --steps 200 --warmup 10 > gpurun_out/mgpu/bench_2gpu_sync.json 2> gpurun_out/mgpu/bench_2gpu_sync.err; grep '^{' gpurun_out/mgpu/bench_2gpu_sync.json | cut -c1-330
