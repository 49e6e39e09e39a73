# N-GPU call (N = 4 or 8): multi-rank tests at N, bench sync/async, sweep, DDP baseline.  Everything bounded.
N=$1
set -x
mkdir -p gpurun_out/mgpu
timeout 300 python -m pytest tests/test_multi_gpu.py -x -q -k "ipc and $N" --timeout 140 -s > gpurun_out/mgpu/tests$N.log 2>&1; tail -8 gpurun_out/mgpu/tests$N.log
TR="timeout 150 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node $N --master-port 29811 bench.py --gpus $N --steps 200 --warmup 10 > gpurun_out/mgpu/bench_${N}gpu_sync.json 2> gpurun_out/mgpu/bench_${N}gpu_sync.err; cut -c1-330 gpurun_out/mgpu/bench_${N}gpu_sync.json
$TR --nproc-per-node $N --master-port 29812 bench.py --gpus $N --steps 200 --warmup 10 --mode async > gpurun_out/mgpu/bench_${N}gpu_async.json 2> gpurun_out/mgpu/bench_${N}gpu_async.err; cut -c1-330 gpurun_out/mgpu/bench_${N}gpu_async.json
$TR --nproc-per-node $N --master-port 29813 bench_tools/pushpull_sweep.py --out gpurun_out/mgpu/pushpull_sweep_${N}gpu.json > gpurun_out/mgpu/sweep$N.log 2>&1; tail -3 gpurun_out/mgpu/sweep$N.log
$TR --nproc-per-node $N --master-port 29814 bench_tools/ddp_baseline.py --graph --steps 100 > gpurun_out/mgpu/ddp_${N}gpu.json 2> gpurun_out/mgpu/ddp_${N}gpu.err; cat gpurun_out/mgpu/ddp_${N}gpu.json; tail -2 gpurun_out/mgpu/ddp_${N}gpu.err
if [ "$N" = "8" ]; then
  timeout 120 python bench.py --gpus 1 --steps 200 --warmup 10 > gpurun_out/mgpu/bench_1gpu_sync_samebox.json 2>/dev/null; cut -c1-200 gpurun_out/mgpu/bench_1gpu_sync_samebox.json
fi
