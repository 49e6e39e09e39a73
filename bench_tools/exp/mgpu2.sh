# 2-GPU validation call: multi-rank tests, recovery, NVLink ncu capture, bench, sweep, DDP baseline.  Everything bounded.
set -x
mkdir -p gpurun_out/mgpu
timeout 420 python -m pytest tests/test_multi_gpu.py -x -q -k "(ipc and 2) or death" --timeout 200 -s > gpurun_out/mgpu/tests2.log 2>&1; tail -15 gpurun_out/mgpu/tests2.log
timeout 120 python bench_tools/nvlink_capture.py > gpurun_out/mgpu/nvlink_capture.json 2> gpurun_out/mgpu/nvlink_capture.err && \
timeout 300 ncu --set full --metrics nvltx__bytes.sum,nvlrx__bytes.sum,nvltx__bytes_data_user.sum,nvlrx__bytes_data_user.sum --clock-control none --import-source on \
    -k regex:'push_sync_kernel|pull_seg_kernel' --launch-skip 11 --launch-count 4 -f -o gpurun_out/mgpu/prof_nvlink \
    python bench_tools/nvlink_capture.py > gpurun_out/mgpu/ncu_nvlink.log 2>&1
cat gpurun_out/mgpu/nvlink_capture.json; tail -3 gpurun_out/mgpu/ncu_nvlink.log
TR="timeout 180 python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1"
$TR --nproc-per-node 2 --master-port 29801 bench.py --gpus 2 --steps 200 --warmup 10 > gpurun_out/mgpu/bench_2gpu_sync.json 2> gpurun_out/mgpu/bench_2gpu_sync.err; cut -c1-330 gpurun_out/mgpu/bench_2gpu_sync.json
$TR --nproc-per-node 2 --master-port 29802 bench.py --gpus 2 --steps 200 --warmup 10 --mode async > gpurun_out/mgpu/bench_2gpu_async.json 2> gpurun_out/mgpu/bench_2gpu_async.err; cut -c1-330 gpurun_out/mgpu/bench_2gpu_async.json
timeout 120 python bench.py --gpus 1 --steps 200 --warmup 10 --mode async > gpurun_out/mgpu/bench_1gpu_async.json 2>/dev/null; cut -c1-330 gpurun_out/mgpu/bench_1gpu_async.json
$TR --nproc-per-node 2 --master-port 29803 bench_tools/pushpull_sweep.py --out gpurun_out/mgpu/pushpull_sweep_2gpu.json > gpurun_out/mgpu/sweep2.log 2>&1; tail -3 gpurun_out/mgpu/sweep2.log
timeout 120 python bench_tools/ddp_baseline.py --graph --steps 100 > gpurun_out/mgpu/ddp_1gpu.json 2> gpurun_out/mgpu/ddp_1gpu.err; cat gpurun_out/mgpu/ddp_1gpu.json
$TR --nproc-per-node 2 --master-port 29804 bench_tools/ddp_baseline.py --graph --steps 100 > gpurun_out/mgpu/ddp_2gpu.json 2> gpurun_out/mgpu/ddp_2gpu.err; cat gpurun_out/mgpu/ddp_2gpu.json; tail -2 gpurun_out/mgpu/ddp_2gpu.err
