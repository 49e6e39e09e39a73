# 2-GPU call, short: NVLS reduction of the sync push (sources stage locally, owners multimem.ld_reduce) -- correctness
# through the public API, then the push as a stand-alone transfer.  Everything bounded.
set -x
mkdir -p gpurun_out/nvls
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node 2"
DPS_NVLS=on DPS_NVLS_REDUCE=1 timeout 90 $TR --master-port 29821 bench_tools/mgpu_check.py sync gpurun_out/nvls/check_sync_reduce_2.json 10 > gpurun_out/nvls/check_sync_reduce_2.log 2>&1
cut -c1-900 gpurun_out/nvls/check_sync_reduce_2.json || tail -20 gpurun_out/nvls/check_sync_reduce_2.log
grep -q '"nvls_reduce": true' gpurun_out/nvls/check_sync_reduce_2.json || { echo "NVLS REDUCE NOT ACTIVE"; tail -30 gpurun_out/nvls/check_sync_reduce_2.log; exit 0; }
timeout 70 $TR --master-port 29822 bench_tools/pushpull_sweep.py --nvls on --nvls-reduce --no-nccl --iters 10 --kb 1024,16384,65536 --out gpurun_out/nvls/sweep_2gpu_reduce.json > gpurun_out/nvls/sweep_reduce.log 2>&1; grep payload gpurun_out/nvls/sweep_reduce.log | cut -c1-220
