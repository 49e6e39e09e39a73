# 2-GPU NVLS call: correctness of the multimem.st pull (replicas + fp32 prefix identical on every rank, first sync step
# equal to the single-process mean step), then the same-box A/B of pull time and step time.  Everything bounded.
set -x
mkdir -p gpurun_out/nvls
N=${1:-2}
TR="python -m torch.distributed.run --nnodes=1 --master-addr 127.0.0.1 --nproc-per-node $N"
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT DPS_NVLS=auto timeout 150 $TR --master-port 29811 bench_tools/mgpu_check.py sync gpurun_out/nvls/check_sync_$N.json 20 > gpurun_out/nvls/check_sync_$N.log 2>&1
grep -i "nvls" gpurun_out/nvls/check_sync_$N.log | head -5; grep -i "NVLS pull not available\|Error\|error" gpurun_out/nvls/check_sync_$N.log | head -5
cat gpurun_out/nvls/check_sync_$N.json | cut -c1-600
grep -q '"nvls_pull": true' gpurun_out/nvls/check_sync_$N.json || { echo "NVLS NOT ACTIVE"; tail -30 gpurun_out/nvls/check_sync_$N.log; exit 0; }
for m in on off; do
timeout 100 $TR --master-port 29812 bench_tools/pushpull_sweep.py --nvls $m --no-nccl --kb 64,1024,16384,65536 --out gpurun_out/nvls/sweep_${N}gpu_$m.json > gpurun_out/nvls/sweep_$m.log 2>&1; grep payload gpurun_out/nvls/sweep_$m.log | cut -c1-200
done
for m in on off on off; do
timeout 120 $TR --master-port 29813 bench.py --gpus $N --steps 200 --warmup 10 --nvls $m 2> gpurun_out/nvls/bench_$m.err | grep '^{' > gpurun_out/nvls/bench_${N}gpu_nvls_$m.json; python -c "import json;d=json.load(open('gpurun_out/nvls/bench_${N}gpu_nvls_$m.json'));print('nvls=$m',d['ms_per_step'],d['e2e']['ms_per_step'],d['config']['nvls_pull'])"
done
DPS_NVLS=on timeout 150 $TR --master-port 29814 bench_tools/mgpu_check.py async gpurun_out/nvls/check_async_$N.json 20 > gpurun_out/nvls/check_async_$N.log 2>&1; cut -c1-400 gpurun_out/nvls/check_async_$N.json
