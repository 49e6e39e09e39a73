set -x
mkdir -p gpurun_out/mgpu
timeout 420 python -m pytest tests/test_multi_gpu.py -q -k "(ipc and 2) or death" --timeout 200 -s > gpurun_out/mgpu/tests2.log 2>&1; tail -15 gpurun_out/mgpu/tests2.log
