timeout 300 python -m pytest tests/test_device_door_gpu.py tests/test_ps_kernels.py -x -q 2>&1 | tail -15
python bench.py --steps 100 --warmup 10 --no-e2e 2>/dev/null | cut -c1-260
