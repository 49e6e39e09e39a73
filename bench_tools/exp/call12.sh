timeout 280 python -X faulthandler -m pytest tests/test_device_door_gpu.py tests/test_ps_kernels.py -x -q --timeout 120 > gpurun_out/door_test.log 2>&1
tail -30 gpurun_out/door_test.log
