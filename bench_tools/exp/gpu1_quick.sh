set -x
timeout 300 python -m pytest tests/test_conv_gpu.py -x -q --timeout 200 -k "training_batch" 2>&1 | tail -6
timeout 200 python -m pytest tests/test_convergence_gpu.py -x -q --timeout 200 -s 2>&1 | grep -E "^step|passed|failed|Error" | head -20
