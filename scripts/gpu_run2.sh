mkdir -p gpurun_out
T=${1:-x}
timeout 120 python -m pytest tests/test_gpu_tc.py -x -q -m gpu -s -k "weight_gradients or linear_backward" 2>&1 | grep -v "^$" | tail -60 > gpurun_out/${T}_tn_tests.log
grep -h "dW\|db\|passed\|failed\|Error\|error" gpurun_out/${T}_tn_tests.log | cut -c1-160 | head -40
