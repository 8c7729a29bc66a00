mkdir -p gpurun_out
T=${1:-x}
timeout 300 python -m pytest tests/test_gpu_tc.py -x -q -m gpu -s 2>&1 | grep -v "^$" | tail -80 > gpurun_out/${T}_tc_tests.log
grep -h "tc gemm\|passed\|failed\|dW\|dX\|db\|layer fwd\|Error" gpurun_out/${T}_tc_tests.log | cut -c1-150
timeout 300 python -m pytest tests/test_gpu_gcn.py tests/test_gpu_msi.py tests/test_gpu_scale.py -x -q -m gpu 2>&1 | tail -4
timeout 200 python profiles/tc_pipeline_trace.py gpurun_out/${T}_tc_trace.json > gpurun_out/${T}_tc_trace.log 2>&1; grep "median\|epilogue tile\|interval" gpurun_out/${T}_tc_trace.log | head -8
timeout 300 python bench.py --steps 20 --warmup 5 --no-rmat --no-rwr --no-rmat20 --no-cpu-baseline > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; python - <<PY
import json;d=json.loads(open('gpurun_out/${T}_bench.json').read().strip().splitlines()[-1]);print('epochs/s',d['value'],'no-hoist',d['config'].get('epochs_per_s_without_hoist'),'e2e',d['e2e']['value'])
PY
