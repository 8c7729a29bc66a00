mkdir -p gpurun_out
python bench.py --steps 2 --warmup 3 --no-rmat --no-rwr --no-rmat20 --no-cpu-baseline > gpurun_out/ll_bench.json 2> gpurun_out/ll_bench.err || exit 1
ncu --metrics gpu__time_duration.sum,sm__inst_executed_pipe_tensor.sum,sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed --clock-control none -c 900 --csv --log-file gpurun_out/launches_${1:-x}.csv python bench.py --steps 1 --warmup 3 --no-rmat --no-rwr --no-rmat20 --no-cpu-baseline > gpurun_out/ncu_ll.log 2>&1
tail -3 gpurun_out/ncu_ll.log
