#!/bin/bash
# Round-3 ncu captures with this round's binary (each target is first run once WITHOUT ncu and must exit 0).
set -u
mkdir -p gpurun_out
run() {  # name, python args...
  local name=$1; shift
  python "$@" > gpurun_out/${name}_plain.log 2>&1 || { echo "$name: plain run failed"; tail -5 gpurun_out/${name}_plain.log; return 1; }
  timeout 900 ncu --set full --clock-control none --import-source on --profile-from-start off -f -o gpurun_out/${name} \
      python "$@" > gpurun_out/${name}_ncu.log 2>&1 || { echo "$name: ncu run failed"; tail -5 gpurun_out/${name}_ncu.log; return 1; }
  ncu -i gpurun_out/${name}.ncu-rep --page raw --csv > gpurun_out/${name}_full_raw.csv 2> /dev/null
  ls -la gpurun_out/${name}.ncu-rep gpurun_out/${name}_full_raw.csv
  rm -f gpurun_out/${name}.ncu-rep          # gpurun_out/ travels back only below 64 MiB: keep the CSV export
}
run spmm_rmat24_r3 profiles/profile_iteration.py rmat 24 128
run msi_iteration_r3 profiles/profile_iteration.py msi
run gram_topk_rmat20_r3 profiles/ncu_gram_topk.py 1048576 256
# launch list of the default bench step (share of the step per kernel)
ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/launches_msi_iteration_r3.csv \
    python bench.py --steps 1 --warmup 3 --no-rmat --no-rwr --no-rmat20 --no-cpu-baseline > gpurun_out/ncu_ll_r3.log 2>&1
tail -2 gpurun_out/ncu_ll_r3.log | cut -c1-300
