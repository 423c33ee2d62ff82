#!/bin/bash
# ncu evidence only (the bench/test part of tools/profile_r01.sh already ran): launch list, full capture of the K2 main
# pass (with source), full capture of the other kernels of one step (no source import: gpurun_out/ returns <= 64 MiB).
# Summaries are produced on the box so that they come back even when a report is too large.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-r01f}
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench rc=$?"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches_raw.csv \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_l.log 2>&1
python tools/ncu_summarise.py launches gpurun_out/${TAG}_launches_raw.csv > gpurun_out/${TAG}_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mi_i8_kernel -s 3 -c 1 -o gpurun_out/${TAG}_mi_i8_full -f \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_f.log 2>&1
timeout 900 ncu --set full --clock-control none -k regex:'hamming_csa|mi_refine_coop|aracne_bitmap|gbits_build|mi_compact' -s 13 -c 13 -o gpurun_out/${TAG}_others_full -f \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_o.log 2>&1
python tools/ncu_summarise.py full gpurun_out/${TAG}_mi_i8_full.ncu-rep gpurun_out/${TAG}_others_full.ncu-rep > gpurun_out/${TAG}_ncu_full_summary.csv
ls -la gpurun_out | tail -12
[ $(stat -c %s gpurun_out/${TAG}_others_full.ncu-rep) -gt 40000000 ] && rm -f gpurun_out/${TAG}_others_full.ncu-rep
du -sh gpurun_out
