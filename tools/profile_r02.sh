#!/bin/bash
# Round-2 evidence run on one B200 (gpurun -- 'bash tools/profile_r02.sh TAG [notest]'):
#   1. GPU tests (whole suite, then the no-alloc test alone in a fresh process)
#   2. bench.py at N=1 (config 2) -> gpurun_out/TAG_bench_n1.json
#   3. ncu launch list of one bench run (gpu__time_duration.sum, --clock-control none)
#   4. ncu --set full: K2 main pass with source, then one launch of each other kernel of a step
# Summaries are produced on the box (reports may exceed what gpurun_out/ brings back).
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-r02g}
if [ "$2" != "notest" ]; then
  timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/${TAG}_pytest.log
  timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/${TAG}_pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -2 gpurun_out/${TAG}_pytest_noalloc_alone.log
fi
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench rc=$?"
python - <<EOF
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_n1.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "mi_aracne_wall_ms")}, d["e2e"]["value"], d["stages_ms"], d["roofline"]["frac"], d["config"]["saved_edges"], d["config"]["direct_edges"])
except Exception as e:
    print("bench json:", e)
EOF
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/${TAG}_launches_raw.csv \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-peaks > gpurun_out/${TAG}_ncu_l.log 2>&1
python tools/ncu_summarise.py launches gpurun_out/${TAG}_launches_raw.csv > gpurun_out/${TAG}_launches.csv
head -30 gpurun_out/${TAG}_launches.csv
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mi_i8_kernel -s 3 -c 1 -o gpurun_out/${TAG}_mi_i8_full -f \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-peaks > gpurun_out/${TAG}_ncu_f.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:'hamming_csa|mi_refine_coop|aracne_hash|lists_and_hash|gbits_build|mi_compact|i8_prep|audit_append|scatter_segment|sample_pairs' -s 24 -c 18 -o gpurun_out/${TAG}_others_full -f \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-peaks > gpurun_out/${TAG}_ncu_o.log 2>&1
python tools/ncu_summarise.py full gpurun_out/${TAG}_mi_i8_full.ncu-rep gpurun_out/${TAG}_others_full.ncu-rep > gpurun_out/${TAG}_ncu_full_summary.csv
ls -la gpurun_out | tail -14
[ -f gpurun_out/${TAG}_others_full.ncu-rep ] && [ $(stat -c %s gpurun_out/${TAG}_others_full.ncu-rep) -gt 30000000 ] && rm -f gpurun_out/${TAG}_others_full.ncu-rep
du -sh gpurun_out
