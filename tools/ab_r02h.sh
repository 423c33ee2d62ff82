#!/bin/bash
# A/B on one box: SUPER (i-tiles per super-row) variants and the mainloop-only experiment, then the new bench with the MMA peak
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
bash tools/ab_bench.sh base super32 super64 super96 skipepi 2>&1 | tee gpurun_out/r02h_ab.txt
timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r02h_bench_n1.json 2> gpurun_out/r02h_bench_n1.err; echo "bench rc=$?"
python -c "
import json; d=json.load(open('gpurun_out/r02h_bench_n1.json')); r=d['roofline']; print(d['value'], r['frac'], r['peak'], r['achieved'], r['traffic'], {k:v for k,v in r['peaks_measured_in_run'].items() if 'tops' in k})"
tail -3 gpurun_out/r02h_bench_n1.err
