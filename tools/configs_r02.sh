#!/bin/bash
# One-GPU evidence for the other BASELINE configs and the CLI (gpurun -- 'bash tools/configs_r02.sh'):
#   CLI end to end at config 2; bench at configs 5, 1 and 4; refine-kernel ncu capture with source; config-4 fullsize test log
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
show() { python -c "
import json,sys; d=json.load(open(sys.argv[1])); print(sys.argv[1], '%.2f Gpairs/s' % d['value'], '%.2f ms' % d['ms_per_step'], 'e2e %.2f' % d['e2e']['value'], 'wall %.1f' % d['mi_aracne_wall_ms'], d['config']['saved_edges'], d['config']['direct_edges'], {k: round(v, 2) for k, v in d['stages_ms'].items()}, 'cpu', (d.get('cpu_baseline') or {}).get('value'))" $1 || tail -5 ${1%.json}.err; }
timeout 900 bash tools/cli_config2.sh 1 > gpurun_out/r02_cli_config2.txt 2>&1; echo "cli rc=$?"; grep -E "total|read |filtered|threshold =|done in|Elapsed" gpurun_out/r02_cli_config2.txt | tail -14
timeout 900 python bench.py --steps 3 --warmup 3 --n 3000 --L 50000 --mi-values 50000000 > gpurun_out/r02_bench_config5_n1.json 2> gpurun_out/r02_bench_config5_n1.err; echo "config5 rc=$?"; show gpurun_out/r02_bench_config5_n1.json
timeout 600 python bench.py --steps 5 --warmup 3 --n 200 --L 5000 > gpurun_out/r02_bench_config1_n1.json 2> gpurun_out/r02_bench_config1_n1.err; echo "config1 rc=$?"; show gpurun_out/r02_bench_config1_n1.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'mi_refine_coop' -s 9 -c 9 -o gpurun_out/r02_refine_full -f \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-peaks > gpurun_out/r02_ncu_refine.log 2>&1; echo "ncu refine rc=$?"
ls -la gpurun_out/r02_refine_full.ncu-rep
timeout 2400 python bench.py --steps 3 --warmup 3 --n 20000 --L 250000 --linear-genome > gpurun_out/r02_bench_config4_n1.json 2> gpurun_out/r02_bench_config4_n1.err; echo "config4 rc=$?"; show gpurun_out/r02_bench_config4_n1.json
SPK_FULLSIZE=4 timeout 3000 python -m pytest tests/test_gpu_fullsize.py -m gpu -q -s -k config4 > gpurun_out/r02_fullsize_config4.log 2>&1; echo "fullsize config 4 rc=$?"; tail -12 gpurun_out/r02_fullsize_config4.log
du -sh gpurun_out
