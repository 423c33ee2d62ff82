#!/bin/bash
# A/B of library variants on one GPU box: tools/ab_bench.sh name1 name2 ... ("base" = the in-tree library)
# variants are built into variants/lib_<name>.so (git-ignored; they travel with gpurun snapshots)
cd "$(dirname "$0")/.."
for v in "$@"; do
  if [ "$v" = base ]; then unset SPYDRPICK_B200_LIB; else export SPYDRPICK_B200_LIB=$PWD/variants/lib_$v.so; fi
  timeout 300 python bench.py --steps 3 --warmup 3 --no-e2e --no-cpu-baseline 2>/dev/null | python -c "
import sys, json
for l in sys.stdin:
    if l.startswith('{'):
        d = json.loads(l); s = d['stages_ms']
        print('$v', 'Gpairs/s %.2f' % d['value'], 'step %.2f' % d['ms_per_step'], 'K1 %.2f' % s['hamming_ms'], 'K2 %.2f' % s['gemm_ms'], 'refine %.2f' % s['refine_ms'], 'cand', d['candidates'], 'edges', d['config']['saved_edges'])
"
done
