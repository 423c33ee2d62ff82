#!/bin/bash
# Multi-GPU evidence on one box with N GPUs (gpurun --gpus N -- 'bash tools/scale_r02.sh TAG N'):
#   bench at 1, 2, 4, .. N GPUs (config 2), config 4 and config 5 at N GPUs, the 2-GPU tests, the CLI with --gpus N
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-r02s}; NG=${2:-8}
PORT=29611
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv,noheader | head -8
timeout 600 python -m pytest tests -m gpu -q -k "two_gpus or collective or multi" > gpurun_out/${TAG}_pytest_multi.log 2>&1; echo "multi-GPU tests rc=$?"; tail -3 gpurun_out/${TAG}_pytest_multi.log
for N in 1 2 4 8; do
  [ $N -gt $NG ] && break
  if [ $N = 1 ]; then
    timeout 600 python bench.py --gpus 1 --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err
  else
    PORT=$((PORT+1))
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_n$N.json 2> gpurun_out/${TAG}_bench_n$N.err
  fi
  echo "bench N=$N rc=$?"
  python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_n$N.json"))
    print("N=$N", "%.2f Gpairs/s" % d["value"], "%.2f ms" % d["ms_per_step"], "e2e %.2f" % d["e2e"]["value"], {k: round(v, 2) for k, v in d["stages_ms"].items()}, d["config"]["saved_edges"], d["config"]["direct_edges"])
except Exception as e:
    print("bench N=$N:", e); print(open("gpurun_out/${TAG}_bench_n$N.err").read()[-1500:])
PY
done
if [ $NG -gt 1 ]; then
  PORT=$((PORT+1))
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $NG --steps 3 --warmup 3 --samples 3000 --columns 50000 --mi-values 50000000 > gpurun_out/${TAG}_bench_config5_n$NG.json 2> gpurun_out/${TAG}_bench_config5_n$NG.err; echo "config 5 N=$NG rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench_config5_n$NG.json')); print('config5', d['value'], d['ms_per_step'], d['e2e']['value'], d['config']['saved_edges'], d['config']['direct_edges'])"
  PORT=$((PORT+1))
  timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $NG --master-addr 127.0.0.1 --master-port $PORT bench.py --gpus $NG --steps 3 --warmup 3 --samples 20000 --columns 250000 --linear-genome > gpurun_out/${TAG}_bench_config4_n$NG.json 2> gpurun_out/${TAG}_bench_config4_n$NG.err; echo "config 4 N=$NG rc=$?"
  python -c "
import json; d=json.load(open('gpurun_out/${TAG}_bench_config4_n$NG.json')); print('config4', d['value'], d['ms_per_step'], d['e2e']['value'], d['mi_aracne_wall_ms'], d['config']['saved_edges'], d['config']['direct_edges'])" || tail -5 gpurun_out/${TAG}_bench_config4_n$NG.err
  timeout 600 bash tools/cli_config2.sh $NG > gpurun_out/${TAG}_cli_config2_gpus$NG.txt 2>&1; echo "cli rc=$?"; grep -E "total|read |wrote|done in" gpurun_out/${TAG}_cli_config2_gpus$NG.txt | tail -12
fi
