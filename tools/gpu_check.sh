#!/bin/bash
# quick GPU confirmation of a change: whole GPU suite, then the N=1 bench (TAG = $1)
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-chk}
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_pytest.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/${TAG}_pytest.log
timeout 900 python bench.py --steps 5 --warmup 3 ${BENCH_ARGS} > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err; echo "bench rc=$?"
python - <<PY
import json
try:
    d = json.load(open("gpurun_out/${TAG}_bench_n1.json"))
    print({k: d[k] for k in ("value", "ms_per_step", "mi_aracne_wall_ms")}, "e2e", d.get("e2e", {}).get("value"), d["stages_ms"], "frac", d["roofline"]["frac"], d["config"]["saved_edges"], d["config"]["direct_edges"], "cand", d["candidates"])
except Exception as e:
    print("bench json:", e); print(open("gpurun_out/${TAG}_bench_n1.err").read()[-2000:])
PY
