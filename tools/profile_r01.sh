#!/bin/bash
# Round-1 evidence run on ONE B200 (launched through gpurun): GPU tests, plain bench, ncu launch list, ncu --set full captures.
# Outputs land in gpurun_out/; the summaries that are judged are copied into profiles/ afterwards.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
TAG=${1:-r01f}
timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/${TAG}_tests.log 2>&1; echo "tests rc=$?" ; tail -3 gpurun_out/${TAG}_tests.log
SPK_BENCH_E2E_BREAKDOWN=1 timeout 600 python bench.py --steps 5 --warmup 3 > gpurun_out/${TAG}_bench_n1.json 2> gpurun_out/${TAG}_bench_n1.err || { echo bench failed; tail -5 gpurun_out/${TAG}_bench_n1.err; exit 1; }
cat gpurun_out/${TAG}_bench_n1.json | cut -c1-600; tail -2 gpurun_out/${TAG}_bench_n1.err
timeout 300 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/${TAG}_bench_ref.json 2> gpurun_out/${TAG}_bench_ref.err; echo "ref rc=$?"
python - <<'PY'
import torch, time
d = torch.empty(330_000_000, dtype=torch.uint8, device="cuda"); h = torch.empty(330_000_000, dtype=torch.uint8).pin_memory()
for _ in range(2):
    torch.cuda.synchronize(); t = time.perf_counter(); h.copy_(d, non_blocking=True); torch.cuda.synchronize(); dt = time.perf_counter() - t
print("D2H 330 MB pinned: %.2f ms (%.1f GB/s)" % (dt * 1e3, 0.33 / dt))
PY
# launch list of the same command (durations are cold-cache/serialised: shares, not absolutes)
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/${TAG}_launches_raw.csv \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_l.log 2>&1
# full capture of the K2 main-pass launch of the timed step (launches of mi_i8_kernel: seed, main per step; threshold estimate launches none)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mi_i8_kernel -s 3 -c 1 -o gpurun_out/${TAG}_mi_i8_full -f \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_f.log 2>&1
# the other kernels of the timed step (second step of the command: skip the first step's launches)
timeout 900 ncu --set full --clock-control none --import-source on -k regex:'hamming_csa|mi_refine_coop|aracne_bitmap|gbits_build|mi_compact' -s 13 -c 13 -o gpurun_out/${TAG}_others_full -f \
  python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline > gpurun_out/${TAG}_ncu_o.log 2>&1
ls -la gpurun_out | tail -8
