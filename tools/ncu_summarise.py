#!/usr/bin/env python
"""Turns the ncu outputs of tools/profile_r01.sh into the small CSV summaries kept under profiles/.

  python tools/ncu_summarise.py launches gpurun_out/X_launches_raw.csv  > profiles/X_launches.csv
  python tools/ncu_summarise.py full gpurun_out/A.ncu-rep [B.ncu-rep..]  > profiles/X_ncu_full_summary.csv
  python tools/ncu_summarise.py index profiles/X_ncu_full_summary.csv ENGINE N L NGPUS   (adds / replaces the entry in profiles/index.json
                                                                                         that bench.py reads its `traffic` values from)
"""
import csv
import io
import subprocess
import sys
from collections import OrderedDict

METRICS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
           "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed",
           "sm__inst_executed.avg.per_cycle_elapsed", "lts__t_sector_hit_rate.pct",
           "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
           "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
           "launch__grid_size", "launch__block_size", "smsp__cycles_active.avg", "sm__cycles_elapsed.max",
           "sm__inst_executed.sum", "lts__t_bytes.sum", "l1tex__t_sector_hit_rate.pct",
           "smsp__inst_executed_pipe_fp64.sum", "sm__inst_executed_pipe_alu.sum", "sm__cycles_active.avg"]


def launches(path):
    rows = [l for l in open(path) if not l.startswith("==")]
    rd = csv.DictReader(io.StringIO("".join(rows)))
    agg = OrderedDict()
    for r in rd:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", ""))
        u = r["Metric Unit"]
        ms = v * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(u, 1e-6)
        k = r["Kernel Name"].split("(")[0].strip()
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += ms
    tot = sum(a[1] for a in agg.values())
    print("# ncu --metrics gpu__time_duration.sum --clock-control none: python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline "
          "(3000x100000; setup + threshold estimate + 2 steps; cold-cache serialised times: compare shares)")
    print("kernel,launches,total_ms,share_pct")
    for k, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print("%s,%d,%.3f,%.1f" % (k.replace(",", ";"), a[0], a[1], 100 * a[1] / tot))


def full(paths):
    print("# ncu --set full --clock-control none --import-source on, one launch each, 3000x100000 workload "
          "(python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline); per-launch values")
    for p in paths:
        out = subprocess.check_output(["ncu", "-i", p, "--page", "raw", "--csv"], text=True)
        rd = list(csv.reader(io.StringIO(out)))
        hdr, units = rd[0], rd[1]
        idx = {h: i for i, h in enumerate(hdr)}
        for h, i in list(idx.items()):                      # "TPC.TriageCompute.sm__pipe_tensor..." -> bare metric name
            idx.setdefault(h.split(".", 2)[-1] if h.split(".")[0] in ("TPC", "SM_C", "SM_A", "SM_B", "GPC", "FBP", "HUB") else h, i)
        seen = set()
        for r in rd[2:]:
            name = r[idx["Kernel Name"]].split("(")[0].strip().replace(",", ";")
            if name in seen:
                continue
            seen.add(name)
            parts = [name]
            for m in METRICS:
                if m in idx:
                    parts.append("%s=%s %s" % (m, r[idx[m]].replace(",", ""), units[idx[m]]))
            print(",".join(parts))


def _bytes(txt):
    v, u = txt.split()
    return float(v) * {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}[u]


def _ms(txt):
    v, u = txt.split()
    return float(v) * {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[u]


def index(summary, engine, n, L, ngpus):
    import json
    import os
    kernels = {}
    for line in open(summary):
        if line.startswith("#") or not line.strip():
            continue
        parts = line.strip().split(",")
        d = dict(x.split("=", 1) for x in parts[1:])
        name = parts[0].replace("void ", "").replace("; ", ",").strip()
        kernels[name] = {"dram_bytes_per_launch": _bytes(d["dram__bytes_read.sum"]) + _bytes(d["dram__bytes_write.sum"]),
                         "ncu_ms": _ms(d["gpu__time_duration.sum"]),
                         "tensor_pipe_active_pct": float(d["sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed"].split()[0]),
                         "ipc": float(d["sm__inst_executed.avg.per_cycle_elapsed"].split()[0])}
    main = next(k for k in kernels if k.startswith("mi_i8_kernel") or k.startswith("mi_gemm_kernel"))
    ent = {"engine": engine, "n": n, "L": L, "n_gpus": ngpus, "file": os.path.relpath(summary, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..")),
           "kernel": main, "dram_bytes_per_launch": kernels[main]["dram_bytes_per_launch"], "tensor_pipe_active_pct": kernels[main]["tensor_pipe_active_pct"],
           "kernels": kernels}
    path = os.path.join(os.path.dirname(os.path.abspath(summary)), "index.json")
    idx = json.load(open(path)) if os.path.exists(path) else {"ncu_full": []}
    idx["ncu_full"] = [e for e in idx["ncu_full"] if (e["engine"], e["n"], e["L"], e.get("n_gpus", 1)) != (engine, n, L, ngpus)] + [ent]
    json.dump(idx, open(path, "w"), indent=1, sort_keys=True)
    print("wrote", path, "(%d kernels)" % len(kernels))


if __name__ == "__main__":
    if sys.argv[1] == "launches":
        launches(sys.argv[2])
    elif sys.argv[1] == "index":
        index(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]), int(sys.argv[5]), int(sys.argv[6]))
    else:
        full(sys.argv[2:])
