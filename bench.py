#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path: weighted all-pairs MI scan + ARACNE.

Workload (BASELINE.json configs[1]): synthetic bacterial-like alignment, 3,000 sequences x 100,000 filtered SNP columns.
One STEP = one pass of the hot path: K1 neighbour counts -> weights -> K2 scan (tcgen05 filter + audit + fp64 refine +
compaction) -> edge sort -> [multi-GPU: NCCL exchange + run merge] -> K3 ARACNE. `value` = column pairs per second with the
alignment already resident in HBM; `e2e` = the same step through the host-buffer C-ABI (alignment H2D and edges / flags D2H
inside the timed region). N > 1: one process per GPU (torchrun); every rank owns one spk context that belongs to the library's
own NCCL communicator (spk_comm_init_rank) and makes the same collective C-ABI calls -- the exchange is inside the library.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                      (the REAL reference compiled in place, bounded sample, host cores)
  python bench.py --n 3000 --L 50000 --mi-values 50000000   (config 5)   --n 20000 --L 250000 --linear-genome (config 4)
"""
import argparse
import glob
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "weighted_mi_scan_gpairs_per_s"
UNIT = "Gpairs/s"
EPS = 2.220446049250313e-16


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", "--samples", dest="n", type=int, default=3000)        # (--samples / --columns: torchrun's own parser trips over "--n")
    ap.add_argument("--L", "--columns", dest="L", type=int, default=100000)
    ap.add_argument("--seed", type=int, default=20260102)
    ap.add_argument("--n-terms", type=int, default=0)
    ap.add_argument("--engine", type=int, default=0, help="0 = default int8 engine (= 4), 3 = int8 compact one-hot GEMM with {hi, lo} 16-bit weight rows, 4 = factorised byte weights (half the tensor work), 2 = dense bf16 GEMM")
    ap.add_argument("--mi-values", type=int, default=0, help="saved pairs target (0 = min(1e7, 100*L))")
    ap.add_argument("--linear-genome", action="store_true")
    ap.add_argument("--cpu-sample-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-peaks", action="store_true", help="skip the in-run int8 / bf16 tensor peak measurement")
    return ap.parse_args()


def driver_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_tflops_sustained": d.get("bf16_tflops_sustained"), "bf16_tflops_burst": d.get("bf16_tflops"), "hbm_gbs": d.get("hbm_gbs", 6534.8),
                "source": "MEASURED_PEAKS.json (driver-written)"}
    return {"bf16_tflops_sustained": 1400.0, "bf16_tflops_burst": 1650.0, "hbm_gbs": 6500.0, "source": "fallback (B200_PROFILING.md)"}


def measure_tensor_peaks(torch, device, seconds=2.0):
    """Dense int8 (s8 x s8 -> s32, torch._int_mm = cuBLASLt) and bf16 peaks on THIS box in THIS run: 8192^3, best of 10 (burst)
    and back to back for `seconds` (sustained). The MI filter issues tcgen05 kind::i8 MMAs, so its utilisation is quoted against
    the measured int8 number; the driver's bf16 figures are printed next to it."""
    out = {}
    M = 8192
    ops = 2.0 * M * M * M
    try:
        a = torch.randint(-8, 8, (M, M), dtype=torch.int8, device=device); b = torch.randint(-8, 8, (M, M), dtype=torch.int8, device=device)
        f = lambda: torch._int_mm(a, b)
        for name, fn in (("int8", f),):
            for _ in range(3):
                fn()
            torch.cuda.synchronize(device)
            best = 1e9
            for _ in range(10):
                e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
                e0.record(); fn(); e1.record(); torch.cuda.synchronize(device)
                best = min(best, e0.elapsed_time(e1))
            e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
            reps = max(10, int(seconds * 1e3 / best))
            e0.record()
            for _ in range(reps):
                fn()
            e1.record(); torch.cuda.synchronize(device)
            out["int8_tops_burst"] = ops / (best * 1e-3) / 1e12
            out["int8_tops_sustained"] = ops * reps / (e0.elapsed_time(e1) * 1e-3) / 1e12
        out["int8_source"] = "torch._int_mm 8192^3 (cuBLASLt s8 x s8 -> s32) in this run: best of 10 (burst), back to back for %.0f s (sustained)" % seconds
        del a, b
    except Exception as ex:                               # no int8 GEMM available: fall back to 2 x the driver's bf16 figure, say so
        out["int8_error"] = str(ex)[:200]
    return out


def measure_mma_peak(device_index, seconds=2.0):
    """Peak of the instruction the filter issues, measured with the library's MMA-only kernel (spk_debug_i8_peak: tcgen05.mma
    cta_group::2 kind::i8 M256 N256 K32 on shared-memory operands, one CTA pair per TPC, no loads / epilogue): best of 8 launches
    of ~30 ms (burst) and launches back to back for `seconds` (sustained)."""
    from spydrpick_b200 import Context
    out = {}
    try:
        with Context(device=device_index) as c:
            kb = 100000
            c.debug_i8_peak(kb, 1)
            best, best_ms = 0.0, None
            for _ in range(8):
                t, ms = c.debug_i8_peak(kb, 1)
                if t > best:
                    best, best_ms = t, ms
            reps = max(4, int(seconds * 1e3 / best_ms))
            ts, mss = c.debug_i8_peak(kb, reps)
            out["i8_mma_tops_burst"] = best
            out["i8_mma_tops_sustained"] = ts
            out["i8_mma_source"] = ("spk_debug_i8_peak in this run: tcgen05.mma.cta_group::2.kind::i8 M256 N256 K32 from shared memory, 74 CTA pairs, "
                                    "%d K blocks per launch; best of 8 launches (%.1f ms, burst), %d launches back to back (%.0f ms, sustained)" % (kb, best_ms, reps, mss))
    except Exception as ex:
        out["i8_mma_error"] = str(ex)[:200]
    return out


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.index = index; self.samples = []; self.stop = False; self.t = None

    def _run(self):
        while not self.stop:
            try:
                out = subprocess.check_output(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                               "--format=csv,noheader,nounits"], timeout=5).decode().strip()
                self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.t = threading.Thread(target=self._run, daemon=True); self.t.start(); return self

    def __exit__(self, *a):
        self.stop = True; self.t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples if s[0].replace(".", "").isdigit())
        mx = [float(s[1]) for s in self.samples if s[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [nm for k, nm in enumerate(names) if any(len(s) > 2 + k and s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons}


def make_workload(args):
    from spydrpick_b200.synth import generate_alignment
    return generate_alignment(args.n, args.L, seed=args.seed)


def workload_name(args):
    tag = {(3000, 100000): "config 2/3", (20000, 250000): "config 4", (3000, 50000): "config 5", (200, 5000): "config 1"}.get((args.n, args.L), "custom")
    return "%dx%d synthetic SNP alignment (%s): K1 reweighting + weighted MI scan + sort + ARACNE" % (args.n, args.L, tag)


# ------------------------------------------------------------------------------------------------
# CPU baseline: the REAL reference (oracle/_ref, compiled in place from /root/reference), all host threads, bounded sample
# ------------------------------------------------------------------------------------------------
def reference_sample(states, pos, threshold, seconds, linear=False, fixed_cols=None):
    """Times cache_sample_weights + get_MI_network + Graph::sort + run_ARACNE of the compiled reference (oracle/_ref/libspkref.so)
    on the first Ls columns (all samples) of the workload; Ls sized for ~`seconds` of work. Falls back to the oracle port if the
    reference was never built. Returns the cpu_baseline dict."""
    n, L = states.shape
    try:
        import oracle.ref as ref
        have_ref = ref.available()
    except Exception:
        have_ref = False
    if have_ref:
        cores = os.cpu_count() or 1
        opts = ["--mi-threshold", repr(threshold), "-t", str(cores)] + (["--linear-genome"] if linear else [])
        ref.configure(*opts)

        def run(Ls):
            sub = np.ascontiguousarray(states[:, :Ls])
            A = ref.Alignment.from_states(sub, "bench", pos[:Ls])          # parse + block compression: outside the timed region
            t0 = time.time()
            A.cache_sample_weights()
            t1 = time.time()
            net = A.scan(threshold)
            t2 = time.time()
            net.sort()
            t3 = time.time()
            net.aracne()
            t4 = time.time()
            return t4 - t0, (t1 - t0, t2 - t1, t3 - t2, t4 - t3), int(len(net.edges()[0]))
        kind, label = "reference", "santeripuranen/SpydrPick f6bc110 compiled in place (oracle/_ref/libspkref.so; tbb::parallel_* on std::thread)"
    else:
        import oracle.oracle as orc
        cores = orc.hardware_threads()

        def run(Ls):
            sub = np.ascontiguousarray(states[:, :Ls])
            t0 = time.time()
            w = orc.weights_from_counts(orc.neighbour_counts(sub, int(0.9 * Ls), True))
            t1 = time.time()
            r = orc.mi_scan(sub, w, pos[:Ls], threshold=threshold, linear_genome=linear, circle_size=Ls)
            t2 = time.time()
            e = orc.sort_edges(r["edges"])
            t3 = time.time()
            orc.aracne_closed(e)
            t4 = time.time()
            return t4 - t0, (t1 - t0, t2 - t1, t3 - t2, t4 - t3), len(e)
        kind, label = "port", "oracle/oracle.cpp (CPU restatement; oracle/_ref not built)"
    if fixed_cols:
        Ls = int(min(L, fixed_cols))
    else:
        Lc = int(min(L, 600))
        dt, _, _ = run(Lc)
        rate = (Lc * (Lc - 1) / 2) / max(dt, 1e-6)
        Ls = int(min(L, max(Lc, (2 * seconds * rate) ** 0.5)))
    dt, stages, ne = run(Ls)
    pairs = Ls * (Ls - 1) / 2
    # rate on the full workload: the reweighting pass grows with L, scan / sort / ARACNE with the number of pairs -- a column subset
    # over-represents the former, so the sample is extrapolated stage by stage (the raw sample rate is kept next to it)
    full_pairs = L * (L - 1) / 2
    t_full = stages[0] * (L / Ls) + (stages[1] + stages[2] + stages[3]) * (full_pairs / pairs)
    return {"value": full_pairs / t_full / 1e9, "unit": UNIT, "cores": cores, "kind": kind, "implementation": label,
            "sample": "first %d of %d columns (all %d samples): %d pairs; weights %.2f s + scan %.2f s + sort %.2f s + ARACNE %.2f s = %.2f s; %d edges; "
                      "alignment parse / block compression outside the timed region; value = pairs/s of the full %d-column workload extrapolated stage by stage "
                      "(weights ~ L, the rest ~ pairs: %.0f s)" % (Ls, L, n, pairs, stages[0], stages[1], stages[2], stages[3], dt, ne, L, t_full),
            "sample_rate": pairs / dt / 1e9, "seconds": dt, "columns": Ls, "full_workload_seconds_extrapolated": t_full}


def run_reference(args):
    """--impl reference: the reference's own CPU implementation of the path on the box's host cores, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    states, pos, _ = make_workload(args)
    n, L = states.shape
    # the same saving threshold as the B200 arm uses for this workload (its estimator is deterministic for the default seed and
    # mi_values; recorded in profiles/index.json), 0.05 for a workload without a record
    thr = 0.05
    try:
        if args.seed == 20260102 and not args.mi_values:
            thr = float(json.load(open(os.path.join(ROOT, "profiles", "index.json")))["workload_thresholds"]["values"].get("%dx%d" % (n, L), thr))
    except Exception:
        pass
    # one calibration run sizes the sample so that warmup + steps end within a few minutes
    first = reference_sample(states, pos, thr, 20.0 / max(1, args.warmup + args.steps) + 2.0, linear=args.linear_genome)
    cols = first["columns"]
    times = []
    for it in range(args.warmup + args.steps):
        r = reference_sample(states, pos, thr, 0, linear=args.linear_genome, fixed_cols=cols)
        if it >= args.warmup:
            times.append(r["seconds"])
    ms = 1e3 * sum(times) / len(times)
    v = r["value"] * r["seconds"] / (ms / 1e3)            # the extrapolated full-workload rate, on the mean step time
    cb = dict(r); cb["value"] = v
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                      "dtype": "f64", "data": "synthetic",
                      "config": {"workload": workload_name(args) + " -- CPU reference on a bounded sample", "n": n, "L": L, "mi_threshold": thr, "sample": cb["sample"]},
                      "cpu_baseline": cb,
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
def ncu_profile_entry(engine, n, L, world):
    """traffic (dram bytes read + written per launch) and tensor-pipe-active % of the main-pass launch from the committed
    ncu --set full summary that matches this engine / shape (profiles/*ncu_full_summary*.csv via profiles/index.json), else None."""
    idx = os.path.join(ROOT, "profiles", "index.json")
    if not os.path.exists(idx):
        return None
    try:
        for ent in json.load(open(idx)).get("ncu_full", []):
            if ent.get("engine") == engine and ent.get("n") == n and ent.get("L") == L and ent.get("n_gpus", 1) == world:
                return ent
    except Exception:
        return None
    return None


def run_b200(args):
    import torch
    import torch.distributed as dist
    from spydrpick_b200 import Context
    from spydrpick_b200.capi import pinned_copy, pinned_empty
    world = int(os.environ.get("WORLD_SIZE", "1")); rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- the B200 path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    device = torch.device("cuda", local_rank)
    # stdout carries the ONE JSON line: anything native libraries print (NCCL's version banner) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1); os.dup2(2, 1)
    if world > 1:
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"            # keep stdout to the single JSON line
        dist.init_process_group("nccl", device_id=device)

    n, L = args.n, args.L
    i8 = args.engine in (0, 3, 4)
    eng = args.engine if args.engine else 4
    tensor_peaks = {}
    if rank == 0 and not args.no_peaks:
        tensor_peaks = measure_tensor_peaks(torch, device)
        tensor_peaks.update(measure_mma_peak(local_rank))

    # ---- workload: rank 0 generates, alignment broadcast ONCE over NCCL (outside the timed region)
    if rank == 0:
        states, pos, gsize = make_workload(args)
    if world > 1:
        st = torch.from_numpy(states).to(device) if rank == 0 else torch.empty((n, L), dtype=torch.uint8, device=device)
        ps = torch.from_numpy(pos.astype(np.int64)).to(device) if rank == 0 else torch.empty(L, dtype=torch.int64, device=device)
        dist.broadcast(st, 0); dist.broadcast(ps, 0)
        states = st.cpu().numpy(); pos = ps.cpu().numpy().astype(np.uint64)
        del st, ps
    pairs_total = L * (L - 1) // 2
    mi_values = args.mi_values or min(10_000_000, 100 * L)

    ctx = Context(device=local_rank)
    # ONE stream for torch's events and the library's kernels
    stream = torch.cuda.Stream(device)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    if world > 1:
        # the library's own communicator: rank 0 makes the id, torch.distributed only carries the 128 bytes
        uid = torch.zeros(128, dtype=torch.uint8, device=device)
        if rank == 0:
            uid.copy_(torch.frombuffer(bytearray(Context.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(uid, 0)
        ctx.comm_init_rank(bytes(uid.cpu().numpy().tobytes()), rank, world)
    states_pin = pinned_copy(states)                  # the caller's alignment lives in page-locked host memory
    pos_pin = pinned_copy(pos)
    ctx.set_alignment(states_pin, pos_pin)            # collective for world > 1: n/G rows per rank + all-gather over NVLink
    min_ident = int(0.9 * float(L))
    weights, counts = ctx.reweight(min_ident, True)
    # save threshold from one K2b estimate with the reference's sampler (deterministic: every rank computes the same value)
    threshold = ctx.determine_mi_threshold(mi_values, iterations=10, seed=args.seed)
    max_edges = int(mi_values * 2.5) // world + 65536
    scan_kw = dict(max_edges=max_edges, sort_desc=True, n_terms=args.n_terms, engine=args.engine, linear_genome=args.linear_genome,
                   n_parts=0 if world > 1 else 1)
    host_flags = None

    def step_device():
        """alignment resident in HBM; results stay in HBM."""
        ctx.reweight(min_ident, True)
        r = ctx.mi_scan(threshold, keep_on_device=True, **scan_kw)
        ne = r["n_edges"]
        flags = torch.empty(max(ne, 1), dtype=torch.uint8, device=device)
        ctx.aracne_device(r["d_edges"], ne, flags.data_ptr(), EPS, part_index=0, n_parts=0 if world > 1 else 1)
        return r, ne, flags

    def step_wall():
        """SURVEY 8(d) metric (ii): K1 + threshold estimate + scan + sort + ARACNE, results in HBM."""
        ctx.reweight(min_ident, True)
        thr = ctx.determine_mi_threshold(mi_values, iterations=10, seed=args.seed, sampler="device")
        r = ctx.mi_scan(thr, keep_on_device=True, **scan_kw)
        ne = r["n_edges"]
        flags = torch.empty(max(ne, 1), dtype=torch.uint8, device=device)
        ctx.aracne_device(r["d_edges"], ne, flags.data_ptr(), EPS, part_index=0, n_parts=0 if world > 1 else 1)
        return r, ne, flags

    def step_e2e():
        """host buffers through the C-ABI: H2D of the alignment (pinned), D2H of edges + flags inside the step."""
        nonlocal host_flags
        tw = [time.perf_counter()]
        ctx.set_alignment(states_pin, pos_pin); tw.append(time.perf_counter())
        ctx.reweight(min_ident, True); tw.append(time.perf_counter())
        r = ctx.mi_scan(threshold, keep_on_device=False, copy=False, async_download=True, **scan_kw); tw.append(time.perf_counter())
        ne = r["n_edges"]
        if world == 1 or rank == 0:
            edges = r["edges"]                        # view of the library's pinned result buffer (download in flight)
            if host_flags is None or host_flags.shape[0] < ne:
                host_flags = pinned_empty(int(ne * 1.25) + 4096, np.uint8)
            if world == 1:
                flags = ctx.aracne(edges, EPS, out=host_flags)   # same pointer -> device copy reused; overlaps the edge download
            else:
                flags = ctx.aracne_collective(ne, EPS, out=host_flags, edges=edges)
            ctx.wait_download()                       # edges + MI-without-gaps lists are now complete on the host
            ndirect = int(np.count_nonzero(flags))    # the step's result is read on the host
            d2h = edges.nbytes + r["edges_wo_gaps"].nbytes + 8 * L + 8 * n + ne
        else:
            ctx.aracne_collective(ne, EPS)
            ndirect = 0; d2h = 8 * L + 8 * n
        rows = (n + world - 1) // world
        h2d = min(rows, max(0, n - rank * rows)) * L + pos.nbytes + 8 * n
        tw.append(time.perf_counter())
        if os.environ.get("SPK_BENCH_E2E_BREAKDOWN") and rank == 0:
            tmx = ctx.timings()
            sys.stderr.write("e2e wall ms: set_alignment %.1f (upload %.1f prep %.1f) | reweight %.1f | scan %.1f | aracne + wait %.1f\n" % (
                1e3 * (tw[1] - tw[0]), tmx["upload_ms"], tmx["prep_ms"], 1e3 * (tw[2] - tw[1]), 1e3 * (tw[3] - tw[2]), 1e3 * (tw[4] - tw[3])))
        return h2d, d2h, ndirect

    def timed(fn, steps, warmup, with_clocks=False):
        for _ in range(warmup):
            out = fn()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        sampler = ClockSampler(local_rank) if with_clocks else None
        if sampler:
            sampler.__enter__()
        lt0 = ctx.timings()["kernel_launches"]
        stage_sum = {}
        e0.record(stream)
        for _ in range(steps):
            out = fn()
            tmk = ctx.timings()
            for k, v in tmk.items():
                if k.endswith("_ms"):
                    stage_sum[k] = stage_sum.get(k, 0.0) + v
        e1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        if sampler:
            sampler.__exit__()
        ms = e0.elapsed_time(e1) / steps
        launches = (ctx.timings()["kernel_launches"] - lt0) // max(steps, 1)
        if world > 1:
            t = torch.tensor([ms], dtype=torch.float64, device=device); dist.all_reduce(t, op=dist.ReduceOp.MAX); ms = float(t.item())
        return ms, out, launches, (sampler.summary() if sampler else None), {k: v / steps for k, v in stage_sum.items()}

    ms, (r, ne, dflags), launches, clocks, stages = timed(step_device, args.steps, args.warmup, with_clocks=True)
    ndirect = int(dflags[:ne].sum().item())
    tm = ctx.timings()
    value = pairs_total / (ms / 1e3) / 1e9
    wms, (rw, new, _), _, _, wstages = timed(step_wall, max(1, min(args.steps, 3)), 1)

    # ---- roofline of the dominant kernel (K2 main pass): issued tensor ops / measured peak of the pipe it uses
    pk = driver_peaks()
    tiles = tm["gemm_tiles"]
    pairs_launch = tm["gemm_pairs"]                      # pairs one launch (this rank) covers, incl. padding / masked diagonal halves
    k_ms = stages.get("gemm_ms", 0.0) - stages.get("seed_ms", 0.0)   # main pass = filter interval minus the seed pass
    n_pad = (n + 127) // 128 * 128
    issued = 2.0 * 128 * 256 * n_pad * tiles * (1 if i8 else (args.n_terms or 2))   # tensor ops actually issued by one launch
    issued_tops = issued / (k_ms / 1e3) / 1e12 if k_ms > 0 else 0.0
    alg_flops = 2.0 * 16.0 * n * pairs_launch           # 32*n flop per column pair (dense 4x4 cells)
    if i8:
        # denominator: the measured peak of the tcgen05 kind::i8 instruction itself (MMA-only kernel, burst = highest rate the pipe
        # reached in this run, so frac <= 1 by construction); cuBLASLt's s8 GEMM, reported next to it, reaches only ~3/4 of it
        peak = tensor_peaks.get("i8_mma_tops_burst")
        peak_src = tensor_peaks.get("i8_mma_source")
        if not peak:
            peak = tensor_peaks.get("int8_tops_burst")
            peak_src = tensor_peaks.get("int8_source")
        if not peak:
            peak = 2.0 * (pk["bf16_tflops_burst"] or 1650.0)
            peak_src = "2 x the driver's burst bf16 figure (no int8 measurement in this run: %s)" % tensor_peaks.get("i8_mma_error", tensor_peaks.get("int8_error", "--no-peaks"))
        unit = "TOP/s"
    else:
        peak = pk["bf16_tflops_sustained"]; peak_src = pk["source"]; unit = "TFLOP/s"
    prof = ncu_profile_entry(eng, n, L, world)
    roofline = {"bound": "tensor", "achieved": issued_tops, "peak": peak, "unit": unit, "frac": issued_tops / peak if peak else None,
                "traffic": prof.get("dram_bytes_per_launch") if prof else None,
                "kernel": ("mi_i8_kernel<2,%d> (main pass)" % (1 if eng == 4 else 2)) if i8 else "mi_gemm_kernel", "kernel_ms": k_ms,
                "seed_pass_ms": stages.get("seed_ms", 0.0), "tiles_per_launch": int(tiles), "pairs_per_launch": int(pairs_launch),
                "peak_source": peak_src,
                "peaks_measured_in_run": tensor_peaks, "peaks_driver": pk,
                "frac_of_sustained_mma_peak": issued_tops / tensor_peaks["i8_mma_tops_sustained"] if i8 and tensor_peaks.get("i8_mma_tops_sustained") else None,
                "ratio_to_cublaslt_int8_gemm": issued_tops / tensor_peaks["int8_tops_burst"] if i8 and tensor_peaks.get("int8_tops_burst") else None,
                "algorithmic_equiv_tflops": alg_flops / (k_ms / 1e3) / 1e12 if k_ms > 0 else 0.0,
                "ncu_tensor_pipe_active_pct": prof.get("tensor_pipe_active_pct") if prof else None,
                "ncu_source": prof.get("file") if prof else None,
                "note": "achieved = u8 x u8 tensor ops ISSUED by one main-pass launch (2 * 128 * 256 * n_pad per M128 N256 tile; the one-hot row compaction "
                        "issues ~3.4 cells per column pair instead of 16) / launch time, against the peak of that instruction measured in this run (MMA-only kernel); "
                        "algorithmic_equiv_tflops is SURVEY 8(d)'s 32*n flop/pair figure (no fraction: it counts work the compact layout never issues)"}
    # ---- secondary kernels: achieved rates against the roofline that bounds each
    hbm = pk["hbm_gbs"]
    W32 = (L + 31) // 32
    k1_ms = stages.get("hamming_ms", 0.0)
    k1_words = n * (n - 1) / 2.0 * W32 / world
    sm_mhz = (clocks or {}).get("sm_mhz") or 1900.0
    popc_peak = 16.0 * 148 * sm_mhz * 1e6 / 1e9          # 16 POPC / clk / SM
    ncand = r["n_candidates"]
    ref_ms = stages.get("refine_ms", 0.0)
    kernels = {
        "K1 hamming_csa_kernel": {"ms": k1_ms, "bound": "issue (POPC/LOP3)", "gpopc_per_s": k1_words / (k1_ms * 1e-3) / 1e9 if k1_ms else None,
                                  "popc_peak_gpopc_per_s": popc_peak, "frac": (k1_words / (k1_ms * 1e-3) / 1e9) / popc_peak if k1_ms else None,
                                  "algorithmic_GBps": 3.0 * n * W32 * 4 / (k1_ms * 1e-3) / 1e9 if k1_ms else None},
        "K2r mi_refine_coop_kernel x9": {"ms": ref_ms, "bound": "L2 bandwidth (planes are L2-resident)", "candidates": int(ncand),
                                          "algorithmic_GBps": ncand * 2.5e3 / (ref_ms * 1e-3) / 1e9 if ref_ms else None, "hbm_peak_GBps": hbm},
        "sort (CUB radix x2 + gather)": {"ms": stages.get("sort_ms", 0.0), "bound": "hbm",
                                         "algorithmic_GBps": (ne / world) * 16 * 6 / (stages["sort_ms"] * 1e-3) / 1e9 if stages.get("sort_ms") else None, "hbm_peak_GBps": hbm},
        "K3 adjacency build (degree histogram + scan + one stable radix sort + per-node hash tables)": {"ms": stages.get("aracne_build_ms", 0.0), "bound": "hbm",
                                        "algorithmic_GBps": 2 * ne * (8 + 4) * 2 * 2 / (stages["aracne_build_ms"] * 1e-3) / 1e9 if stages.get("aracne_build_ms") else None, "hbm_peak_GBps": hbm},
        "K3 aracne_hash_kernel": {"ms": stages.get("aracne_ms", 0.0), "bound": "L2 latency / issue (sorted neighbour lists + hash tables are L2-resident)"},
        "exchange (NCCL + run merge)": {"ms": stages.get("exchange_ms", 0.0)},
        "filter audit": {"ms": stages.get("audit_ms", 0.0), "pairs": r["audit_pairs"], "max_error": r["audit_max_error"]},
    }

    # measured DRAM traffic per launch of the secondary kernels, from the same committed ncu --set full summary (null without one)
    pk_k = (prof or {}).get("kernels", {})
    def _traffic(*names):
        hits = [v["dram_bytes_per_launch"] for k, v in pk_k.items() if any(k.startswith(nm) for nm in names)]
        return sum(hits) if hits else None
    kernels["K1 hamming_csa_kernel"]["traffic"] = _traffic("hamming_csa_kernel")
    kernels["K2r mi_refine_coop_kernel x9"]["traffic"] = _traffic("mi_refine_coop_kernel")
    kernels["K3 aracne_hash_kernel"]["traffic"] = _traffic("aracne_hash_kernel")
    kernels["K3 adjacency build (degree histogram + scan + one stable radix sort + per-node hash tables)"]["traffic_lists_and_hash_kernel"] = _traffic("lists_and_hash_kernel")
    kernels["filter audit"]["traffic"] = _traffic("audit_append_kernel")
    kernels["operand prep i8_prep_kernel"] = {"ms": stages.get("expand_ms", 0.0), "bound": "hbm (write)", "algorithmic_GBps": (n * L + 2.0 * 1.7 * L * ((n + 127) // 128 * 128)) / (stages["expand_ms"] * 1e-3) / 1e9 if stages.get("expand_ms") else None,
                                              "hbm_peak_GBps": hbm, "traffic": _traffic("i8_prep_kernel")}
    kernels["compaction mi_compact_kernel"] = {"ms": stages.get("compact_ms", 0.0), "bound": "hbm", "traffic": _traffic("mi_compact_kernel")}

    e2e = None
    if not args.no_e2e:
        ems, (h2d, d2h, nd2), _, _, _ = timed(step_e2e, max(1, min(args.steps, 3)), 2)
        e2e = {"value": pairs_total / (ems / 1e3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": ems}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = reference_sample(states, pos, threshold, args.cpu_sample_seconds, linear=args.linear_genome)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "u8 x u8 -> s32 tensor filter + f64 refine" if i8 else "bf16 tensor filter + f64 refine",
                "data": "synthetic",
                "config": {"workload": workload_name(args),
                           "n": n, "L": L, "pairs": pairs_total, "saved_edges": int(ne), "direct_edges": int(ndirect), "mi_threshold": threshold,
                           "linear_genome": bool(args.linear_genome),
                           "engine": ("int8 compact one-hot, 16-bit weights as {hi, lo} rows" if eng == 3 else "int8 compact one-hot, factorised byte weights a*b") if i8 else "dense bf16 one-hot, n_terms=%d" % (args.n_terms or 2),
                           "l2_policy": "inputs (>1.5 GB one-hot operands) larger than L2",
                           "parallelism": "tile-space partition x%d; exchange inside the library (spk_comm_*: NCCL all-gather of edge lists, all-reduce of colmax / K1 counts)" % world},
                "clocks": clocks, "gpu_launches": int(launches),
                "stages_ms": {k: stages.get(k, 0.0) for k in ("hamming_ms", "stats_ms", "expand_ms", "gemm_ms", "seed_ms", "audit_ms", "refine_ms", "compact_ms",
                                                              "sort_ms", "exchange_ms", "aracne_build_ms", "aracne_ms")},
                "mi_aracne_wall_ms": wms,
                "mi_aracne_wall": {"ms": wms, "definition": "SURVEY 8(d) metric (ii): t(K1) + t(threshold estimate, 10 x %d pairs, device sampler) + t(scan) + t(sort) + t(ARACNE), "
                                                             "device-timed, results in HBM" % (100000), "threshold_ms": wstages.get("threshold_ms", 0.0),
                                   "saved_edges": int(new)},
                "candidates": r["n_candidates"], "max_filter_error": r["max_filter_error"],
                "audit": {"pairs": r["audit_pairs"], "violations": r["audit_violations"], "max_error": r["audit_max_error"], "band": r["band"]},
                "roofline": roofline, "kernels": kernels}
        if e2e:
            e2e["path"] = ("spk_set_alignment (pinned host states%s) -> spk_reweight -> spk_mi_scan -> spk_aracne, edges + MI-without-gaps + flags downloaded%s" %
                           (", n/G rows per rank + NCCL all-gather" if world > 1 else "", " on rank 0 (h2d bytes are per rank)" if world > 1 else ""))
            line["e2e"] = e2e
        if cpu:
            line["cpu_baseline"] = cpu
        os.write(json_fd, (json.dumps(line) + "\n").encode())
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_b200(args)


if __name__ == "__main__":
    main()
