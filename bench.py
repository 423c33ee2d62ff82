#!/usr/bin/env python
"""bench.py -- headline benchmark of the hot path: weighted all-pairs MI scan + ARACNE.

Workload (BASELINE.json configs[1]): synthetic bacterial-like alignment, 3,000 sequences x 100,000
filtered SNP columns. One STEP = one pass of the hot path: K1 neighbour counts -> weights -> K2 scan
(tcgen05 filter + fp64 refine + compaction) -> edge sort -> K3 ARACNE. `value` = column pairs per
second with the alignment already resident in HBM; `e2e` = the same step through the host-buffer
C-ABI (alignment H2D and edges / flags D2H inside the timed region).

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torch.distributed.run)
  python bench.py --impl reference ...                      (CPU port of the reference, bounded sample)
"""
import argparse
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


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--n", type=int, default=3000)
    ap.add_argument("--L", type=int, default=100000)
    ap.add_argument("--seed", type=int, default=20260102)
    ap.add_argument("--n-terms", type=int, default=0)
    ap.add_argument("--engine", type=int, default=0, help="0 = default int8 engine (= 4), 3 = int8 compact one-hot GEMM with {hi, lo} 16-bit weight rows, 4 = factorised byte weights (half the tensor work), 2 = dense bf16 GEMM")
    ap.add_argument("--mi-values", type=int, default=0, help="saved pairs target (0 = min(1e7, 100*L))")
    ap.add_argument("--cpu-sample-seconds", type=float, default=15.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


# from the committed ncu --set full captures of the main-pass launch at 3000 x 100000
# (engine 4: profiles/r01f_ncu_full_summary.csv; engine 3: profiles/r01_ncu_full_summary.csv):
# dram__bytes_read.sum + dram__bytes_write.sum per launch, sm__pipe_tensor_cycles_active %, per int8 engine
NCU_TRAFFIC = {3: 52.2e9, 4: 26.7e9}
NCU_TENSOR_PCT = {3: 90.4, 4: 79.1}


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return {"bf16_tflops": d.get("bf16_tflops_sustained", d.get("bf16_tflops", 1400.0)), "hbm_gbs": d.get("hbm_gbs", 6650.0),
                "bf16_burst_tflops": d.get("bf16_tflops", 1650.0),
                "source": "measured (MEASURED_PEAKS.json, sustained bf16)"}
    return {"bf16_tflops": 1400.0, "bf16_burst_tflops": 1650.0, "hbm_gbs": 6650.0, "source": "fallback (B200_PROFILING.md)"}


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


# ------------------------------------------------------------------------------------------------
# CPU baseline: the oracle port of the reference, all host threads, bounded sample of the workload
# ------------------------------------------------------------------------------------------------
def cpu_baseline(states, pos, weights, threshold, seconds):
    import oracle.oracle as orc
    n, L = states.shape
    cores = orc.hardware_threads()
    # calibrate on a small column subset, then size the sample for ~`seconds` of work
    Lc = min(L, 400)
    t0 = time.time(); orc.mi_scan(states[:, :Lc], weights, pos[:Lc], threshold=threshold, circle_size=Lc); dt = time.time() - t0
    rate = (Lc * (Lc - 1) / 2) / max(dt, 1e-6)
    Ls = int(min(L, max(Lc, (2 * seconds * rate) ** 0.5)))
    t0 = time.time()
    r = orc.mi_scan(states[:, :Ls], weights, pos[:Ls], threshold=threshold, circle_size=Ls)
    e = orc.sort_edges(r["edges"]); orc.aracne_closed(e)
    dt = time.time() - t0
    pairs = Ls * (Ls - 1) / 2
    return {"value": pairs / dt / 1e9, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": "first %d of %d columns (all %d samples): %d pairs, scan+sort+ARACNE in %.1f s" % (Ls, L, n, pairs, dt)}


def run_reference(args):
    """--impl reference: the reference's CPU path (oracle port: the reference binary needs Boost + TBB,
    absent here), all host threads, bounded sample per step."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle.oracle as orc
    states, pos, _ = make_workload(args)
    n, L = states.shape
    Ls = min(L, 1500)          # bounded sample: ~1.1M pairs per step
    sub = np.ascontiguousarray(states[:, :Ls])
    thr = 0.05
    times = []
    for it in range(args.warmup + args.steps):
        t0 = time.time()
        counts = orc.neighbour_counts(sub, int(0.9 * Ls), True)
        w = orc.weights_from_counts(counts)
        r = orc.mi_scan(sub, w, pos[:Ls], threshold=thr, circle_size=Ls)
        e = orc.sort_edges(r["edges"]); orc.aracne_closed(e)
        if it >= args.warmup:
            times.append(time.time() - t0)
    pairs = Ls * (Ls - 1) / 2
    ms = 1e3 * sum(times) / len(times)
    v = pairs / (ms / 1e3) / 1e9
    sample = "first %d of %d columns, all %d samples (%d pairs per step)" % (Ls, L, n, pairs)
    print(json.dumps({"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
                      "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                      "dtype": "f64", "data": "synthetic",
                      "config": {"workload": "3000x100000 synthetic SNP alignment, MI scan + ARACNE (CPU port on a bounded sample)",
                                 "n": n, "L": L, "sample": sample},
                      "cpu_baseline": {"value": v, "unit": UNIT, "cores": orc.hardware_threads(), "kind": "port", "sample": sample},
                      "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}))


# ------------------------------------------------------------------------------------------------
# GPU arm
# ------------------------------------------------------------------------------------------------
class _DevView:
    """Wraps a raw device pointer so torch can view it (__cuda_array_interface__)."""

    def __init__(self, ptr, nbytes):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 2}


def dev_bytes(torch, ptr, nbytes, device):
    if nbytes == 0:
        return torch.empty(0, dtype=torch.uint8, device=device)
    return torch.as_tensor(_DevView(ptr, nbytes), device=device)


def run_b200(args):
    import torch
    import torch.distributed as dist
    from spydrpick_b200 import Context
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

    # ---- workload: rank 0 generates, alignment broadcast ONCE over NCCL (outside the timed region)
    n, L = args.n, args.L
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
    # ONE stream for torch's own ops (allocator fills, NCCL hand-offs, CUDA events) and the library's kernels: torch's
    # default current stream is the legacy stream 0, which spk_set_stream reads as "use your own stream"
    stream = torch.cuda.Stream(device)
    torch.cuda.set_stream(stream)
    ctx.set_stream(stream.cuda_stream)
    ctx.set_alignment(states, pos)
    min_ident = int(0.9 * float(L))
    counts = ctx.neighbour_counts(min_ident, True)
    weights = 1.0 / counts.astype(np.float64)
    ctx.set_weights(weights)
    # pinned save threshold from one K2b estimate (untimed; the reference's estimate is clock-seeded)
    t0 = time.time()
    threshold = ctx.determine_mi_threshold(mi_values, iterations=10, seed=args.seed)
    t_threshold = time.time() - t0
    if world > 1:
        tt = torch.tensor([threshold], dtype=torch.float64, device=device); dist.broadcast(tt, 0); threshold = float(tt.item())
    max_edges = int(mi_values * 2.5) + 65536
    EPS = 2.220446049250313e-16

    def merge_across_ranks(r):
        """Edge allgather over NCCL/NVLink before ARACNE (+ colmax allreduce-max). Returns (keepalive, ptr, n_edges)."""
        from spydrpick_b200 import dist as sdist
        ne = r["n_edges"]
        if world == 1:
            return None, r["d_edges"], ne
        mine = dev_bytes(torch, r["d_edges"], ne * 16, device)
        merged, total, counts = sdist.allgather_edges(mine, ne)
        sdist.allreduce_colmax(dev_bytes(torch, r["d_colmax"], L * 8, device).view(torch.float64))
        # every rank sorted its own edges inside the scan: the gathered list is G sorted runs -> G-1 merges, not a full sort.
        # torch filled `merged` on ITS current stream (the legacy default stream, which the library's non-blocking stream
        # does not wait for): the list must be complete before the pointer is handed over
        torch.cuda.current_stream(device).synchronize()
        ctx.merge_sorted_runs_device(merged.data_ptr(), counts)
        return merged, merged.data_ptr(), total

    def sharded_counts():
        """K1: every rank evaluates its share of the sample-tile pairs; partial counts are summed over NCCL."""
        if world == 1:
            return ctx.neighbour_counts(min_ident, True)
        from spydrpick_b200 import dist as sdist
        return sdist.allreduce_neighbour_counts(ctx.neighbour_counts(min_ident, True, part_index=rank, n_parts=world), device)

    def step_device():
        """alignment resident in HBM; results stay in HBM."""
        c = sharded_counts()
        ctx.set_weights(1.0 / c.astype(np.float64))
        r = ctx.mi_scan(threshold, max_edges=max_edges, sort_desc=True, n_terms=args.n_terms, engine=args.engine,
                        part_index=rank, n_parts=world, keep_on_device=True)
        keep, d_edges, ne = merge_across_ranks(r)
        flags = torch.zeros(max(ne, 1), dtype=torch.uint8, device=device)
        ctx.aracne_device(d_edges, ne, flags.data_ptr(), EPS, part_index=rank, n_parts=world)
        if world > 1:
            from spydrpick_b200 import dist as sdist
            flags = sdist.allgather_flag_shards(flags, ne)       # edge-range partition of K3 -> every rank
        return r, ne, int(flags[:ne].sum().item())

    from spydrpick_b200.capi import pinned_copy, pinned_empty
    states_pin = pinned_copy(states)                  # the caller's alignment lives in page-locked host memory
    pos_pin = pinned_copy(pos)
    states_pin_t = torch.from_numpy(states_pin)       # same page-locked memory, as a tensor for the row-shard upload

    host_edges = host_flags = side = None

    def step_e2e():
        """host buffers through the C-ABI: H2D of the alignment (pinned), D2H of edges + flags inside the step."""
        tw = [time.perf_counter()]
        if world == 1:
            ctx.set_alignment(states_pin, pos_pin)
        else:
            # every rank uploads n/G rows over its own PCIe link; the slices are all-gathered over NVLink
            from spydrpick_b200 import dist as sdist
            ra, rb = sdist.row_shard(n, world, rank)
            d_states = sdist.allgather_alignment_rows(states_pin_t[ra:rb].reshape(-1), n, L, device)
            ctx.set_alignment_device(d_states.data_ptr(), n, L, pos_pin)
            del d_states
        tw.append(time.perf_counter())
        c = sharded_counts()
        ctx.set_weights(1.0 / c.astype(np.float64)); tw.append(time.perf_counter())
        if world == 1:
            r = ctx.mi_scan(threshold, max_edges=max_edges, sort_desc=True, n_terms=args.n_terms, engine=args.engine,
                            keep_on_device=False, copy=False, async_download=True); tw.append(time.perf_counter())
            edges = r["edges"]                        # view of the library's pinned result buffer (download in flight)
            nonlocal host_flags
            if host_flags is None or host_flags.shape[0] < len(edges):
                host_flags = pinned_empty(int(len(edges) * 1.25) + 4096, np.uint8)
            flags = ctx.aracne(edges, EPS, out=host_flags)   # same pointer -> device copy reused; overlaps the edge download
            t_ar = time.perf_counter()
            ctx.wait_download()                       # edges + MI-without-gaps lists are now complete on the host
            t_wd = time.perf_counter()
            ndirect = int(np.count_nonzero(flags))    # the step's result is read on the host
            if os.environ.get("SPK_BENCH_E2E_BREAKDOWN"):
                tma = ctx.timings()
                sys.stderr.write("e2e aracne phase ms: spk_aracne %.2f (device: CSR build %.2f, K3 %.2f) | wait_download %.2f | count flags %.2f\n" % (
                    1e3 * (t_ar - tw[-1]), tma["aracne_build_ms"], tma["aracne_ms"], 1e3 * (t_wd - t_ar), 1e3 * (time.perf_counter() - t_wd)))
            h2d = states.nbytes + pos.nbytes + 8 * n + len(flags)
            d2h = edges.nbytes + r["edges_wo_gaps"].nbytes + 8 * L + 8 * n + len(flags)
        else:
            # N > 1: the partial edge lists never visit the host -- NCCL allgather over NVLink, device sort + K3, then ONE
            # D2H of the merged list and the flags into pinned host buffers (what a caller of the multi-GPU path receives)
            r = ctx.mi_scan(threshold, max_edges=max_edges, sort_desc=True, n_terms=args.n_terms, engine=args.engine,
                            part_index=rank, n_parts=world, keep_on_device=True); tw.append(time.perf_counter())
            keep, d_edges, ne = merge_across_ranks(r)   # returns with the merged, sorted list complete in HBM
            nonlocal host_edges, side
            if host_edges is None or host_edges.numel() < ne * 16:
                host_edges = torch.empty(int(ne * 16 * 1.25) + 4096, dtype=torch.uint8).pin_memory()
                host_flags = torch.empty(int(ne * 1.25) + 4096, dtype=torch.uint8).pin_memory()
                side = torch.cuda.Stream(device)
            if rank == 0:                             # the caller's process receives the merged list: download overlaps K3
                with torch.cuda.stream(side):
                    host_edges[:ne * 16].copy_(dev_bytes(torch, d_edges, ne * 16, device), non_blocking=True)
            dflags = torch.zeros(max(ne, 1), dtype=torch.uint8, device=device)
            ctx.aracne_device(d_edges, ne, dflags.data_ptr(), EPS, part_index=rank, n_parts=world)
            from spydrpick_b200 import dist as sdist
            dflags = sdist.allgather_flag_shards(dflags, ne)
            if rank == 0:
                host_flags[:ne].copy_(dflags[:ne], non_blocking=True)
            torch.cuda.synchronize()
            ndirect = int(host_flags[:ne].numpy().sum()) if rank == 0 else 0
            h2d = (rb - ra) * L + pos.nbytes + 8 * n  # per rank; whole job = the alignment once + G x (pos, weights)
            d2h = ne * 16 + ne + 8 * L + 8 * n        # rank 0
        tw.append(time.perf_counter())
        if os.environ.get("SPK_BENCH_E2E_BREAKDOWN"):
            tmx = ctx.timings()
            sys.stderr.write("e2e wall ms: set_alignment %.1f (upload %.1f prep %.1f) | counts+weights %.1f | scan %.1f | aracne %.1f\n" % (
                1e3 * (tw[1] - tw[0]), tmx["upload_ms"], tmx["prep_ms"], 1e3 * (tw[2] - tw[1]), 1e3 * (tw[3] - tw[2]), 1e3 * (tw[4] - tw[3])))
        return h2d, d2h

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
        e0.record(stream)
        for _ in range(steps):
            out = fn()
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
        return ms, out, launches, (sampler.summary() if sampler else None)

    ms, (r, ne, ndirect), launches, clocks = timed(step_device, args.steps, args.warmup, with_clocks=True)
    tm = ctx.timings()
    value = pairs_total / (ms / 1e3) / 1e9

    # roofline of the dominant kernel (K2 tcgen05 GEMM): algorithmic flops = 2*16*n per pair
    pk = peaks()
    tiles = tm["gemm_tiles"]
    pairs_launch = tm["gemm_pairs"]                      # pairs one launch (this rank) covers, incl. padding / masked diagonal halves
    flops = 2.0 * 16.0 * n * pairs_launch
    # the int8 engine launches the kernel twice per scan: a small colmax seed pass (seed_ms) and the main pass the
    # tile / pair counts refer to; the roofline is that of the main-pass launch
    k_ms = tm["gemm_ms"] - tm.get("seed_ms", 0.0)
    achieved = flops / (k_ms / 1e3) / 1e12 if k_ms > 0 else 0.0
    i8 = args.engine in (0, 3, 4)
    n_pad = (n + 127) // 128 * 128
    executed = 2.0 * 128 * 256 * n_pad * tiles * (1 if i8 else (args.n_terms or 2))   # tensor ops actually issued
    executed_tops = executed / (k_ms / 1e3) / 1e12 if k_ms > 0 else 0.0
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE launch from the committed ncu --set full capture
    # (profiles/r01f_ncu_full_summary.csv); only valid for the workload it was captured on
    traffic = NCU_TRAFFIC.get(args.engine if args.engine else 4) if (i8 and world == 1 and n == 3000 and L == 100000) else None
    roofline = {"bound": "tensor", "achieved": achieved, "peak": pk["bf16_tflops"], "unit": "TFLOP/s",
                "frac": achieved / pk["bf16_tflops"], "traffic": traffic,
                "kernel": "mi_i8_kernel<2> (main pass)" if i8 else "mi_gemm_kernel", "kernel_ms": k_ms, "seed_pass_ms": tm.get("seed_ms", 0.0),
                "peak_source": pk["source"],
                "executed_tops": executed_tops,
                "executed_frac_of_pipe_peak": executed_tops / (pk["bf16_burst_tflops"] * (2.0 if i8 else 1.0)),
                "executed_frac_of_nominal_pipe": executed_tops / (4500.0 if i8 else 2250.0),
                "ncu_tensor_pipe_active_pct": NCU_TENSOR_PCT.get(args.engine if args.engine else 4) if traffic else None,   # main pass, profiles/r01f_ncu_full_summary.csv
                "note": ("achieved = SURVEY 8(d)'s algorithmic 32*n flop/pair vs the measured bf16 peak; the int8 compact engine "
                         "issues FEWER tensor ops than that (only the one-hot rows a column needs, u8 x u8 -> s32 at 2x the bf16 "
                         "rate), so frac exceeds 1 -- the honest utilisation figures are executed_frac_of_pipe_peak = issued "
                         "int8 ops / (2 x measured bf16 burst peak), executed_frac_of_nominal_pipe = issued / 4.5 POP/s, and "
                         "ncu's sm__pipe_tensor_cycles_active (profiles/r01f_ncu_full_summary.csv)") if i8 else
                        "algorithmic 32*n flop/pair; executed tensor flops are n_terms x that (bf16 weight split)"}

    e2e = None
    if not args.no_e2e:
        ems, (h2d, d2h), _, _ = timed(step_e2e, max(1, min(args.steps, 3)), 2)
        e2e = {"value": pairs_total / (ems / 1e3) / 1e9, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
               "ms_per_step": ems}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline(states, pos, weights, threshold, args.cpu_sample_seconds)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
                "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "u8 x u8 -> s32 tensor filter + f64 refine" if i8 else "bf16 tensor filter + f64 refine",
                "data": "synthetic",
                "config": {"workload": "%dx%d synthetic SNP alignment: K1 reweighting + weighted MI scan + sort + ARACNE" % (n, L),
                           "n": n, "L": L, "pairs": pairs_total, "saved_edges": int(ne), "direct_edges": int(ndirect), "mi_threshold": threshold,
                           "engine": ("int8 compact one-hot, 16-bit weights as {hi, lo} rows" if args.engine == 3 else "int8 compact one-hot, factorised byte weights a*b") if i8 else "dense bf16 one-hot, n_terms=%d" % (args.n_terms or 2),
                           "l2_policy": "inputs (>1.5 GB one-hot operands) larger than L2",
                           "threshold_estimate_s": t_threshold, "parallelism": "tile-space partition x%d + NCCL edge allgather" % world},
                "clocks": clocks, "gpu_launches": int(launches),
                "stages_ms": {k: tm[k] for k in ("hamming_ms", "stats_ms", "expand_ms", "gemm_ms", "refine_ms", "compact_ms",
                                                 "sort_ms", "aracne_build_ms", "aracne_ms")},
                "candidates": r["n_candidates"], "max_filter_error": r["max_filter_error"],
                "roofline": roofline}
        if e2e:
            e2e["path"] = ("spk_set_alignment (pinned host states) -> K1 -> scan -> sort -> ARACNE, edges + flags downloaded" if world == 1 else
                           "per rank: H2D of n/G alignment rows + NVLink all-gather -> spk_set_alignment_device -> sharded K1/scan -> "
                           "per-rank edge sort -> NCCL edge all-gather -> run merge -> sharded ARACNE; rank 0 downloads the merged edges + flags "
                           "(h2d bytes are per rank, d2h bytes are rank 0's)")
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
