"""One-GPU emulation of the multi-GPU edge exchange: scan the partitions one after the other (each sorted by the scan),
concatenate in rank order, merge the runs, compare with the unpartitioned sorted scan."""
import sys
import numpy as np
import torch
from spydrpick_b200.capi import Context
from spydrpick_b200.synth import generate_alignment

n = int(sys.argv[1]); L = int(sys.argv[2]); parts = int(sys.argv[3])
s, pos, _ = generate_alignment(n, L, seed=20260102)
ctx = Context(device=0)
ctx.set_stream(torch.cuda.current_stream(0).cuda_stream)
ctx.set_alignment(s, pos)
c = ctx.neighbour_counts(int(0.9 * L), True)
ctx.set_weights(1.0 / c.astype(np.float64))
thr = ctx.determine_mi_threshold(min(10_000_000, 100 * L), iterations=10, seed=1)
cap = int(min(10_000_000, 100 * L) * 2.5) + 65536


class V:
    def __init__(self, p, nb):
        self.__cuda_array_interface__ = {"shape": (nb,), "typestr": "|u1", "data": (p, False), "version": 3}


whole = ctx.mi_scan(thr, max_edges=cap, sort_desc=True, keep_on_device=True)
ref = torch.as_tensor(V(whole["d_edges"], whole["n_edges"] * 16), device="cuda:0").clone()
print("whole", whole["n_edges"], flush=True)
runs, cnt = [], []
for k in range(parts):
    r = ctx.mi_scan(thr, max_edges=cap, sort_desc=True, part_index=k, n_parts=parts, keep_on_device=True)
    print("part", k, r["n_edges"], flush=True)
    runs.append(torch.as_tensor(V(r["d_edges"], r["n_edges"] * 16), device="cuda:0").clone()); cnt.append(r["n_edges"])
cat = torch.cat(runs)
torch.cuda.synchronize()
print("merging", cnt, flush=True)
ctx.merge_sorted_runs_device(cat.data_ptr(), cnt)
torch.cuda.synchronize()
print("equal:", bool(torch.equal(cat, ref)), "merge ms:", ctx.timings()["sort_ms"])
