"""Large-size check of spk_merge_sorted_runs_device against spk_sort_edges_device (one GPU)."""
import sys
import numpy as np
import torch
from spydrpick_b200.capi import Context, EDGE_DTYPE

n = int(sys.argv[1]) if len(sys.argv) > 1 else 10_000_000
runs = int(sys.argv[2]) if len(sys.argv) > 2 else 2
ctx = Context(device=0)
rng = np.random.default_rng(3)
e = np.zeros(n, dtype=EDGE_DTYPE)
e["i"] = rng.integers(0, 100000, n); e["j"] = e["i"] + 1 + rng.integers(0, 100000, n)
e["mi"] = rng.random(n)
full = torch.from_numpy(e.view(np.uint8).copy()).cuda()
ctx.sort_edges_device(full.data_ptr(), n)
bounds = [n * k // runs for k in range(runs + 1)]
parts = []
for k in range(runs):
    t = torch.from_numpy(e[bounds[k]:bounds[k + 1]].view(np.uint8).copy()).cuda()
    ctx.sort_edges_device(t.data_ptr(), bounds[k + 1] - bounds[k])
    parts.append(t)
cat = torch.cat(parts)
torch.cuda.synchronize()
ctx.merge_sorted_runs_device(cat.data_ptr(), [bounds[k + 1] - bounds[k] for k in range(runs)])
torch.cuda.synchronize()
print("equal:", bool(torch.equal(cat, full)), "sort_ms(merge):", ctx.timings()["sort_ms"])
