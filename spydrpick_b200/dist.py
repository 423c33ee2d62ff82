"""The CALLER-DRIVEN multi-GPU recipe over torch.distributed (NCCL on GPUs, gloo in CPU tests). The product path no longer
needs it: the exchange lives inside the library (csrc/spk_comm.cu; `spk_comm_init_rank` + collective `spk_mi_scan` /
`spk_aracne`, what bench.py and the CLI's --gpus use). Kept for hosts that own their communication and for the gloo tests of
its CPU-side logic (tests/test_dist_gloo.py).

The scan shards the upper-triangle tile space (`part_index = rank`); the one exchange step before ARACNE
is an all-gather of the per-rank edge lists plus a max-reduction of colmax. ARACNE shards
the sorted edge range; its flag shards are all-gathered. Tensors are uint8 views of spk_edge[] (16 B).
"""
import torch
import torch.distributed as dist

EDGE_BYTES = 16


def allgather_edges(mine: torch.Tensor, n_edges: int, group=None):
    """mine: 1-D uint8 tensor whose first n_edges*16 bytes are this rank's edges.
    Returns (merged uint8 tensor [total*16], total, per-rank counts) -- identical on every rank (rank order)."""
    world = dist.get_world_size(group)
    dev = mine.device
    cnt = torch.tensor([n_edges], dtype=torch.int64, device=dev)
    allc = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(allc, cnt, group=group)
    counts = [int(c.item()) for c in allc]
    mx = max(max(counts), 1)
    pad = torch.zeros(mx * EDGE_BYTES, dtype=torch.uint8, device=dev)
    if n_edges:
        pad[:n_edges * EDGE_BYTES].copy_(mine[:n_edges * EDGE_BYTES])
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    merged = torch.cat([parts[k][:counts[k] * EDGE_BYTES] for k in range(world)])
    return merged, sum(counts), counts


def allreduce_neighbour_counts(partial, device=None, group=None):
    """K1 sharded over sample-tile pairs (spk_neighbour_counts_part): element-wise SUM of the per-rank partial counts.
    partial: numpy uint64 [n] (host). Returns numpy uint64 [n], identical on every rank."""
    import numpy as np
    t = torch.from_numpy(partial.astype(np.int64))
    if device is not None:
        t = t.to(device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return t.cpu().numpy().astype(np.uint64)


def allreduce_colmax(colmax: torch.Tensor, group=None):
    """In-place element-wise max over ranks of the float64 per-column maxima (MI_solver::join, mi.hpp:355-361)."""
    dist.all_reduce(colmax, op=dist.ReduceOp.MAX, group=group)
    return colmax


def shard_bounds(n_edges: int, world: int):
    """Edge-range partition used by spk_aracne(part_index, n_parts)."""
    return [n_edges * k // world for k in range(world + 1)]


def allgather_flag_shards(flags: torch.Tensor, n_edges: int, group=None):
    """flags: uint8 [>= n_edges] with this rank's shard filled in. Returns the full flag vector on every rank."""
    world = dist.get_world_size(group); rank = dist.get_rank(group)
    b = shard_bounds(n_edges, world)
    mx = max(max(b[k + 1] - b[k] for k in range(world)), 1)
    mine = torch.zeros(mx, dtype=torch.uint8, device=flags.device)
    mine[:b[rank + 1] - b[rank]].copy_(flags[b[rank]:b[rank + 1]])
    parts = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(parts, mine, group=group)
    return torch.cat([parts[k][:b[k + 1] - b[k]] for k in range(world)])


def row_shard(n: int, world: int, rank: int):
    """Rows of the alignment rank `rank` uploads: equal blocks of ceil(n / world) rows (the last may be short or empty)."""
    per = (n + world - 1) // world
    return min(n, rank * per), min(n, (rank + 1) * per)


def allgather_alignment_rows(my_rows: torch.Tensor, n: int, L: int, device=None, group=None):
    """"Alignment broadcast once"  as G small uploads + one all-gather: `my_rows` is this rank's
    row_shard() slice of the row-major uint8[n][L] states in (pinned) host memory. It goes up this rank's own PCIe link,
    the slices are all-gathered over NVLink, and every rank ends with the full uint8[n, L] tensor on its device
    (feed its data_ptr() to spk_set_alignment_device)."""
    world = dist.get_world_size(group); rank = dist.get_rank(group)
    per = (n + world - 1) // world
    a, b = row_shard(n, world, rank)
    assert my_rows.dtype == torch.uint8 and my_rows.numel() == (b - a) * L
    dev = device if device is not None else my_rows.device
    full = torch.empty((world * per, L), dtype=torch.uint8, device=dev)
    mine = full[rank * per:(rank + 1) * per]
    if b > a:
        mine[:b - a].copy_(my_rows.view(b - a, L), non_blocking=True)
    if b - a < per:
        mine[b - a:].zero_()
    dist.all_gather_into_tensor(full, mine, group=group)
    if full.is_cuda:
        # the library copies on ITS stream: the gathered rows must be complete before the pointer is handed over
        torch.cuda.current_stream(full.device).synchronize()
    return full[:n]
