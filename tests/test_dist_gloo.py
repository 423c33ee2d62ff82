"""world_size-2 gloo tests (CPU) of the multi-GPU host logic: edge all-gather, colmax max-reduce, ARACNE
flag shards. The per-rank scan results are produced by the oracle on a column-block partition of the pair
space, the merged pipeline must equal the single-process one."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import oracle.oracle as orc
from spydrpick_b200 import dist as sdist
from spydrpick_b200.synth import generate_alignment


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    s, pos, _ = generate_alignment(60, 240, seed=11)
    w = orc.weights_from_counts(orc.neighbour_counts(s, int(0.9 * 240)))
    # K1 sharded over sample pairs: partial neighbour counts (partition 0 carries the sample itself) summed over ranks
    counts_full = orc.neighbour_counts(s, int(0.9 * 240), quirks=orc.QUIRK_LASTBLOCK)   # the pair rule only (the sample-0 extra counts ride on partition 0)
    owner = (np.add.outer(np.arange(60) // 8, np.arange(60) // 8 * 3)) % world
    owner = np.triu(owner, 1); owner = owner + owner.T                       # symmetric owner of the unordered pair
    partial = np.zeros(60, dtype=np.uint64)
    for a in range(60):
        for b_ in range(60):
            if a != b_ and owner[a, b_] == rank and orc.neighbour_counts(s[[a, b_]], int(0.9 * 240), quirks=orc.QUIRK_LASTBLOCK)[0] == 2:
                partial[a] += 1
    if rank == 0:
        partial += 1
    counts_sum = sdist.allreduce_neighbour_counts(partial)
    k1_ok = np.array_equal(counts_sum, counts_full.astype(np.uint64))
    thr = 0.05
    full = orc.mi_scan(s, w, pos, threshold=thr, circle_size=240, threads=2)
    # emulate the tile-space partition: rank r owns pairs whose i-block (32 columns) index % world == r
    e = full["edges"]
    mine = e[(e["i"] // 32) % world == rank]
    colmax = np.full(240, -np.finfo(np.float64).max)
    masks = orc.state_masks(s)
    for i in range(240):
        if (i // 32) % world != rank:
            continue
        for j in range(i + 1, 240):
            m = orc.mi_pair(s, w, i, j, masks=masks)
            d = abs(int(pos[i]) - int(pos[j]))
            if min((240 - d) % (2 ** 64), d) > 0:      # circular distance with circle = L, size_t wrap included
                colmax[i] = max(colmax[i], m); colmax[j] = max(colmax[j], m)
    t = torch.from_numpy(np.frombuffer(mine.tobytes(), dtype=np.uint8).copy())
    merged, total, counts = sdist.allgather_edges(t, len(mine))
    cm = sdist.allreduce_colmax(torch.from_numpy(colmax.copy()))
    me = orc.sort_edges(np.frombuffer(merged.numpy().tobytes(), dtype=orc.EDGE_DTYPE))
    # ARACNE edge-range shards
    b = sdist.shard_bounds(total, world)
    flags_full = orc.aracne_closed(me)
    shard = np.zeros(total, dtype=np.uint8); shard[b[rank]:b[rank + 1]] = flags_full[b[rank]:b[rank + 1]]
    flags = sdist.allgather_flag_shards(torch.from_numpy(shard), total).numpy()
    ref = orc.sort_edges(e)
    # alignment row shards: every rank contributes its rows, all ranks end with the whole alignment (uneven: 61 rows)
    s61 = np.ascontiguousarray(np.vstack([s, s[:1]]))
    ra, rb = sdist.row_shard(61, world, rank)
    whole = sdist.allgather_alignment_rows(torch.from_numpy(s61[ra:rb].copy()).reshape(-1), 61, 240)
    rows_ok = np.array_equal(whole.numpy(), s61) and sdist.row_shard(5, 8, 7) == (5, 5)
    ok = (total == len(e) and sum(counts) == total and rows_ok and me.tobytes() == ref.tobytes()
          and np.array_equal(flags, orc.aracne_closed(ref)) and np.allclose(cm.numpy(), full["colmax"], rtol=1e-12)
          and k1_ok)
    out[rank] = bool(ok)
    dist.destroy_process_group()


def test_two_rank_merge_equals_single_process():
    world = 2
    mgr = mp.Manager(); out = mgr.dict()
    mp.spawn(_worker, args=(world, _free_port(), out), nprocs=world, join=True)
    assert dict(out) == {0: True, 1: True}
