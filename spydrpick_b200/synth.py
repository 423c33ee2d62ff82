"""Synthetic bacterial-like alignments: population structure, mostly biallelic SNP
columns, per-column gap rates, short-range LD by pattern copying, planted long-range couplings.

States are the reference's codes a=0,t=1,c=2,g=3,GAP=4
(ref: externals/apegrunt/src/StateVector_state_types.cpp:33-44). Deterministic given the seed.
"""
import numpy as np

STATE_CHARS = np.frombuffer(b"ATCG-", dtype=np.uint8)


def generate_alignment(n, L, seed=20260102, n_lineages=None, ld_copy_prob=0.6, n_planted=2000,
                       mean_spacing=22, chunk=8192, all_pass_filter=True):
    """Returns (states uint8[n, L], pos uint64[L] strictly increasing original positions, genome_size)."""
    rng = np.random.Generator(np.random.PCG64(seed))
    G = n_lineages or max(8, n // 100)
    # two-level structure: lineages -> clone groups (~10 samples each) -> samples
    n_clones = max(G, n // 10)
    clone_lineage = rng.integers(0, G, size=n_clones)
    lineage_flip = rng.uniform(0.05, 0.35, size=G).astype(np.float32)
    clone_flip_p = lineage_flip[clone_lineage][:, None]           # [n_clones,1] clone centroid vs lineage centroid
    clone = rng.integers(0, n_clones, size=n)
    sample_flip = rng.uniform(0.002, 0.12, size=n_clones).astype(np.float32)
    flip_p = sample_flip[clone][:, None]                          # [n,1] sample vs clone centroid

    states = np.empty((n, L), dtype=np.uint8)
    # column classes: 90% biallelic, 8% triallelic, 2% four-allelic
    ncls = rng.choice([2, 3, 4], size=L, p=[0.90, 0.08, 0.02])
    maf = np.clip(rng.beta(0.5, 2.0, size=L), 0.02 if all_pass_filter else 0.0, 0.5)
    gap_rate = np.clip(rng.beta(1.0, 40.0, size=L), 0.0, 0.12)
    gap_rate[rng.random(L) < 0.30] = 0.0                            # 30% gap-free columns
    # allele labels: a random permutation of the 4 nucleotides per column
    perm = np.argsort(rng.random((L, 4)), axis=1).astype(np.uint8)  # [L,4]
    # cumulative allele probabilities per column (major first)
    probs = np.zeros((L, 4), dtype=np.float64)
    probs[:, 0] = 1.0 - maf
    for k in (2, 3, 4):
        m = ncls == k
        share = maf[m][:, None] * np.array({2: [1.0], 3: [0.7, 0.3], 4: [0.6, 0.25, 0.15]}[k])[None, :]
        probs[m, 1:k] = share
    cum = np.cumsum(probs, axis=1).astype(np.float32)

    def draw(cols, shape_rows):
        """i.i.d. allele indices for the given columns: [rows, len(cols)] in 0..3"""
        u = rng.random((shape_rows, len(cols)), dtype=np.float32)
        c = cum[cols]
        return ((u >= c[None, :, 0]).astype(np.uint8) + (u >= c[None, :, 1]) + (u >= c[None, :, 2])).astype(np.uint8)

    for c0 in range(0, L, chunk):
        c1 = min(L, c0 + chunk)
        cols = np.arange(c0, c1)
        cent = draw(cols, G)                                        # [G, Lc] lineage centroids
        ccent = cent[clone_lineage]                                 # [n_clones, Lc]
        cflip = rng.random((n_clones, c1 - c0), dtype=np.float32) < clone_flip_p
        ccent = np.where(cflip, draw(cols, n_clones), ccent)
        x = ccent[clone]                                            # [n, Lc]
        flip = rng.random((n, c1 - c0), dtype=np.float32) < flip_p
        fresh = draw(cols, n)
        x = np.where(flip, fresh, x)
        states[:, c0:c1] = x                                        # allele INDICES (0 = major) until relabelled below

    # short-range LD: column c copies the sample pattern of its run head with prob ld_copy_prob (5% noise)
    copy = rng.random(L) < ld_copy_prob
    copy[0] = False
    head = np.where(~copy, np.arange(L), 0)
    head = np.maximum.accumulate(head)
    cc = np.nonzero(copy)[0]
    for k0 in range(0, len(cc), chunk):
        idx = cc[k0:k0 + chunk]
        keep_own = rng.random((n, len(idx)), dtype=np.float32) < 0.02
        # copy the allele-index pattern (not the labels), folded into the copying column's own allele count
        src = np.minimum(states[:, head[idx]], (ncls[idx] - 1)[None, :].astype(np.uint8))
        states[:, idx] = np.where(keep_own, states[:, idx], src)

    # planted long-range couplings
    n_planted = min(n_planted, L // 50)
    if n_planted > 0:
        a = rng.choice(L, size=n_planted, replace=False)
        b = (a + rng.integers(L // 4, L // 2, size=n_planted)) % L
        relabel = np.argsort(rng.random((n_planted, 4)), axis=1).astype(np.uint8)
        noise = rng.random((n, n_planted), dtype=np.float32) < 0.10
        src = states[:, a]
        mapped = relabel[np.arange(n_planted)[None, :], src] % ncls[b][None, :].astype(np.uint8)
        states[:, b] = np.where(noise, states[:, b], mapped)

    # allele indices -> nucleotide labels (per-column permutation), then gaps last (so copied columns keep
    # their own gap rates)
    for c0 in range(0, L, chunk):
        c1 = min(L, c0 + chunk)
        states[:, c0:c1] = perm[c0:c1][np.arange(c1 - c0)[None, :], states[:, c0:c1]]
        g = rng.random((n, c1 - c0), dtype=np.float32) < gap_rate[c0:c1][None, :].astype(np.float32)
        blk = states[:, c0:c1]
        blk[g] = 4

    spacing = rng.geometric(1.0 / mean_spacing, size=L).astype(np.uint64)
    pos = np.cumsum(spacing) - 1
    genome_size = int(pos[-1]) + int(mean_spacing)
    return states, pos.astype(np.uint64), genome_size


def to_fasta(states, path, width=0):
    """Write states as FASTA (sequence names s0..s{n-1})."""
    with open(path, "wb") as f:
        for i in range(states.shape[0]):
            f.write(b">s%d\n" % i)
            f.write(STATE_CHARS[states[i]].tobytes())
            f.write(b"\n")
