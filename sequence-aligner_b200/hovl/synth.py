"""Synthetic read sets of BASELINE.json's configs.

All reads are forward strand (the reference has no reverse complement, src/ObjectStore.scala:
121-123), sampled uniformly from an iid-uniform ACGT genome with substitution errors only.
Everything is seeded, so every rank / every run sees the same bytes.
"""
import numpy as np

_ACGT = np.frombuffer(b"ACGT", np.uint8)

# name -> (genome bp, reads, read bp, substitution rate, k, seeds)
WORKLOADS = {
    # configs[2]: 200 k x 1 kb over 5 Mbp is 40x coverage (BASELINE says 20x; the read shape wins)
    "K3": dict(genome=5_000_000, reads=200_000, length=1000, sub=0.01, k=16, seeds=(1001, 1002, 1003)),
    # configs[3]: 10 M x 150 bp over 100 Mbp (15x), no error rate named -> error-free
    "K4": dict(genome=100_000_000, reads=10_000_000, length=150, sub=0.0, k=12, seeds=(2001, 2002, 2003)),
    # configs[4]: 2 M x 10 kb over 1 Gbp (20x), 1 % substitutions
    "K5": dict(genome=1_000_000_000, reads=2_000_000, length=10_000, sub=0.01, k=16, seeds=(3001, 3002, 3003)),
}


def scaled(name, scale=1.0):
    """The named workload with genome and read count multiplied by `scale` (coverage, read
    length, error rate and k unchanged)."""
    w = dict(WORKLOADS[name])
    w["genome"] = max(int(round(w["genome"] * scale)), 4 * w["length"])
    w["reads"] = max(int(round(w["reads"] * scale)), 2)
    w["name"] = name if scale == 1.0 else "%s*%g" % (name, scale)
    return w


def generate(w, out=None, chunk_reads=None):
    """Returns (bases uint8[n*L] ASCII, offsets uint64[n+1]).  `out` may be a preallocated
    uint8 array (e.g. a view of pinned memory) of n*L bytes."""
    G, n, L, sub = w["genome"], w["reads"], w["length"], w["sub"]
    sg, ss, se = w["seeds"]
    genome = np.random.default_rng(sg).integers(0, 4, G, dtype=np.uint8)
    starts = np.random.default_rng(ss).integers(0, G - L + 1, n, dtype=np.int64)
    rng_e = np.random.default_rng(se)
    if out is None:
        out = np.empty(n * L, np.uint8)
    assert out.size == n * L
    step = chunk_reads or max(1, (1 << 24) // L)
    col = np.arange(L, dtype=np.int64)[None, :]
    for lo in range(0, n, step):
        hi = min(n, lo + step)
        seg = genome[starts[lo:hi, None] + col]                       # (m, L) codes 0..3
        if sub > 0:
            flip = rng_e.random(seg.shape, dtype=np.float32) < sub
            add = rng_e.integers(1, 4, seg.shape, dtype=np.uint8)
            seg = np.where(flip, (seg + add) & 3, seg)
        out[lo * L:hi * L] = _ACGT[seg].reshape(-1)
    offsets = (np.arange(n + 1, dtype=np.uint64) * np.uint64(L))
    return out, offsets


def shape_flags(w, shape):
    """Settings overrides for a workload.  `default` = the Project4 defaults (tuned for 100 bp
    reads: --max-ignore 90 and --max-collisions 222 reject nearly every true overlap of longer
    reads); `wide` = --max-ignore L --max-collisions 2147483647, the shape in which the reads
    are actually overlapped.  The named workloads all have more than 32767 reads, i.e. the
    wide pair key and canonical record order; scaled-down runs of them keep that form (the reference-order
    replay of a table with millions of buckets is host work that the full-size workloads never do)."""
    kw = dict(k=w["k"], id_mode=1, order=1)
    if shape == "wide":
        kw.update(max_ignore=w["length"], max_coll=2147483647)
    return kw
