"""Host-side (numpy) synthetic Illumina-shaped reads for tests and small parity cases.

Same statistical models as the device generators in csrc/fqk_synth.cu
 but a different random stream; the bench generates its
workload on the device and samples it back, it does not use this module.
"""
import numpy as np

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, dtype=np.uint8)
for a, b in zip(b"ACGTN", b"TGCAN"):
    _COMP[a] = b
TRUSEQ_R1 = b"AGATCGGAAGAGCACACGTCTGAACTCCAGTCA"      # /root/reference/src/knownadapters.h:14
TRUSEQ_R2 = b"AGATCGGAAGAGCGTCGTGTAGGGAAAGAGTGT"      # /root/reference/src/knownadapters.h:15


def _fixed_off(n, L):
    return (np.arange(n + 1, dtype=np.uint64) * np.uint64(L))


def se_hiseq(n, L=150, seed=1, n_rate=0.002, collapse_p=0.08):
    """cfg2 model: mean Q 38->30 along the read, sd 4, clamp [2,41]; 8 % of reads
    collapse to Q2 over their last <=15 cycles; N with p=0.002 (Q2)."""
    rng = np.random.default_rng(seed)
    seq = _ACGT[rng.integers(0, 4, size=(n, L))]
    mean = 38.0 - 8.0 * np.arange(L) / max(L - 1, 1)
    q = np.clip(np.rint(rng.normal(mean, 4.0, size=(n, L))), 2, 41).astype(np.int32)
    coll = rng.random(n) < collapse_p
    start = L - np.minimum(rng.geometric(0.2, size=n), 15)
    q = np.where(coll[:, None] & (np.arange(L)[None, :] >= start[:, None]), 2, q)
    isn = rng.random((n, L)) < n_rate
    seq = np.where(isn, np.uint8(ord("N")), seq)
    q = np.where(isn, 2, q)
    return seq.reshape(-1).astype(np.uint8), (q + 33).reshape(-1).astype(np.uint8), _fixed_off(n, L)


def _novaseq_q(rng, n, L):
    lv = np.array([2, 12, 23, 37])
    p3 = np.linspace(0.0, 0.12, L)                      # degrade toward 3'
    pr = np.stack([0.01 + p3 * 0.2, 0.03 + p3 * 0.3, 0.08 + p3 * 0.5, 0.88 - p3], axis=1)
    cdf = np.cumsum(pr, axis=1)
    u = rng.random((n, L))
    k = (u[:, :, None] > cdf[None, :, :]).sum(axis=2).clip(0, 3)
    return lv[k]


def pe_novaseq(n, L=150, seed=2, frag_mean=300.0, frag_sd=90.0, err=0.003):
    """cfg3/cfg5 model: fragment ~ N(mean, sd) clamped [35, 700]; R2 = revcomp of the
    fragment end; short fragments read through into TruSeq adapters then polyG."""
    rng = np.random.default_rng(seed)
    frag = np.clip(np.rint(rng.normal(frag_mean, frag_sd, size=n)), 35, 700).astype(np.int64)
    s1 = np.empty((n, L), dtype=np.uint8); s2 = np.empty((n, L), dtype=np.uint8)
    a1 = np.frombuffer(TRUSEQ_R1, dtype=np.uint8); a2 = np.frombuffer(TRUSEQ_R2, dtype=np.uint8)
    for i in range(n):
        f = _ACGT[rng.integers(0, 4, size=frag[i])]
        rc = _COMP[f[::-1]]
        for dst, src, ad in ((s1, f, a1), (s2, rc, a2)):
            r = np.concatenate([src[:L], ad, np.full(L, ord("G"), dtype=np.uint8)])[:L] if len(src) < L else src[:L]
            dst[i] = r
    for s in (s1, s2):
        e = rng.random((n, L)) < err
        s[e] = _ACGT[rng.integers(0, 4, size=int(e.sum()))]
    q1 = _novaseq_q(rng, n, L) + 33; q2 = _novaseq_q(rng, n, L) + 33
    off = _fixed_off(n, L)
    return (s1.reshape(-1), q1.reshape(-1).astype(np.uint8), off), (s2.reshape(-1), q2.reshape(-1).astype(np.uint8), off.copy())


def se_adapter_polyx(n, L=100, seed=3, adapter=TRUSEQ_R1, adapter_p=0.12, polyx_p=0.03, dup_p=0.10):
    """cfg4 model: 12 % adapter at a uniform position >= 20 followed by polyA, 3 % polyX
    tails (10-40 bases, 1 mismatch / 10), 10 % of reads copy an earlier read."""
    rng = np.random.default_rng(seed)
    seq = _ACGT[rng.integers(0, 4, size=(n, L))].copy()
    ad = np.frombuffer(adapter, dtype=np.uint8)
    for i in np.nonzero(rng.random(n) < adapter_p)[0]:
        p = int(rng.integers(20, L))
        tail = np.concatenate([ad, np.full(L, ord("A"), dtype=np.uint8)])[:L - p]
        seq[i, p:] = tail
    for i in np.nonzero(rng.random(n) < polyx_p)[0]:
        k = int(rng.integers(10, 41)); x = _ACGT[rng.integers(0, 4)]
        t = np.full(k, x, dtype=np.uint8)
        mm = rng.random(k) < 0.1
        t[mm] = _ACGT[rng.integers(0, 4, size=int(mm.sum()))]
        seq[i, L - k:] = t
    for i in np.nonzero(rng.random(n) < dup_p)[0]:
        if i > 0:
            seq[i] = seq[int(rng.integers(0, i))]
    mean = 37.0 - 6.0 * np.arange(L) / max(L - 1, 1)
    q = np.clip(np.rint(rng.normal(mean, 4.0, size=(n, L))), 2, 41).astype(np.int32)
    isn = rng.random((n, L)) < 0.002
    seq = np.where(isn, np.uint8(ord("N")), seq)
    q = np.where(isn, 2, q)
    return seq.reshape(-1).astype(np.uint8), (q + 33).reshape(-1).astype(np.uint8), _fixed_off(n, L)


def ragged(n, seed=4, lmin=0, lmax=220, n_rate=0.03, lowq_rate=0.25):
    """Adversarial variable-length reads: empty reads, many N, polyG/polyX tails, low-quality runs."""
    rng = np.random.default_rng(seed)
    lens = rng.integers(lmin, lmax + 1, size=n)
    tiny = rng.random(n) < 0.05
    lens[tiny] = rng.integers(0, 8, size=int(tiny.sum()))
    off = np.zeros(n + 1, dtype=np.uint64)
    np.cumsum(lens, out=off[1:].view(np.int64))
    tot = int(off[-1])
    seq = _ACGT[rng.integers(0, 4, size=tot)].copy()
    q = rng.integers(2, 42, size=tot)
    low = rng.random(tot) < lowq_rate
    q = np.where(low, rng.integers(2, 20, size=tot), q)
    for i in range(n):
        a, b = int(off[i]), int(off[i + 1])
        L = b - a
        if L == 0:
            continue
        u = rng.random()
        if u < 0.2:                                     # homopolymer tail with sparse mismatches
            k = int(rng.integers(1, L + 1)); x = b"ACGTN"[int(rng.integers(0, 5))]
            t = np.full(k, x, dtype=np.uint8); mm = rng.random(k) < 0.08
            t[mm] = _ACGT[rng.integers(0, 4, size=int(mm.sum()))]
            seq[b - k:b] = t
        elif u < 0.3:                                   # low-quality head and tail
            k = int(rng.integers(1, L + 1))
            q[a:a + k // 2] = rng.integers(2, 12, size=k // 2)
            q[b - k // 2:b] = rng.integers(2, 12, size=k // 2)
    isn = rng.random(tot) < n_rate
    seq = np.where(isn, np.uint8(ord("N")), seq).astype(np.uint8)
    return seq, (q + 33).astype(np.uint8), off


def revcomp_reads(reads):
    """Reverse-complement every read (helper for building overlapping pairs in tests)."""
    seq, qual, off = reads
    out_s = np.empty_like(seq); out_q = np.empty_like(qual)
    for i in range(len(off) - 1):
        a, b = int(off[i]), int(off[i + 1])
        out_s[a:b] = _COMP[seq[a:b][::-1]]
        out_q[a:b] = qual[a:b][::-1]
    return out_s, out_q, off.copy()
