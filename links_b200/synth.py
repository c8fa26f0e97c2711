"""Seeded synthetic inputs of the shapes BASELINE.json names.

genome : i.i.d. uniform ACGT plus ~5 % of its length in 2-5 kb repeats copied to random loci
contigs: the genome cut at exponential lengths around a target mean, a 0-2 kb gap dropped between
         neighbours, 50 % reverse-complemented, ~1 % shorter than 500 bp, a few with an N run
reads  : uniform starts, log-normal lengths around the target mean, random strand, ONT-like 5 %
         errors (sub:ins:del = 1:1:1), 0.1 % of reads carrying an N
Everything is numpy; the same seed gives the same bytes on every machine.
"""
from __future__ import annotations

import numpy as np

from .seqio import SeqSet

_ACGT = np.frombuffer(b"ACGT", dtype=np.uint8)
_COMP = np.zeros(256, dtype=np.uint8)
for _a, _b in zip(b"ACGTNacgtn", b"TGCANtgcan"):
    _COMP[_a] = _b


def revcomp(a: np.ndarray) -> np.ndarray:
    return _COMP[a[::-1]]


def make_genome(n: int, seed: int = 1, repeat_frac: float = 0.05) -> np.ndarray:
    rng = np.random.default_rng(seed)
    g = _ACGT[rng.integers(0, 4, size=n, dtype=np.uint8)]
    budget = int(n * repeat_frac)
    while budget > 0 and n > 12000:
        ln = int(rng.integers(2000, 5001))
        src = int(rng.integers(0, n - ln))
        dst = int(rng.integers(0, n - ln))
        g[dst:dst + ln] = g[src:src + ln].copy()
        budget -= ln
    return g


def make_contigs(genome: np.ndarray, n_contigs: int, seed: int = 3, short_frac: float = 0.01,
                 n_run_frac: float = 0.01) -> SeqSet:
    rng = np.random.default_rng(seed)
    n = len(genome)
    mean_len = max(n / max(n_contigs, 1) - 1000, 200)
    recs: list[bytes] = []
    pos = 0
    while pos < n:
        if rng.random() < short_frac:
            ln = int(rng.integers(20, 500))
        else:
            ln = int(max(200, rng.exponential(mean_len)))
        seg = genome[pos:pos + ln].copy()
        if len(seg) == 0:
            break
        if rng.random() < n_run_frac and len(seg) > 400:
            a = int(rng.integers(100, len(seg) - 100))
            seg[a:a + int(rng.integers(1, 60))] = ord("N")
        if rng.random() < 0.5:
            seg = revcomp(seg)
        recs.append(seg.tobytes())
        pos += ln + int(rng.integers(0, 2001))
    return SeqSet.from_records(recs)


def make_reads(genome: np.ndarray, coverage: float, mean_len: int, seed: int = 2, err: float = 0.05,
               n_read_frac: float = 0.001, sigma: float = 0.35) -> SeqSet:
    rng = np.random.default_rng(seed)
    n = len(genome)
    target = int(n * coverage)
    mu = np.log(mean_len) - sigma * sigma / 2
    chunks: list[np.ndarray] = []
    lens: list[int] = []
    total = 0
    while total < target:
        ln = int(min(max(rng.lognormal(mu, sigma), 200), n))
        st = int(rng.integers(0, n - ln + 1))
        seg = genome[st:st + ln]
        # errors: 0 = none, 1 = substitution, 2 = insertion after, 3 = deletion
        e = rng.random(ln) < err
        kind = np.zeros(ln, dtype=np.uint8)
        ne = int(e.sum())
        kind[e] = rng.integers(1, 4, size=ne, dtype=np.uint8)
        rep = np.ones(ln, dtype=np.int64)
        rep[kind == 2] = 2
        rep[kind == 3] = 0
        out = np.repeat(seg, rep)
        idx = np.cumsum(rep) - 1
        sub = kind == 1
        if sub.any():  # a different base: rotate within ACGT by 1..3
            cur = out[idx[sub]]
            code = (cur >> 1) & 3  # A0 C1 T2 G3
            lut = np.frombuffer(b"ACTG", dtype=np.uint8)
            out[idx[sub]] = lut[(code + rng.integers(1, 4, size=int(sub.sum()))) & 3]
        ins = kind == 2
        if ins.any():
            out[idx[ins]] = _ACGT[rng.integers(0, 4, size=int(ins.sum()))]
        if rng.random() < n_read_frac and len(out) > 10:
            out[int(rng.integers(0, len(out)))] = ord("N")
        if rng.random() < 0.5:
            out = revcomp(out)
        chunks.append(out)
        lens.append(len(out))
        total += len(out)
    off = np.zeros(len(lens) + 1, dtype=np.uint64)
    off[1:] = np.cumsum(lens, dtype=np.uint64)
    seq = np.concatenate(chunks) if chunks else np.zeros(0, np.uint8)
    return SeqSet(seq, off, [f"r{i}" for i in range(len(lens))])


def unique_lengths(contigs: SeqSet) -> SeqSet:
    """Trim bases off the end of later contigs until no two contigs have the same length.  The reference's layout stage
    seeds scaffolds longest contig first and breaks ties in Perl hash order (bin/LINKS.pl:673), which changes from run to
    run, so an input meant to pin its output must not contain ties."""
    seen: set[int] = set()
    recs = []
    for i in range(contigs.n):
        r = contigs.record(i)
        while len(r) in seen and len(r) > 1:
            r = r[:-1]
        seen.add(len(r))
        recs.append(r)
    return SeqSet.from_records(recs, list(contigs.names))


def candidate_pairs(read_off: np.ndarray, k: int, dist: list[int], step: list[int], offset: int = 0) -> int:
    """Analytic count of (A,B) k-mer pairs visited for N-free reads:
    sum_reads sum_d floor((len-k+1-o-d)/t)  (reference src/LINKS.hpp:525-543)."""
    lens = np.diff(read_off.astype(np.int64))
    total = 0
    for d, t in zip(dist, step):
        npos = lens - k + 1 - offset - d
        total += int((np.maximum(npos, 0) // t).sum())
    return total


CONFIGS = {
    # name: (genome bp, n contigs, coverage, mean read len, k, distances, steps)
    "c2_bacterial_5mbp": (5_000_000, 500, 50.0, 10_000, 15, [4000], [2]),
    "c3_celegans_100mbp": (100_000_000, 1500, 30.0, 15_000, 15, [1000, 2500, 5000], [20, 10, 5]),
    "c5_step_sweep_100mbp": (100_000_000, 1500, 5.0, 15_000, 15, [10000], [1]),
    "tiny": (200_000, 20, 10.0, 8_000, 15, [1000, 2000], [2, 3]),
}


def make_config(name: str, scale: float = 1.0):
    g, nc, cov, rl, k, d, t = CONFIGS[name]
    g = int(g * scale)
    nc = max(int(nc * scale), 2)
    genome = make_genome(g, seed=1)
    contigs = make_contigs(genome, nc, seed=3)
    reads = make_reads(genome, cov, rl, seed=2)
    return genome, contigs, reads, k, d, t
