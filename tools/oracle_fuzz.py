#!/usr/bin/env python
"""Differential fuzz of the CPU oracle (oracle/links_oracle.c) against the UNMODIFIED reference binary.

    python tools/oracle_fuzz.py [--cases 100] [--seed 1] [--keep]      (build container only: needs oracle/_ref/LINKS_CPP)

The oracle is what the GPU parity tests check the CUDA path against; the goldens pin it on a fixed set of inputs.  This
widens the pin: random small jobs -- genome 15-120 kb with repeats, contigs with N runs / lower case / IUPAC codes / records
shorter than k, reads of 3-12x coverage with 0-10 % errors, k 8..40, one to four distances with their own step sizes, an
offset now and then, two insert-size tolerances -- run through `LINKS_CPP -j 1` and through the oracle, and the three output
files must agree: .bloom byte for byte, .tigpair_checkpoint.tsv after a line sort, .pairing_issues after a record sort (the
same normalisation the reference's CI applies, azure-pipelines.yml:43-47).  Exit code 1 if any case differs.
"""
from __future__ import annotations

import argparse
import os
import shutil
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def dirty(rng: np.random.Generator, seqset):
    """Lower case stretches, IUPAC codes, N runs, a 'U': what the reference's readers pass through and ntHash skips or folds."""
    seq = seqset.seq.copy()
    n = len(seq)
    if n < 100:
        return seqset
    for _ in range(int(rng.integers(0, 4))):
        a = int(rng.integers(0, n - 50))
        b = a + int(rng.integers(1, 400))
        seq[a:b] = np.frombuffer(seq[a:b].tobytes().lower(), np.uint8)
    for _ in range(int(rng.integers(0, 4))):
        a = int(rng.integers(0, n - 50))
        seq[a:a + int(rng.integers(1, 40))] = ord("N")
    for _ in range(int(rng.integers(0, 6))):
        seq[int(rng.integers(0, n))] = int(rng.choice(np.frombuffer(b"RYKMSWnUu-.", np.uint8)))
    from links_b200.seqio import SeqSet
    return SeqSet(seq, seqset.off, seqset.names)


def make_case(seed: int, wide: bool = False):
    from links_b200 import synth
    rng = np.random.default_rng(seed)
    G = int(rng.integers(15_000, 120_000))
    genome = synth.make_genome(G, seed=seed * 7 + 1, repeat_frac=float(rng.choice([0.0, 0.05, 0.2])))
    contigs = synth.make_contigs(genome, int(rng.integers(2, 25)), seed=seed * 7 + 2, short_frac=float(rng.choice([0.0, 0.05, 0.2])),
                                 n_run_frac=float(rng.choice([0.0, 0.1, 0.5])))
    reads = synth.make_reads(genome, float(rng.uniform(3, 12)), int(rng.integers(400, 6000)), seed=seed * 7 + 3,
                             err=float(rng.choice([0.0, 0.01, 0.05, 0.1])), n_read_frac=float(rng.choice([0.0, 0.01, 0.2])))
    if rng.random() < 0.5:
        contigs = dirty(rng, contigs)
    if rng.random() < 0.5:
        reads = dirty(rng, reads)
    contigs.names = [str(i + 1) for i in range(contigs.n)]
    reads.names = [f"r{i}" for i in range(reads.n)]
    k = int(rng.choice([8, 11, 15, 16, 17, 20, 26, 31, 32, 33, 40]))
    if wide:  # (a separate switch: the seeded cases of the test suites keep the jobs they were validated with)
        k = int(rng.choice([1, 2, 3, 5, 48, 63, 64, 65, 100, 127, 200]))
    nd = int(rng.integers(1, 5))
    d = sorted(int(x) for x in rng.choice(np.arange(100, 6000, 50), size=nd, replace=False))
    if rng.random() < 0.3:
        rng.shuffle(d)  # first distance wins: the order given is the order tried
        d = [int(x) for x in d]
    t = [int(rng.choice([1, 2, 3, 5, 10, 20]))] * nd if rng.random() < 0.5 else [int(rng.choice([1, 2, 3, 5, 10, 20])) for _ in range(nd)]
    opts = {"k": k, "d": d, "t": t, "o": int(rng.choice([0, 0, 0, 1, 7, 50])), "e": float(rng.choice([0.1, 0.3])), "p": 0.001}
    return contigs, reads, opts


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=100)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--keep", action="store_true")
    ap.add_argument("--wide", action="store_true", help="k from {1, 2, 3, 5, 48, 63, 64, 65, 100, 127, 200} instead of 8..40")
    a = ap.parse_args()
    from make_golden import REF_BIN, fasta_bytes, run_ref, sort_issue_records
    from oracle import oracle_py as oracle
    if not os.path.exists(REF_BIN):
        print(f"{REF_BIN} is missing: make -C oracle ref")
        return 2
    oracle.load()
    n_bad = 0
    n_rows = n_issues = 0
    for i in range(a.cases):
        contigs, reads, o = make_case(a.seed * 100_003 + i, a.wide)
        wd = tempfile.mkdtemp(prefix="oracle_fuzz_")
        cfa = fasta_bytes(contigs)
        open(os.path.join(wd, "contigs.fa"), "wb").write(cfa)
        open(os.path.join(wd, "reads.fa"), "wb").write(fasta_bytes(reads))
        try:
            ref = run_ref(wd, os.path.join(wd, "contigs.fa"), [os.path.join(wd, "reads.fa")], o)
        except Exception as e:  # the reference itself refused the job (it prints why): not a parity case
            print(f"case {i}: reference failed ({str(e)[:120]}), opts {o}")
            shutil.rmtree(wd, ignore_errors=True)
            continue
        _m, nbytes, h = oracle.bf_size(len(cfa), o["p"])
        bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, o["k"])
        bloom = oracle.bf_header(o["k"], h, nbytes) + bits.tobytes()
        rows, issues, _st, _ = oracle.run(o["k"], o["d"], o["t"], o["o"], o["e"], bits, h, contigs.seq, contigs.off, reads.seq, reads.off)
        tsv = b"".join(sorted(oracle.format_rows(rows).splitlines(keepends=True)))
        iss = sort_issue_records(oracle.format_issues(issues))
        diff = [n for n, x, y in (("bloom", bloom, ref["bloom"]), ("tsv", tsv, ref["tsv"]), ("issues", iss, ref["issues"])) if x != y]
        n_rows += len(rows)
        n_issues += len(issues)
        if diff or h != ref["hash_num"]:
            n_bad += 1
            print(f"case {i} (seed {a.seed}) differs in {diff}: opts {o}, {contigs.n} contigs, {reads.n} reads" + (f"  kept {wd}" if a.keep else ""), flush=True)
            if a.keep:
                continue
        shutil.rmtree(wd, ignore_errors=True)
    print(f"{a.cases} cases: {n_bad} differ ({n_rows} link rows and {n_issues} pairing issues compared)")
    return 1 if n_bad else 0


if __name__ == "__main__":
    sys.exit(main())
