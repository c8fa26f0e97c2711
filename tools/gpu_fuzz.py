#!/usr/bin/env python
"""The CUDA path against the CPU oracle on random small jobs (tools/oracle_fuzz.py's generator), as many as fit a time budget.

    python tools/gpu_fuzz.py [--first 2000] [--cases 200] [--seconds 30]        (needs a GPU; TEST INFRASTRUCTURE)

tests/test_gpu_parity.py::test_fuzz_jobs_match_oracle runs a seeded two dozen of these in the suite; this is the wider
one-off sweep.  Alternates the contig table between AUTO and MATES.  One JSON line: jobs run, jobs that differed (with
their seeds), rows and issues compared.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--first", type=int, default=2000)
    ap.add_argument("--cases", type=int, default=200)
    ap.add_argument("--seconds", type=float, default=30.0)
    a = ap.parse_args()
    from oracle_fuzz import make_case
    from links_b200 import Links, LinksOptions, bf_size
    from oracle import oracle_py as oracle
    oracle.load()
    srt = lambda b: b"".join(sorted(b.splitlines(keepends=True)))  # noqa: E731

    def recs(b: bytes) -> bytes:
        lines = b.split(b"\n")
        return b"".join(sorted(b"\n".join(lines[i:i + 4]) + b"\n" for i in range(0, len(lines) - 1, 4)))

    t0 = time.time()
    bad, n, n_rows, n_issues = [], 0, 0, 0
    for seed in range(a.first, a.first + a.cases):
        if time.time() - t0 > a.seconds:
            break
        contigs, reads, o = make_case(seed)
        fb = sum(len(contigs.names[i]) + 2 + int(contigs.off[i + 1] - contigs.off[i]) + 1 for i in range(contigs.n))
        _m, nbytes, h = bf_size(fb, o["p"])
        with Links(LinksOptions(k=o["k"], distances=o["d"], step_sizes=o["t"], offset=o["o"], insert_stdev=o["e"])) as L:
            L.table_mode("mates" if seed & 1 else "auto")
            L.bf_alloc(nbytes, h)
            L.push_contigs(contigs)
            L.push_reads(reads)
            bits = L.bf_bits()
            rows, issues, stats = L.finalize()
            tsv, iss = L.format_rows(rows), L.format_issues(issues)
        ok = np.array_equal(bits, oracle.bf_build(contigs.seq, contigs.off, nbytes, h, o["k"]))
        orows, oiss, ost, _ = oracle.run(o["k"], o["d"], o["t"], o["o"], o["e"], bits, h, contigs.seq, contigs.off, reads.seq, reads.off)
        ok = ok and srt(tsv) == srt(oracle.format_rows(orows)) and recs(iss) == recs(oracle.format_issues(oiss))
        ok = ok and all(stats[k] == ost[k] for k in ("n_candidates", "n_survivors", "n_unique_pairs"))
        if not ok:
            bad.append(seed)
        n += 1
        n_rows += len(rows)
        n_issues += len(issues)
    print(json.dumps({"probe": "gpu_vs_oracle_fuzz", "first_seed": a.first, "jobs": n, "differ": bad, "rows_compared": n_rows,
                      "issues_compared": n_issues, "seconds": round(time.time() - t0, 1)}), flush=True)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
