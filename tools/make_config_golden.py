#!/usr/bin/env python
"""Goldens for the BASELINE.json configurations at a size the unmodified reference finishes in seconds.

    python tools/make_config_golden.py        (build container only: needs oracle/_ref/LINKS_CPP)

configs[2] (several distances with their own step sizes), configs[3] (k 20, d 5000, t 5) and configs[4]
(the step-size sweep at d 10000, t = 1 ... 100) have no fixtures in the reference.  Their inputs are
synthetic and SEEDED (links_b200/synth.py regenerates the same bytes anywhere; case.json keeps a sha256
of them), so only the reference's OUTPUT is stored: tests/golden/cfg_<name>/{case.json,
expected.tigpair_checkpoint.tsv, expected.pairing_issues}, produced by `oracle/_ref/LINKS_CPP -j 1`.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
from links_b200 import synth  # noqa: E402
from make_golden import GOLD, REF_BIN, fasta_bytes, run_ref  # noqa: E402

# name: (genome bp, contigs, coverage, mean read length, k, distances, steps)
CASES = {
    "c3_multi_distance": (300_000, 12, 20.0, 15_000, 15, [1000, 2500, 5000], [20, 10, 5]),
    "c4_k20_d5000_t5": (300_000, 10, 10.0, 20_000, 20, [5000], [5]),
    "c5_sweep_t1": (150_000, 8, 6.0, 20_000, 15, [10000], [1]),
    "c5_sweep_t3": (150_000, 8, 6.0, 20_000, 15, [10000], [3]),
    "c5_sweep_t7": (150_000, 8, 6.0, 20_000, 15, [10000], [7]),
    "c5_sweep_t20": (150_000, 8, 6.0, 20_000, 15, [10000], [20]),
    "c5_sweep_t100": (150_000, 8, 6.0, 20_000, 15, [10000], [100]),
}


def make_inputs(name: str):
    g, nc, cov, rl, k, d, t = CASES[name]
    genome = synth.make_genome(g, seed=11)
    contigs = synth.make_contigs(genome, nc, seed=13)
    reads = synth.make_reads(genome, cov, rl, seed=12)
    contigs.names = [str(i + 1) for i in range(contigs.n)]
    reads.names = [f"r{i}" for i in range(reads.n)]
    return contigs, reads, k, d, t


def inputs_digest(contigs, reads) -> str:
    h = hashlib.sha256()
    for a in (contigs.seq, contigs.off, reads.seq, reads.off):
        h.update(a.tobytes())
    return h.hexdigest()


def main() -> None:
    if not os.path.exists(REF_BIN):
        raise SystemExit("build the reference first: make -C oracle ref")
    for name in CASES:
        contigs, reads, k, d, t = make_inputs(name)
        opts = {"k": k, "d": d, "t": t, "o": 0, "e": 0.1, "p": 0.001}
        with tempfile.TemporaryDirectory() as wd:
            cb = fasta_bytes(contigs)
            open(os.path.join(wd, "contigs.fa"), "wb").write(cb)
            open(os.path.join(wd, "reads.fa"), "wb").write(fasta_bytes(reads))
            out = run_ref(wd, os.path.join(wd, "contigs.fa"), [os.path.join(wd, "reads.fa")], opts)
        dst = os.path.join(GOLD, "cfg_" + name)
        shutil.rmtree(dst, ignore_errors=True)
        os.makedirs(dst)
        open(os.path.join(dst, "expected.tigpair_checkpoint.tsv"), "wb").write(out["tsv"])
        open(os.path.join(dst, "expected.pairing_issues"), "wb").write(out["issues"])
        meta = {"opts": opts, "synth": list(CASES[name]), "inputs_sha256": inputs_digest(contigs, reads),
                "contigs_file_bytes": len(cb), "bf_m_bits": out["m"], "bf_hash_num": out["hash_num"],
                "bloom_sha256": hashlib.sha256(out["bloom"]).hexdigest(), "n_rows": out["tsv"].count(b"\n"),
                "read_bases": int(reads.off[-1]), "n_reads": reads.n,
                "how": "oracle/_ref/LINKS_CPP -j 1 (unmodified reference) on links_b200.synth inputs, tools/make_config_golden.py"}
        json.dump(meta, open(os.path.join(dst, "case.json"), "w"), indent=1, sort_keys=True)
        print(f"cfg_{name}: reads={reads.n} bases={meta['read_bases']} rows={meta['n_rows']} "
              f"issues={out['issues'].count(b'Pairs unsat')}")


if __name__ == "__main__":
    main()
