#!/usr/bin/env python
"""Generate tests/golden/ from the reference itself.  Run in the build container only (needs
/root/reference and oracle/_ref/LINKS_CPP built by `make -C oracle ref`):

    python tools/make_golden.py

 c1/        the reference's installation test (reference tests/run_me.sh): its two input FASTA
            files (gzip'd copies of test DATA, not source) and the golden outputs shipped in
            reference tests/expected_outputs/ (.tigpair_checkpoint.tsv, .pairing_issues, sha256 of .bloom)
 adv_*/     adversarial synthetic inputs (Ns, lowercase, short reads/contigs, tandem repeats with
            several -d, even k, k > 32, FASTQ reads, non-zero -o) with the outputs of the UNMODIFIED
            reference binary run with -j 1 (deterministic first-wins order).
Each case directory holds case.json (options + sha256/size of the .bloom + BF sizing the binary printed).
"""
from __future__ import annotations

import gzip
import hashlib
import json
import os
import re
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from links_b200 import synth  # noqa: E402
from links_b200.seqio import SeqSet, write_fasta  # noqa: E402

REF = "/root/reference"
REF_BIN = os.path.join(ROOT, "oracle", "_ref", "LINKS_CPP")
GOLD = os.path.join(ROOT, "tests", "golden")


def gz_write(path: str, data: bytes) -> None:
    with open(path, "wb") as raw:
        with gzip.GzipFile(filename="", mode="wb", fileobj=raw, mtime=0, compresslevel=9) as f:
            f.write(data)


def sort_issue_records(text: bytes) -> bytes:
    lines = text.split(b"\n")
    recs = [b"\n".join(lines[i:i + 4]) for i in range(0, len(lines) - 1, 4)]
    return b"".join(r + b"\n" for r in sorted(recs))


def run_ref(workdir: str, contigs_fa: str, read_files: list[str], opts: dict) -> dict:
    fof = os.path.join(workdir, "reads.fof")
    with open(fof, "w") as f:
        f.write("\n".join(os.path.basename(p) for p in read_files) + "\n")
    cmd = [REF_BIN, "-f", os.path.basename(contigs_fa), "-s", "reads.fof", "-b", "out", "-j", "1",
           "-d", ",".join(map(str, opts["d"])), "-t", ",".join(map(str, opts["t"])), "-k", str(opts["k"]),
           "-o", str(opts.get("o", 0)), "-e", str(opts.get("e", 0.1)), "-p", str(opts.get("p", 0.001))]
    res = subprocess.run(cmd, cwd=workdir, capture_output=True, text=True, check=True)
    m = re.search(r"- m: (\d+)", res.stdout)
    h = re.search(r"- hashFct: (\d+)", res.stdout)
    bloom = open(os.path.join(workdir, "out.bloom"), "rb").read()
    tsv = open(os.path.join(workdir, "out.tigpair_checkpoint.tsv"), "rb").read()
    issues = open(os.path.join(workdir, "out.pairing_issues"), "rb").read()
    return {"m": int(m.group(1)), "hash_num": int(h.group(1)), "bloom": bloom,
            "tsv": b"".join(sorted(tsv.splitlines(keepends=True))), "issues": sort_issue_records(issues)}


def save_case(name: str, contigs: bytes, reads: dict[str, bytes], opts: dict, out: dict) -> None:
    d = os.path.join(GOLD, name)
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    gz_write(os.path.join(d, "contigs.fa.gz"), contigs)
    for fn, data in reads.items():
        gz_write(os.path.join(d, fn + ".gz"), data)
    with open(os.path.join(d, "expected.tigpair_checkpoint.tsv"), "wb") as f:
        f.write(out["tsv"])
    with open(os.path.join(d, "expected.pairing_issues"), "wb") as f:
        f.write(out["issues"])
    hdr_end = out["bloom"].index(b"[HeaderEnd]\n") + len(b"[HeaderEnd]\n")
    meta = {"opts": opts, "read_files": list(reads.keys()), "contigs_file_bytes": len(contigs),
            "bf_m_bits": out["m"], "bf_hash_num": out["hash_num"], "bloom_file_bytes": len(out["bloom"]),
            "bloom_sha256": hashlib.sha256(out["bloom"]).hexdigest(),
            "bloom_header": out["bloom"][:hdr_end + 50 + len("  <binary data>")].decode(),
            "n_rows": out["tsv"].count(b"\n"), "how": "oracle/_ref/LINKS_CPP -j 1 (unmodified reference, tools/make_golden.py)"}
    with open(os.path.join(d, "case.json"), "w") as f:
        json.dump(meta, f, indent=1, sort_keys=True)
    print(f"{name}: rows={meta['n_rows']} issues={out['issues'].count(b'Pairs unsat')} bloom={meta['bloom_file_bytes']}")


def fasta_bytes(s: SeqSet) -> bytes:
    return b"".join(b">" + s.names[i].encode() + b"\n" + s.record(i) + b"\n" for i in range(s.n))


def fastq_bytes(s: SeqSet) -> bytes:
    return b"".join(b"@" + s.names[i].encode() + b"\n" + s.record(i) + b"\n+\n" + b"I" * len(s.record(i)) + b"\n"
                    for i in range(s.n))


def case_from_sets(name: str, contigs: SeqSet, read_sets: dict[str, tuple[SeqSet, str]], opts: dict) -> None:
    with tempfile.TemporaryDirectory() as wd:
        cb = fasta_bytes(contigs)
        open(os.path.join(wd, "contigs.fa"), "wb").write(cb)
        reads_b = {}
        for fn, (s, fmt) in read_sets.items():
            reads_b[fn] = fastq_bytes(s) if fmt == "fastq" else fasta_bytes(s)
            open(os.path.join(wd, fn), "wb").write(reads_b[fn])
        out = run_ref(wd, os.path.join(wd, "contigs.fa"), [os.path.join(wd, fn) for fn in read_sets], opts)
        save_case(name, cb, reads_b, opts, out)


def sprinkle(s: SeqSet, rng, n_frac: float, lower_frac: float, n_runs: int) -> SeqSet:
    seq = s.seq.copy()
    n = len(seq)
    idx = rng.random(n) < n_frac
    seq[idx] = ord("N")
    for _ in range(n_runs):
        a = int(rng.integers(0, max(n - 50, 1)))
        seq[a:a + int(rng.integers(1, 40))] = rng.choice(np.frombuffer(b"NnRYKM-.", np.uint8))
    low = rng.random(n) < lower_frac
    seq[low] = seq[low] | 0x20
    return SeqSet(seq, s.off.copy(), list(s.names))


def main() -> None:
    if not os.path.exists(REF_BIN):
        raise SystemExit("build the reference first: make -C oracle ref")
    os.makedirs(GOLD, exist_ok=True)

    # ---- c1: the reference's installation test, goldens as shipped
    d = os.path.join(GOLD, "c1")
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    contigs = open(f"{REF}/tests/simulated_contigs.fa", "rb").read()
    genome = open(f"{REF}/tests/simulated_genome.fa", "rb").read()
    gz_write(os.path.join(d, "contigs.fa.gz"), contigs)
    gz_write(os.path.join(d, "simulated_genome.fa.gz"), genome)
    exp = f"{REF}/tests/expected_outputs/installation_test"
    tsv = open(exp + ".tigpair_checkpoint.tsv", "rb").read()
    open(os.path.join(d, "expected.tigpair_checkpoint.tsv"), "wb").write(b"".join(sorted(tsv.splitlines(keepends=True))))
    open(os.path.join(d, "expected.pairing_issues"), "wb").write(sort_issue_records(open(exp + ".pairing_issues", "rb").read()))
    bloom = open(exp + ".bloom", "rb").read()
    hdr_end = bloom.index(b"[HeaderEnd]\n") + len(b"[HeaderEnd]\n")
    opts = {"d": [1000, 2000, 3000, 4000, 5000, 6000, 7000, 8000, 9000, 10000, 12000, 14000, 16000, 18000, 20000],
            "t": [2], "k": 15, "o": 0, "e": 0.1, "p": 0.001}
    meta = {"opts": opts, "read_files": ["simulated_genome.fa"], "contigs_file_bytes": len(contigs),
            "bf_m_bits": 10373792, "bf_hash_num": 9, "bloom_file_bytes": len(bloom),
            "bloom_sha256": hashlib.sha256(bloom).hexdigest(),
            "bloom_header": bloom[:hdr_end + 50 + len("  <binary data>")].decode(), "n_rows": tsv.count(b"\n"),
            "how": "copied from reference tests/ and tests/expected_outputs/ (golden files shipped by the reference)"}
    json.dump(meta, open(os.path.join(d, "case.json"), "w"), indent=1, sort_keys=True)
    print(f"c1: rows={meta['n_rows']}")

    rng = np.random.default_rng(11)

    # ---- adv_ns_lower: Ns, IUPAC, lowercase, short reads and contigs, three distances
    g = synth.make_genome(60_000, seed=5)
    contigs_s = sprinkle(synth.make_contigs(g, 12, seed=6, short_frac=0.2, n_run_frac=0.3), rng, 0.0005, 0.05, 3)
    contigs_s = SeqSet.concat([contigs_s, SeqSet.from_records([b"ACGT", b"", b"ACGTACGTACGTAC", b"N" * 40])])
    contigs_s.names = [str(i + 1) for i in range(contigs_s.n)]
    reads_s = sprinkle(synth.make_reads(g, 12, 3000, seed=7, err=0.01, n_read_frac=0.3), rng, 0.002, 0.1, 40)
    reads_s = SeqSet.concat([reads_s, SeqSet.from_records([b"ACGTAC", b"", b"N" * 700, b"ACGTN" * 300])])
    reads_s.names = [f"r{i}" for i in range(reads_s.n)]
    case_from_sets("adv_ns_lower", contigs_s, {"reads.fa": (reads_s, "fasta")},
                   {"d": [500, 1000, 1500], "t": [2, 3, 1], "k": 15, "o": 0, "e": 0.1, "p": 0.001})

    # ---- adv_offset_k20: even k (palindromes possible), -o 7, -t 5, -e 0.25, FASTQ reads in two files
    g = synth.make_genome(80_000, seed=8)
    contigs_s = synth.make_contigs(g, 10, seed=9)
    r1 = synth.make_reads(g, 8, 4000, seed=10, err=0.02)
    r2 = sprinkle(synth.make_reads(g, 4, 2500, seed=12, err=0.0), rng, 0.001, 0.0, 5)
    case_from_sets("adv_offset_k20", contigs_s, {"reads_a.fq": (r1, "fastq"), "reads_b.fa": (r2, "fasta")},
                   {"d": [800, 2000], "t": [5], "k": 20, "o": 7, "e": 0.25, "p": 0.01})

    # ---- adv_tandem: (1) 250-bp units duplicated in the READ genome but collapsed in the contigs, so the
    #      same (h_a,h_b) is observed under several -d and the first-wins rule decides its distance;
    #      (2) tandem arrays + duplicated contigs (multiplicity > 1); exact reads, heavy dedup
    pieces = [synth.make_genome(5000, seed=100 + i, repeat_frac=0) for i in range(7)]
    units = [synth.make_genome(250, seed=200 + i, repeat_frac=0) for i in range(6)]
    g_reads = np.concatenate([np.concatenate([pieces[i], units[i], units[i]]) for i in range(6)] + [pieces[6]])
    recs = [pieces[0].tobytes()] + [np.concatenate([units[i], pieces[i + 1]]).tobytes() for i in range(6)]
    unit = synth.make_genome(250, seed=13, repeat_frac=0)
    flank = synth.make_genome(30_000, seed=14, repeat_frac=0)
    g2 = np.concatenate([flank[:10_000], np.tile(unit, 12), flank[10_000:20_000], np.tile(unit[:200], 10), flank[20_000:]])
    c2 = synth.make_contigs(g2, 8, seed=15, short_frac=0.0, n_run_frac=0.0)
    contigs_s = SeqSet.concat([SeqSet.from_records(recs), c2, SeqSet.from_records([c2.record(1), c2.record(2)[:300]])])
    contigs_s.names = [str(i + 1) for i in range(contigs_s.n)]
    reads_a = synth.make_reads(g_reads, 30, 4000, seed=16, err=0.0, n_read_frac=0.0)
    reads_b = synth.make_reads(g2, 25, 5000, seed=21, err=0.0, n_read_frac=0.0)
    case_from_sets("adv_tandem", contigs_s, {"reads_a.fa": (reads_a, "fasta"), "reads_b.fa": (reads_b, "fasta")},
                   {"d": [750, 500, 1000, 250], "t": [1, 2], "k": 15, "o": 0, "e": 0.1, "p": 0.001})

    # ---- adv_k40: k > 32 (window spans > 64 packed bits), odd distance, -t 3
    g = synth.make_genome(50_000, seed=17)
    contigs_s = sprinkle(synth.make_contigs(g, 6, seed=18), rng, 0.0002, 0.02, 1)
    reads_s = sprinkle(synth.make_reads(g, 10, 6000, seed=19, err=0.005), rng, 0.0005, 0.02, 4)
    case_from_sets("adv_k40", contigs_s, {"reads.fa": (reads_s, "fasta")},
                   {"d": [1501], "t": [3], "k": 40, "o": 3, "e": 0.1, "p": 0.001})


if __name__ == "__main__":
    main()
