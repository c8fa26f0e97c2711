#!/usr/bin/env python
"""Digest goldens for the BASELINE.json configurations at their FULL size, from the unmodified reference.

    python tools/make_fullsize_golden.py [name ...] [-j 1]     (build container only: needs oracle/_ref/LINKS_CPP)

The toy-size goldens (tools/make_config_golden.py) store the reference's output files; at full size those are too
large for the history, so only digests are kept: tests/golden/full_<name>/case.json holds

    tsv_sha256      sha256 of the byte-sorted lines of <b>.tigpair_checkpoint.tsv
    issues_sha256   sha256 of <b>.pairing_issues with its four-line records byte-sorted
    bloom_sha256    sha256 of <b>.bloom (header + bits)
    n_rows / n_issues / the order-independent counters of the PAIRED K-MER STATS block

plus what the run cost on this container (per-stage seconds from the reference's own stdout banners,
src/main.cpp:36-64; process wall; peak RSS): the full-size, same-config CPU baseline of BASELINE.md section 3.
Inputs are links_b200.synth's seeded workloads (bench.py --workload <name> --scale 1 uses the same bytes);
case.json keeps their sha256.  Multi-distance workloads need -j 1 (deterministic first-wins order).
"""
from __future__ import annotations

import argparse
import hashlib
import json
import os
import re
import resource
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from links_b200 import synth  # noqa: E402

REF_BIN = os.path.join(ROOT, "oracle", "_ref", "LINKS_CPP")
GOLD = os.path.join(ROOT, "tests", "golden")

# name -> (synth.CONFIGS key, scale, overrides of (coverage, steps)) ; the c5 sweep reuses one input for every -t
CASES = {
    "c2_bacterial_5mbp": ("c2_bacterial_5mbp", 1.0, {}),
    "c5_dense_t1": ("c5_step_sweep_100mbp", 1.0, {}),
    "c3_celegans_100mbp": ("c3_celegans_100mbp", 1.0, {}),
}


def write_fasta_stream(path: str, seq: np.ndarray, off: np.ndarray, prefix: str, one_based: bool) -> int:
    """One record per two lines; returns the file size.  Names: contigs '1'..'n' (the reference ignores them and
    numbers by ordinal, src/LINKS.hpp:657-659), reads 'r<i>'."""
    n = len(off) - 1
    total = 0
    with open(path, "wb", buffering=1 << 24) as f:
        for i in range(n):
            hdr = b">" + (prefix + str(i + 1 if one_based else i)).encode() + b"\n"
            f.write(hdr)
            f.write(seq[int(off[i]):int(off[i + 1])].tobytes())
            f.write(b"\n")
            total += len(hdr) + int(off[i + 1] - off[i]) + 1
    return total


def sort_issue_records(text: bytes) -> bytes:
    lines = text.split(b"\n")
    recs = [b"\n".join(lines[i:i + 4]) for i in range(0, len(lines) - 1, 4)]
    return b"".join(r + b"\n" for r in sorted(recs))


def parse_duration(s: str) -> float:
    m = re.match(r"(\d+)h (\d+)m (\d+)s (\d+)ms", s.strip())
    return int(m.group(1)) * 3600 + int(m.group(2)) * 60 + int(m.group(3)) + int(m.group(4)) / 1000.0


def parse_banners(stdout: str) -> dict:
    out = {}
    for key, pat in (("bloom_s", r"Bloom filter generated or loaded in: (.*)"), ("reads_s", r"Read fasta stage completed in: (.*)"),
                     ("contigs_s", r"Read contig stage completed in: (.*)"), ("pairing_s", r"Pair contig stage completed in: (.*)")):
        m = re.search(pat, stdout)
        if m:
            out[key] = parse_duration(m.group(1))
    return out


def parse_stats(stdout: str) -> dict:
    pats = {"ct_ok_contig": r"distance on target\): (\d+)", "ct_iz_issues": r"distance\s+out-of-bounds\): (\d+)",
            "ct_illogical": r"<-<- or <-->\): (\d+)", "ct_ok_pairs": r"\(pre-scaffold\): (\d+)",
            "ct_problem_pairs": r"calculated distances out-of-bounds\): (\d+)"}
    flat = stdout.replace("\n", " ")
    return {k: int(re.search(p, flat).group(1)) for k, p in pats.items() if re.search(p, flat)}


def inputs_digest(contigs, reads) -> str:
    h = hashlib.sha256()
    for a in (contigs.seq, contigs.off, reads.seq, reads.off):
        h.update(np.ascontiguousarray(a).tobytes() if a.nbytes < (1 << 31) else memoryview(np.ascontiguousarray(a)))
    return h.hexdigest()


def run_case(name: str, threads: int, keep: bool) -> None:
    cfg, scale, _ = CASES[name]
    t0 = time.time()
    _g, contigs, reads, k, d, t = synth.make_config(cfg, scale)
    gen_s = time.time() - t0
    cand = synth.candidate_pairs(reads.off, k, list(d), list(t))
    wd = tempfile.mkdtemp(prefix=f"links_full_{name}_")
    cbytes = write_fasta_stream(os.path.join(wd, "contigs.fa"), contigs.seq, contigs.off, "", True)
    write_fasta_stream(os.path.join(wd, "reads.fa"), reads.seq, reads.off, "r", False)
    open(os.path.join(wd, "reads.fof"), "w").write("reads.fa\n")
    cmd = [REF_BIN, "-f", "contigs.fa", "-s", "reads.fof", "-b", "out", "-j", str(threads), "-k", str(k),
           "-d", ",".join(map(str, d)), "-t", ",".join(map(str, t))]
    print(f"[{name}] inputs ready in {gen_s:.0f} s ({reads.n} reads, {int(reads.off[-1])} bases, {cand} candidate pairs); running {' '.join(cmd)}",
          flush=True)
    r0 = resource.getrusage(resource.RUSAGE_CHILDREN)
    w0 = time.time()
    res = subprocess.run(cmd, cwd=wd, capture_output=True, text=True)
    wall = time.time() - w0
    r1 = resource.getrusage(resource.RUSAGE_CHILDREN)
    if res.returncode != 0:
        raise SystemExit(f"[{name}] reference failed rc={res.returncode}: {res.stderr[-2000:]}")
    tsv = open(os.path.join(wd, "out.tigpair_checkpoint.tsv"), "rb").read()
    issues = open(os.path.join(wd, "out.pairing_issues"), "rb").read()
    hb = hashlib.sha256()
    with open(os.path.join(wd, "out.bloom"), "rb") as f:
        for blk in iter(lambda: f.read(1 << 24), b""):
            hb.update(blk)
    m = re.search(r"- m: (\d+)", res.stdout)
    h = re.search(r"- hashFct: (\d+)", res.stdout)
    stages = parse_banners(res.stdout)
    stage234 = sum(stages.get(s, 0.0) for s in ("reads_s", "contigs_s", "pairing_s"))
    meta = {
        "workload": cfg, "scale": scale, "opts": {"k": k, "d": list(d), "t": list(t), "o": 0, "e": 0.1, "p": 0.001},
        "inputs_sha256": inputs_digest(contigs, reads), "contigs_file_bytes": cbytes, "n_contigs": contigs.n,
        "n_reads": reads.n, "read_bases": int(reads.off[-1]), "candidate_pairs_closed_form": cand,
        "bf_m_bits": int(m.group(1)), "bf_hash_num": int(h.group(1)),
        "tsv_sha256": hashlib.sha256(b"".join(sorted(tsv.splitlines(keepends=True)))).hexdigest(),
        "issues_sha256": hashlib.sha256(sort_issue_records(issues)).hexdigest(),
        "bloom_sha256": hb.hexdigest(), "n_rows": tsv.count(b"\n"), "n_issues": issues.count(b"Pairs unsat"),
        "stats": parse_stats(res.stdout),
        "cpu": {"threads": threads, "host_cores": os.cpu_count(), "wall_s": round(wall, 2), "stages_s": stages,
                "stage234_s": round(stage234, 2), "pairs_per_s_stage234": cand / stage234 if stage234 else None,
                "pairs_per_s_wall": cand / wall, "peak_rss_gb": round(max(r1.ru_maxrss, r0.ru_maxrss) / 1e6, 2),
                "note": "stages 2 (merge thread) and 4 (pair_contigs) are single-threaded in the reference whatever -j is"},
        "how": f"oracle/_ref/LINKS_CPP -j {threads} (unmodified reference) on links_b200.synth.make_config('{cfg}', {scale}); tools/make_fullsize_golden.py {name}",
    }
    dst = os.path.join(GOLD, "full_" + name)
    os.makedirs(dst, exist_ok=True)
    json.dump(meta, open(os.path.join(dst, "case.json"), "w"), indent=1, sort_keys=True)
    print(f"[{name}] rows={meta['n_rows']} issues={meta['n_issues']} wall={wall:.1f}s stages={stages} rss={meta['cpu']['peak_rss_gb']} GB", flush=True)
    if keep:
        print(f"[{name}] kept {wd}")
    else:
        import shutil
        shutil.rmtree(wd, ignore_errors=True)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("names", nargs="*", default=list(CASES))
    ap.add_argument("-j", type=int, default=1)
    ap.add_argument("--keep", action="store_true")
    a = ap.parse_args()
    if not os.path.exists(REF_BIN):
        raise SystemExit("build the reference first: make -C oracle ref")
    for n in a.names:
        run_case(n, a.j, a.keep)


if __name__ == "__main__":
    main()
