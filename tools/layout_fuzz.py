#!/usr/bin/env python
"""Differential fuzz of the layout stage: links_b200/bin/LINKS.pl (C++, links_layout.hpp) against the reference's own Perl.

    python tools/layout_fuzz.py [--cases 200] [--seed 1] [--keep]      (build container only: needs oracle/_ref/perl/LINKS.pl)

Each case: 20-90 random contigs of distinct lengths, a random "true" layout (order + strand) whose neighbours get well
supported links, plus spurious and competing links (so that -l and -a decide something), rows mirrored the way the
checkpoint writer mirrors them, random row order, random -l / -a / -z.  The Perl runs under three PERL_HASH_SEEDs; a
case whose Perl output depends on the seed has no single right answer and is skipped.  Otherwise .scaffolds,
.scaffolds.fa, .gv (below its date line) and .assembly_correspondence.tsv (below its first line) must be identical.
Prints one line per mismatch and a summary; exit code 1 if any case differs.
"""
from __future__ import annotations

import argparse
import os
import random
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
PERL = os.path.join(ROOT, "oracle", "_ref", "perl", "LINKS.pl")
SEEDS = ("777", "424242", "deadbeef")
BINS = (10, 500, 5000, 10000)


def below(b: bytes, n: int) -> bytes:
    return b"\n".join(b.split(b"\n")[n:])


def make_case(rng: random.Random):
    n = rng.randint(20, 90)
    lens = rng.sample(range(120, 9000), n)  # distinct: equal lengths are ties the Perl breaks by hash order
    seqs = ["".join(rng.choice("ACGT") for _ in range(L)) for L in lens]
    if rng.random() < 0.3:  # lower case and Ns are copied through
        i = rng.randrange(n)
        seqs[i] = seqs[i][:50].lower() + "N" * 7 + seqs[i][57:]
    order = list(range(1, n + 1))
    rng.shuffle(order)
    strand = {t: rng.choice("fr") for t in order}
    rows = {}

    def add(a_or, a, b_or, b, links, gap, binv):
        # one direction + its mirror (reverse both, swap), as the checkpoint writer emits them
        key = (binv, a_or + str(a), b_or + str(b))
        if key in rows or a == b:
            return
        rows[key] = (links, links * gap)
        flip = {"f": "r", "r": "f"}
        rows[(binv, flip[b_or] + str(b), flip[a_or] + str(a))] = (links, links * gap)

    n_chain_breaks = set(rng.sample(range(n - 1), max(1, n // rng.randint(3, 12))))
    for i in range(n - 1):
        if i in n_chain_breaks:
            continue
        a, b = order[i], order[i + 1]
        gap = rng.choice([rng.randint(-80, 0), rng.randint(0, 300), rng.randint(300, 4000)])
        binv = next((x for x in BINS if gap <= x), 10000) if rng.random() < 0.8 else rng.choice(BINS)
        add(strand[a], a, strand[b], b, rng.randint(1, 40), gap, binv)
        if rng.random() < 0.35:  # the same neighbours in a second distance bin
            add(strand[a], a, strand[b], b, rng.randint(1, 25), gap + rng.randint(-30, 30), rng.choice(BINS))
        if rng.random() < 0.3 and i + 2 < n:  # a link that skips one contig
            c = order[i + 2]
            add(strand[a], a, strand[c], c, rng.randint(1, 12), gap + lens[b - 1] + rng.randint(0, 200), rng.choice(BINS))
    for _ in range(rng.randint(0, n)):  # spurious and competing links
        a, b = rng.sample(order, 2)
        add(rng.choice("fr"), a, rng.choice("fr"), b, rng.randint(1, 15), rng.randint(-100, 5000), rng.choice(BINS))
    lines = [f"{k[0]}\t{k[1]}\t{k[2]}\t{v[0]}\t{v[1]}" for k, v in rows.items()]
    rng.shuffle(lines)
    opts = {"l": rng.choice([1, 2, 3, 5, 8]), "a": rng.choice([0.1, 0.3, 0.5, 0.7, 0.9]), "z": rng.choice([100, 200, 500, 1000, 3000])}
    fasta = "".join(f">{i + 1}\n{s}\n" for i, s in enumerate(seqs))
    return fasta, "\n".join(lines) + "\n", opts


def run(exe_cmd, wd, opts, env=None):
    for f in ("scaffolds", "scaffolds.fa", "gv", "assembly_correspondence.tsv", "log"):
        try:
            os.remove(os.path.join(wd, "out." + f))
        except FileNotFoundError:
            pass
    r = subprocess.run(exe_cmd + ["-f", "contigs.fa", "-s", "empty.fof", "-b", "out", "-l", str(opts["l"]), "-a", str(opts["a"]), "-z", str(opts["z"])],
                       cwd=wd, env=env, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if r.returncode != 0:
        return ("rc", r.returncode)
    rd = lambda s: open(os.path.join(wd, "out." + s), "rb").read()  # noqa: E731
    return (rd("scaffolds"), rd("scaffolds.fa"), below(rd("gv"), 2), below(rd("assembly_correspondence.tsv"), 1))


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=200)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--keep", action="store_true", help="keep the working directory of every mismatch")
    a = ap.parse_args()
    if not os.path.exists(PERL):
        print(f"{PERL} is missing: make -C oracle ref")
        return 2
    from links_b200 import build
    ours = build.build_layout()
    n_bad = n_skip = 0
    for i in range(a.cases):
        rng = random.Random(a.seed * 1_000_003 + i)
        fasta, tsv, opts = make_case(rng)
        wd = tempfile.mkdtemp(prefix="layout_fuzz_")
        open(os.path.join(wd, "contigs.fa"), "w").write(fasta)
        open(os.path.join(wd, "out.tigpair_checkpoint.tsv"), "w").write(tsv)
        open(os.path.join(wd, "empty.fof"), "w").close()
        perl = [run(["perl", PERL], wd, opts, dict(os.environ, PERL_HASH_SEED=s, PERL_PERTURB_KEYS="0")) for s in SEEDS]
        if any(p != perl[0] for p in perl):
            n_skip += 1
            shutil.rmtree(wd, ignore_errors=True)
            continue
        got = run([ours], wd, opts)
        if got != perl[0]:
            n_bad += 1
            which = [n for n, x, y in zip(("scaffolds", "scaffolds.fa", "gv", "correspondence"), got, perl[0]) if x != y] if got[0] != "rc" and perl[0][0] != "rc" else [str(got[:2]), str(perl[0][:2])]
            print(f"case {i} (seed {a.seed}) differs in {which}: opts {opts}" + (f"  kept {wd}" if a.keep else ""), flush=True)
            if a.keep:
                continue
        shutil.rmtree(wd, ignore_errors=True)
    print(f"{a.cases} cases: {n_bad} differ, {n_skip} skipped (the Perl's own output depends on its hash seed)")
    return 1 if n_bad else 0


if __name__ == "__main__":
    sys.exit(main())
