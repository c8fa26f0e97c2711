#!/usr/bin/env python
"""Differential fuzz of the command line: links_b200/bin/LINKS_CPP against the unmodified reference binary.

    python tools/cli_fuzz.py [--cases 300] [--seed 1]       (build container only: needs oracle/_ref/LINKS_CPP; no GPU needed)

What can be compared without a GPU is everything InputParser does (reference src/InputParser.hpp): random option lines --
valid and invalid -d / -t lists, missing option values, unknown options, -h, omitted -b (the derived base name), numeric
options with junk -- must give
  * the same "LINKS running options:" block on stdout, and
  * when the reference stops before its first stage (usage, argument errors): the same exit code and the same message.
Past the options block the GPU binary stops here with LINKS_E_NODEVICE ("no CPU path") where the reference goes on to
compute; that part is the GPU parity tests' business.  Exit code 1 if any command line differs.

Left out on purpose: --help (the reference parses it into a flag nobody reads and carries on; this binary prints the usage),
-j 0 (the reference hangs), the wording of the usage text (this binary says what it does with -x and -j).  An -p outside
(0, 1) fails in both, differently: the reference in BloomFilter's constructor, this binary in links_bf_alloc.
"""
from __future__ import annotations

import argparse
import os
import random
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
REF = os.path.join(ROOT, "oracle", "_ref", "LINKS_CPP")
OURS = os.path.join(ROOT, "links_b200", "bin", "LINKS_CPP")
SEQ = "ACGTACGTACGTACGTACGTACGTAAGGCCTTAAGGCCTTACGATCGATCGATCGGGATATATCGCGCGATATCGCGC"


def options_block(out: str) -> str:
    """stdout up to (not including) the first STAGE banner / the first line that is not part of the options block."""
    keep = []
    for ln in out.split("\n"):
        if ln.startswith("STAGE:") or ln.startswith("LINKS_CPP:"):
            break
        keep.append(ln)
    return "\n".join(keep).rstrip("\n")


def make_args(rng: random.Random) -> list[str]:
    def int_list():
        n = rng.randint(1, 4)
        xs = [str(rng.choice([1, 2, 5, 10, 20, 30, 100, 4000])) for _ in range(n)]
        s = ",".join(xs)
        r = rng.random()
        if r < 0.05:
            s += ","
        elif r < 0.1:
            s = s.replace(",", ", ", 1)
        elif r < 0.14:
            s = "x" + s
        return s

    args = []
    if rng.random() < 0.92:
        args += ["-f", rng.choice(["c.fa", "c.fa", "c.fa", "missing.fa"])]
    if rng.random() < 0.92:
        args += ["-s", rng.choice(["r.fof", "r.fof", "r.fof", "empty.fof", "missing.fof"])]
    if rng.random() < 0.7:
        args += ["-d", int_list()]
    if rng.random() < 0.5:
        args += ["-t", int_list()]
    if rng.random() < 0.6:
        args += ["-k", rng.choice(["11", "15", "20", "0", "abc", "15.5", "-3"])]
    for opt, vals in (("-o", ["0", "3", "x"]), ("-e", ["0.1", "0.25", "1", "e"]), ("-l", ["5", "2", "0"]), ("-a", ["0.3", "0.7", "2"]),
                      ("-z", ["500", "100", "z"]), ("-p", ["0.001", "0.01", "0.1"]), ("-x", ["0", "1"]), ("-v", ["0", "1"]),
                      ("-j", ["1", "3", "2"]), ("-r", ["nope.bloom"])):
        if rng.random() < 0.18:
            args += [opt, rng.choice(vals)]
    if rng.random() < 0.6:
        args += ["-b", rng.choice(["out", "sub dir", "o.u.t"])]
    if rng.random() < 0.06:
        args += [rng.choice(["-h", "-q", "-Z", "stray"])]  # (not --help: the reference parses and ignores it, this binary prints the usage)
    if rng.random() < 0.05 and args:
        args = args[:-1] if args[-1] not in ("-h", "--help") else args  # an option that lost its value
    if rng.random() < 0.3:
        i = list(range(0, len(args) - 1, 2))
        rng.shuffle(i)
        args = [y for j in i for y in args[j:j + 2]] + (args[len(args) - (len(args) % 2):] if len(args) % 2 else [])
    return args


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=300)
    ap.add_argument("--seed", type=int, default=1)
    a = ap.parse_args()
    if not os.path.exists(REF):
        print(f"{REF} is missing: make -C oracle ref")
        return 2
    if not os.path.exists(OURS):
        sys.path.insert(0, ROOT)
        from links_b200 import build
        build.build_all()
    n_bad = n_err = 0
    for i in range(a.cases):
        rng = random.Random(a.seed * 1_000_003 + i)
        args = make_args(rng)
        outs = []
        for exe in (REF, OURS):
            with tempfile.TemporaryDirectory() as wd:
                open(os.path.join(wd, "c.fa"), "w").write(">1\n" + SEQ + "\n")
                open(os.path.join(wd, "r.fa"), "w").write(">r0\n" + SEQ + "\n")
                open(os.path.join(wd, "r.fof"), "w").write("r.fa\n")
                open(os.path.join(wd, "empty.fof"), "w").close()
                try:
                    r = subprocess.run([exe] + args, cwd=wd, capture_output=True, text=True, timeout=60)
                    outs.append((r.returncode, r.stdout, r.stderr))
                except subprocess.TimeoutExpired:
                    outs.append((-999, "", "timeout"))
        (rc_r, so_r, se_r), (rc_o, so_o, se_o) = outs
        early = "STAGE:" not in so_r  # the reference stopped in its argument handling
        se_r, se_o = se_r.replace(REF, "LINKS_CPP"), se_o.replace(OURS, "LINKS_CPP")  # getopt's messages start with argv[0]
        why = []
        usage = "Usage:" in so_r  # the usage text is worded differently here (it says what this binary does with -x and -j)
        if (usage != ("Usage:" in so_o)) or (not usage and options_block(so_r) != options_block(so_o)):
            why.append("options block / stdout")
        if early:
            n_err += 1
            if rc_r != rc_o:
                why.append(f"exit code {rc_r} vs {rc_o}")
            if se_r.strip() != se_o.strip():
                why.append("stderr")
        if why:
            n_bad += 1
            print(f"case {i} (seed {a.seed}): {' '.join(repr(x) for x in args)}\n   differs in {why}\n   ref  rc={rc_r} out={so_r[-160:]!r} err={se_r[-200:]!r}\n"
                  f"   ours rc={rc_o} out={so_o[-160:]!r} err={se_o[-200:]!r}", flush=True)
    print(f"{a.cases} command lines: {n_bad} differ ({n_err} of them stopped in the reference's argument handling)")
    return 1 if n_bad else 0


if __name__ == "__main__":
    sys.exit(main())
