#!/usr/bin/env python
"""Goldens for the layout stage (links_b200/csrc/host/links_layout.hpp) from the reference's own Perl.

    python tools/make_layout_golden.py        (build container only: needs /root/reference and perl)

 c1            the installation test's shipped outputs (reference tests/expected_outputs/): .scaffolds,
               sha256 of .scaffolds.fa, .gv below its date line, .assembly_correspondence.tsv
 cfg_<name>    the unmodified bin/LINKS.pl run on the reference's tigpair_checkpoint.tsv of the BASELINE-config
               cases (tests/golden/cfg_*/expected.tigpair_checkpoint.tsv) with the synthetic contigs, options
               -l 5 -a 0.3 -z 500.  Perl's hash order is random per process and decides ties, so LINKS.pl runs
               under five different PERL_HASH_SEEDs; a case whose outputs differ between them is reported and
               NOT stored (there is no single right answer to pin).
 lay_stress_*  300 synthetic contigs of distinct lengths (links_b200.synth, seeded; ties would make the Perl's own
               output vary), links from the unmodified LINKS_CPP at three distances, laid out under three -l/-a/-z
               settings: tens of scaffolds, chains extended on both sides, reverse-strand joins, overlaps (m < 0).
Stored under tests/golden/<case>/layout/.
"""
from __future__ import annotations

import hashlib
import json
import os
import shutil
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tools"))
REF = "/root/reference"
GOLD = os.path.join(ROOT, "tests", "golden")
# hex digits; picked so that `keys` of a small hash really comes out in different orders (1 / 12345 / 987654321 do not)
SEEDS = ("1", "777", "424242", "deadbeef", "0badc0de")


def below_label(gv: bytes) -> bytes:
    return b"\n".join(gv.split(b"\n")[2:])


def store(case: str, files: dict[str, bytes], how: str, opts: dict) -> None:
    d = os.path.join(GOLD, case, "layout")
    shutil.rmtree(d, ignore_errors=True)
    os.makedirs(d)
    open(os.path.join(d, "expected.scaffolds"), "wb").write(files["scaffolds"])
    open(os.path.join(d, "expected.gv_body"), "wb").write(below_label(files["gv"]))
    corr = files["corr"].split(b"\n", 1)[1]  # the first line names the output file
    open(os.path.join(d, "expected.assembly_correspondence_body.tsv"), "wb").write(corr)
    meta = {"scaffolds_fa_sha256": hashlib.sha256(files["fa"]).hexdigest(), "scaffolds_fa_bytes": len(files["fa"]),
            "n_scaffolds": files["scaffolds"].count(b"\n"), "opts": opts, "how": how}
    json.dump(meta, open(os.path.join(d, "layout.json"), "w"), indent=1, sort_keys=True)
    merged = sum(1 for ln in files["scaffolds"].split(b"\n") if b"_" in ln)
    print(f"{case}: {meta['n_scaffolds']} scaffolds, {merged} with more than one contig, fa {len(files['fa'])} bytes")


def run_perl(contigs_fa: str, tsv: bytes, opts: dict, seed: str) -> dict[str, bytes]:
    with tempfile.TemporaryDirectory() as wd:
        shutil.copy(contigs_fa, os.path.join(wd, "contigs.fa"))
        open(os.path.join(wd, "out.tigpair_checkpoint.tsv"), "wb").write(tsv)
        open(os.path.join(wd, "empty.fof"), "w").close()
        env = dict(os.environ, PERL_HASH_SEED=seed, PERL_PERTURB_KEYS="0")
        subprocess.run(["perl", f"{REF}/bin/LINKS.pl", "-f", "contigs.fa", "-s", "empty.fof", "-b", "out", "-l", str(opts["l"]),
                        "-a", str(opts["a"]), "-z", str(opts["z"])], cwd=wd, env=env, check=True, stdout=subprocess.DEVNULL,
                       stderr=subprocess.DEVNULL)
        rd = lambda s: open(os.path.join(wd, "out." + s), "rb").read()  # noqa: E731
        return {"scaffolds": rd("scaffolds"), "fa": rd("scaffolds.fa"), "gv": rd("gv"), "corr": rd("assembly_correspondence.tsv")}


def main() -> None:
    exp = f"{REF}/tests/expected_outputs/installation_test"
    rd = lambda s: open(exp + "." + s, "rb").read()  # noqa: E731
    store("c1", {"scaffolds": rd("scaffolds"), "fa": rd("scaffolds.fa"), "gv": rd("gv"), "corr": rd("assembly_correspondence.tsv")},
          "reference tests/expected_outputs/installation_test.* as shipped", {"l": 5, "a": 0.3, "z": 500})
    # .gv's `type=` keeps the LAST of equally supported orientations (bin/LINKS.pl:1207), so it depends on the row order
    # of the checkpoint file: keep the file in the order the reference shipped it (tests/golden/c1 stores it sorted)
    shutil.copy(exp + ".tigpair_checkpoint.tsv", os.path.join(GOLD, "c1", "layout", "input.tigpair_checkpoint.tsv"))
    from make_config_golden import CASES, make_inputs
    from make_golden import fasta_bytes
    for name in CASES:
        contigs, _reads, _k, _d, _t = make_inputs(name)
        tsv = open(os.path.join(GOLD, "cfg_" + name, "expected.tigpair_checkpoint.tsv"), "rb").read()
        opts = {"l": 5, "a": 0.3, "z": 500}
        with tempfile.TemporaryDirectory() as wd:
            fa = os.path.join(wd, "contigs.fa")
            open(fa, "wb").write(fasta_bytes(contigs))
            runs = [run_perl(fa, tsv, opts, seed) for seed in SEEDS]
        same = all(r["scaffolds"] == runs[0]["scaffolds"] and r["fa"] == runs[0]["fa"] and r["corr"] == runs[0]["corr"]
                   and below_label(r["gv"]) == below_label(runs[0]["gv"]) for r in runs)
        if not same:
            print(f"cfg_{name}: LINKS.pl's own output depends on its hash seed (ties) -- not stored")
            continue
        store("cfg_" + name, runs[0], "perl /root/reference/bin/LINKS.pl (unmodified) on expected.tigpair_checkpoint.tsv, identical "
              "under five PERL_HASH_SEEDs, tools/make_layout_golden.py", opts)
    # the adversarial cases (Ns, lower case, tandem repeats, short contigs) with non-default -l -a -z
    import gzip
    for name in ("adv_k40", "adv_ns_lower", "adv_offset_k20", "adv_tandem"):
        tsv = open(os.path.join(GOLD, name, "expected.tigpair_checkpoint.tsv"), "rb").read()
        opts = {"l": 2, "a": 0.5, "z": 200}
        with tempfile.TemporaryDirectory() as wd:
            fa = os.path.join(wd, "contigs.fa")
            open(fa, "wb").write(gzip.open(os.path.join(GOLD, name, "contigs.fa.gz"), "rb").read())
            runs = [run_perl(fa, tsv, opts, seed) for seed in SEEDS]
        same = all(r["scaffolds"] == runs[0]["scaffolds"] and r["fa"] == runs[0]["fa"] and r["corr"] == runs[0]["corr"]
                   and below_label(r["gv"]) == below_label(runs[0]["gv"]) for r in runs)
        if not same:
            print(f"{name}: LINKS.pl's own output depends on its hash seed (ties) -- not stored")
            continue
        store(name, runs[0], "perl /root/reference/bin/LINKS.pl (unmodified), -l 2 -a 0.5 -z 200, identical under five PERL_HASH_SEEDs, "
              "tools/make_layout_golden.py", opts)
    stress()


STRESS = {"genome": 3_000_000, "contigs": 300, "seeds": [21, 23, 22], "coverage": 8.0, "read_len": 15_000,
          "links_opts": {"k": 15, "d": [2000, 5000, 10000], "t": [10, 10, 10], "o": 0, "e": 0.1, "p": 0.001}}
STRESS_LAYOUTS = {"lay_stress_default": {"l": 5, "a": 0.3, "z": 500}, "lay_stress_loose": {"l": 2, "a": 0.7, "z": 200},
                  "lay_stress_strict": {"l": 10, "a": 0.1, "z": 1000}}


def stress_contigs():
    from links_b200 import synth
    genome = synth.make_genome(STRESS["genome"], seed=STRESS["seeds"][0])
    contigs = synth.make_contigs(genome, STRESS["contigs"], seed=STRESS["seeds"][1])
    contigs.names = [str(i + 1) for i in range(contigs.n)]
    return genome, synth.unique_lengths(contigs)


def stress() -> None:
    from links_b200 import synth
    from make_golden import fasta_bytes, run_ref
    genome, contigs = stress_contigs()
    reads = synth.make_reads(genome, STRESS["coverage"], STRESS["read_len"], seed=STRESS["seeds"][2])
    reads.names = [f"r{i}" for i in range(reads.n)]
    with tempfile.TemporaryDirectory() as wd:
        fa = os.path.join(wd, "contigs.fa")
        open(fa, "wb").write(fasta_bytes(contigs))
        open(os.path.join(wd, "reads.fa"), "wb").write(fasta_bytes(reads))
        tsv = run_ref(wd, fa, [os.path.join(wd, "reads.fa")], STRESS["links_opts"])["tsv"]
        first = None
        for case, opts in STRESS_LAYOUTS.items():
            runs = [run_perl(fa, tsv, opts, seed) for seed in SEEDS]
            same = all(r["scaffolds"] == runs[0]["scaffolds"] and r["fa"] == runs[0]["fa"] and r["corr"] == runs[0]["corr"]
                       and below_label(r["gv"]) == below_label(runs[0]["gv"]) for r in runs)
            if not same:
                print(f"{case}: LINKS.pl's own output depends on its hash seed (ties) -- not stored")
                continue
            store(case, runs[0], "perl /root/reference/bin/LINKS.pl (unmodified) on the unmodified LINKS_CPP's checkpoint for 300 "
                  "synthetic contigs of distinct lengths, identical under five PERL_HASH_SEEDs, tools/make_layout_golden.py", opts)
            d = os.path.join(GOLD, case, "layout")
            meta = json.load(open(os.path.join(d, "layout.json")))
            meta["stress"] = STRESS
            if first is None:
                first = case
                open(os.path.join(d, "input.tigpair_checkpoint.tsv"), "wb").write(tsv)
            else:
                meta["input_from"] = first  # one copy of the 85 KB checkpoint
            json.dump(meta, open(os.path.join(d, "layout.json"), "w"), indent=1, sort_keys=True)


if __name__ == "__main__":
    main()
