"""The layout stage (links_b200/csrc/host/links_layout.hpp = the reference's bin/LINKS.pl in C++) against the
reference's own outputs: the installation test's shipped files, and the unmodified Perl run on the other golden
cases (tools/make_layout_golden.py).  CPU only."""
from __future__ import annotations

import gzip
import hashlib
import json
import os
import shutil
import subprocess

import pytest

from .conftest import GOLDEN, ConfigCase

CASES = sorted(d for d in os.listdir(GOLDEN) if os.path.exists(os.path.join(GOLDEN, d, "layout", "layout.json")))


@pytest.fixture(scope="module")
def layout_bin():
    from links_b200 import build
    return build.build_layout()


def _below_label(gv: bytes) -> bytes:
    return b"\n".join(gv.split(b"\n")[2:])


def _run(layout_bin, wd, contigs_name, opts, exe_name=None):
    exe = layout_bin
    if exe_name:  # called under another name, as LINKS-make calls "LINKS.pl"
        exe = os.path.join(wd, exe_name)
        shutil.copy2(layout_bin, exe)
    open(os.path.join(wd, "empty.fof"), "w").close()
    subprocess.run([exe, "-f", contigs_name, "-s", "empty.fof", "-b", "out", "-l", str(opts["l"]), "-a", str(opts["a"]), "-z", str(opts["z"]),
                    "-e", "0.1", "-p", "0.001", "-v", "0"], cwd=wd, check=True, stdout=subprocess.DEVNULL)
    rd = lambda s: open(os.path.join(wd, "out." + s), "rb").read()  # noqa: E731
    return rd("scaffolds"), rd("scaffolds.fa"), rd("gv"), rd("assembly_correspondence.tsv")


@pytest.mark.parametrize("case", CASES)
def test_layout_matches_reference_perl(layout_bin, tmp_path, case):
    d = os.path.join(GOLDEN, case, "layout")
    meta = json.load(open(os.path.join(d, "layout.json")))
    wd = str(tmp_path)
    d_in = d
    if case.startswith("cfg_"):
        from links_b200.seqio import write_fasta
        contigs, _reads = ConfigCase(case[4:]).inputs()
        contigs.names = [str(i + 1) for i in range(contigs.n)]
        write_fasta(os.path.join(wd, "contigs.fa"), contigs)
        name = "contigs.fa"
    elif "stress" in meta:  # seeded synthetic contigs of distinct lengths (tools/make_layout_golden.py: stress_contigs)
        from links_b200 import synth
        from links_b200.seqio import write_fasta
        st = meta["stress"]
        contigs = synth.make_contigs(synth.make_genome(st["genome"], seed=st["seeds"][0]), st["contigs"], seed=st["seeds"][1])
        contigs.names = [str(i + 1) for i in range(contigs.n)]
        write_fasta(os.path.join(wd, "contigs.fa"), synth.unique_lengths(contigs))
        name = "contigs.fa"
        if "input_from" in meta:
            d_in = os.path.join(GOLDEN, meta["input_from"], "layout")
    elif case == "c1":  # read through the gunzip pipe, as LINKS.pl does for *.gz (bin/LINKS.pl:1253)
        shutil.copy(os.path.join(GOLDEN, case, "contigs.fa.gz"), os.path.join(wd, "contigs.fa.gz"))
        name = "contigs.fa.gz"
    else:
        open(os.path.join(wd, "contigs.fa"), "wb").write(gzip.open(os.path.join(GOLDEN, case, "contigs.fa.gz"), "rb").read())
        name = "contigs.fa"
    tsv = os.path.join(d_in, "input.tigpair_checkpoint.tsv")  # c1: the rows in the order the reference shipped them
    if not os.path.exists(tsv):
        tsv = os.path.join(GOLDEN, case, "expected.tigpair_checkpoint.tsv")
    shutil.copy(tsv, os.path.join(wd, "out.tigpair_checkpoint.tsv"))
    sc, fa, gv, corr = _run(layout_bin, wd, name, meta["opts"], exe_name="LINKS.pl" if case == "c1" else None)
    assert sc == open(os.path.join(d, "expected.scaffolds"), "rb").read()
    assert hashlib.sha256(fa).hexdigest() == meta["scaffolds_fa_sha256"] and len(fa) == meta["scaffolds_fa_bytes"]
    assert _below_label(gv) == open(os.path.join(d, "expected.gv_body"), "rb").read()
    assert corr.split(b"\n", 1)[1] == open(os.path.join(d, "expected.assembly_correspondence_body.tsv"), "rb").read()
    assert sc.count(b"\n") == meta["n_scaffolds"]


def test_layout_option_and_input_quirks(layout_bin, tmp_path):
    """`-l 0`, `-a 0`, `-z 0` leave the defaults (Perl: `$min_links = $opt_l if($opt_l)`, bin/LINKS.pl:52-55); a missing
    checkpoint is not an error (:121-124); rows naming an unknown or empty contig id are skipped (:1166-1168)."""
    wd = str(tmp_path)
    open(os.path.join(wd, "c.fa"), "w").write(">a\n" + "ACGT" * 200 + "\n>b\n" + "TTGCA" * 200 + "\n>c\nACG\n")
    open(os.path.join(wd, "empty.fof"), "w").close()
    r = subprocess.run([layout_bin, "-f", "c.fa", "-b", "out"], cwd=wd, capture_output=True, text=True)
    assert r.returncode == 0 and "No checkpoint provided" in r.stdout and not os.path.exists(os.path.join(wd, "out.scaffolds"))
    rows = ["500\tf1\tf2\t7\t700", "500\tr2\tr1\t7\t700", "500\tf1\tf\t9\t90", "500\tf1\tf3\t50\t50", "500\tf1\tf99\t50\t50"]
    open(os.path.join(wd, "out.tigpair_checkpoint.tsv"), "w").write("\n".join(rows) + "\n")
    subprocess.run([layout_bin, "-f", "c.fa", "-s", "empty.fof", "-b", "out", "-l", "0", "-a", "0", "-z", "0"], cwd=wd, check=True,
                   stdout=subprocess.DEVNULL)
    sc = open(os.path.join(wd, "out.scaffolds")).read().splitlines()
    # b (1000 bp) seeds first, a (800 bp) joins on its left with 7 >= 5 links, mean gap 100; c (3 bp < 500) stays alone
    assert sc == ["scaffold1,1800,f1z800k7a0m100_f2Z1000", "scaffold2,3,f3Z3"]
    fa = open(os.path.join(wd, "out.scaffolds.fa")).read().splitlines()
    assert fa[1] == "ACGT" * 200 + "N" * 100 + "TTGCA" * 200


@pytest.mark.parametrize("case", ["c1", "lay_stress_loose", "adv_ns_lower"])
def test_layout_from_rows_in_memory_equals_layout_from_tsv(tmp_path, case):
    """LINKS_CPP can hand its rows to the layout stage in memory (LINKS_CPP_LAYOUT=1); the files must
    be the ones the stage writes after reading the tsv those rows print as."""
    ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    exe = str(tmp_path / "layout_rows_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "links_b200", "csrc", "host"),
                    os.path.join(ROOT, "tests", "host", "layout_rows_check.cpp"), "-o", exe], check=True)
    d = os.path.join(GOLDEN, case, "layout")
    meta = json.load(open(os.path.join(d, "layout.json")))
    wd = str(tmp_path)
    d_in = os.path.join(GOLDEN, meta.get("input_from", case), "layout")
    tsv = os.path.join(d_in, "input.tigpair_checkpoint.tsv")
    if not os.path.exists(tsv):
        tsv = os.path.join(GOLDEN, case, "expected.tigpair_checkpoint.tsv")
    if "stress" in meta:
        from links_b200 import synth
        from links_b200.seqio import write_fasta
        st = meta["stress"]
        contigs = synth.make_contigs(synth.make_genome(st["genome"], seed=st["seeds"][0]), st["contigs"], seed=st["seeds"][1])
        contigs.names = [str(i + 1) for i in range(contigs.n)]
        write_fasta(os.path.join(wd, "contigs.fa"), synth.unique_lengths(contigs))
    else:
        open(os.path.join(wd, "contigs.fa"), "wb").write(gzip.open(os.path.join(GOLDEN, case, "contigs.fa.gz"), "rb").read())
    o = meta["opts"]
    subprocess.run([exe, "contigs.fa", tsv, "file", "mem", str(o["l"]), str(o["a"]), str(o["z"])], cwd=wd, check=True)
    rd = lambda b, s: open(os.path.join(wd, b + "." + s), "rb").read()  # noqa: E731
    assert rd("mem", "scaffolds") == rd("file", "scaffolds") == open(os.path.join(d, "expected.scaffolds"), "rb").read()
    assert rd("mem", "scaffolds.fa") == rd("file", "scaffolds.fa")
    assert _below_label(rd("mem", "gv")) == _below_label(rd("file", "gv")) == open(os.path.join(d, "expected.gv_body"), "rb").read()
    assert rd("mem", "assembly_correspondence.tsv").split(b"\n", 1)[1] == rd("file", "assembly_correspondence.tsv").split(b"\n", 1)[1]


def test_unmodified_wrapper_chain_finds_our_layout_program(tmp_path):
    """`LINKS` -> `LINKS-make` (the reference's wrappers, installed unmodified into oracle/_ref/bin) with THIS repo's C++
    `LINKS.pl` on PATH and the tsv already there (make skips the LINKS_CPP rule: the checkpoint semantics of
    bin/LINKS-make:121): the layout stage must be found, take LINKS-make's fixed option order (-f -s -b -l -a -e -z -p -v,
    bin/LINKS-make:125) and write the installation test's scaffolds.  CPU only; the GPU half of the chain is
    tests/test_gpu_parity.py::test_unmodified_wrapper_chain_over_gpu_binary."""
    import gzip
    import hashlib
    import json
    import shutil
    import subprocess
    import time
    from links_b200 import build
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    ref = os.path.join(root, "oracle", "_ref")
    if not os.path.exists(os.path.join(ref, "bin", "LINKS-make")) or not shutil.which("make"):
        pytest.skip("oracle/_ref/bin/LINKS-make (the reference's wrappers) or make not available")
    build.build_layout()
    gold = os.path.join(root, "tests", "golden", "c1")
    for fn, dst in (("contigs.fa.gz", "simulated_contigs.fa"), ("simulated_genome.fa.gz", "simulated_genome.fa")):
        open(tmp_path / dst, "wb").write(gzip.open(os.path.join(gold, fn), "rb").read())
    open(tmp_path / "reads.fof", "w").write("simulated_genome.fa\n")
    time.sleep(0.05)
    shutil.copy(os.path.join(gold, "layout", "input.tigpair_checkpoint.tsv"), tmp_path / "installation_test.tigpair_checkpoint.tsv")
    fake = tmp_path / "nobin"  # LINKS_CPP must NOT be needed: make sees an up-to-date checkpoint
    fake.mkdir()
    path = os.pathsep.join([os.path.dirname(build.LAYOUT_BIN), str(fake), os.path.join(ref, "bin"), os.environ.get("PATH", "")])
    r = subprocess.run(["LINKS", "-f", "simulated_contigs.fa", "-s", "reads.fof", "-b", "installation_test"], cwd=tmp_path,
                       capture_output=True, text=True, env=dict(os.environ, PATH=path))
    assert r.returncode == 0, r.stdout[-1500:] + r.stderr[-1500:]
    meta = json.load(open(os.path.join(gold, "layout", "layout.json")))
    assert hashlib.sha256(open(tmp_path / "installation_test.scaffolds.fa", "rb").read()).hexdigest() == meta["scaffolds_fa_sha256"]


def test_layout_differential_fuzz_against_reference_perl():
    """tools/layout_fuzz.py: random contigs, random true layouts with spurious and competing links, random -l / -a / -z; the
    C++ layout must write the files the unmodified Perl writes.  Needs oracle/_ref/perl/LINKS.pl (the reference's file,
    installed by oracle/Makefile in the build container), so it is skipped where that is absent.  1840 cases of the same
    generator were run when this was added: none differed."""
    import sys
    from .conftest import ROOT
    perl = os.path.join(ROOT, "oracle", "_ref", "perl", "LINKS.pl")
    if not os.path.exists(perl) or shutil.which("perl") is None:
        pytest.skip("the reference's LINKS.pl is not installed here (oracle/_ref/perl)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "layout_fuzz.py"), "--cases", "12", "--seed", "5"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "12 cases: 0 differ" in r.stdout
