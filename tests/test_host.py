"""Host-side logic that needs no GPU: option handling, sequence IO, synthetic generator."""
from __future__ import annotations

import gzip
import os

import numpy as np
import pytest

from links_b200 import LinksOptions, SeqSet, read_seqfile, synth, write_fasta


def test_step_padding_matches_inputparser():
    # reference src/InputParser.hpp:224-229,236-240
    assert LinksOptions(distances=[1, 2, 3], step_sizes=[5]).padded_steps() == [5, 5, 5]
    assert LinksOptions(distances=[1, 2, 3], step_sizes=[20, 10]).padded_steps() == [20, 10, 10]
    with pytest.raises(ValueError):
        LinksOptions(distances=[1], step_sizes=[2, 3]).padded_steps()


def test_fasta_fastq_roundtrip(tmp_path):
    recs = [b"ACGTNNacgt", b"", b"GGGG", b"T" * 1000]
    s = SeqSet.from_records(recs)
    p = str(tmp_path / "a.fa")
    write_fasta(p, s)
    assert [read_seqfile(p).record(i) for i in range(4)] == recs
    # multi-line FASTA + CRLF
    open(p, "wb").write(b">x desc\r\nACG\r\nTTA\r\n>y\nGG\n\n")
    t = read_seqfile(p)
    assert [t.record(i) for i in range(t.n)] == [b"ACGTTA", b"GG"] and t.names == ["x", "y"]
    q = str(tmp_path / "a.fq.gz")
    with gzip.open(q, "wb") as f:
        f.write(b"@r1\nACGT\n+\nIIII\n@r2\nGGA\n+r2\n@II\n")
    u = read_seqfile(q)
    assert [u.record(i) for i in range(u.n)] == [b"ACGT", b"GGA"]


def test_concat_offsets():
    a = SeqSet.from_records([b"AC", b"G"])
    b = SeqSet.from_records([b"TTT"])
    c = SeqSet.concat([a, b])
    assert list(c.off) == [0, 2, 3, 6] and c.seq.tobytes() == b"ACGTTT"


def test_synth_is_deterministic_and_shaped():
    g1, c1, r1, k, d, t = synth.make_config("tiny")
    g2, c2, r2, *_ = synth.make_config("tiny")
    assert np.array_equal(g1, g2) and np.array_equal(c1.seq, c2.seq) and np.array_equal(r1.seq, r2.seq)
    assert set(np.unique(r1.seq)) <= set(b"ACGTN")
    cov = len(r1.seq) / len(g1)
    assert 9.5 < cov < 11.5
    assert synth.candidate_pairs(r1.off, k, d, t) > 0


def test_candidate_pair_formula(oracle):
    """The analytic candidate count used for throughput equals the oracle's visit count on N-free reads."""
    g = synth.make_genome(30_000, seed=4)
    reads = synth.make_reads(g, 3, 2500, seed=5, n_read_frac=0.0)
    contigs = synth.make_contigs(g, 4, seed=6, n_run_frac=0.0)
    k, d, t, o = 15, [700, 1500], [2, 7], 3
    _m, nbytes, h = oracle.bf_size(len(contigs.seq), 0.001)
    bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, k)
    stats = oracle.run(k, d, t, o, 0.1, bits, h, contigs.seq, contigs.off, reads.seq, reads.off)[2]
    assert stats["n_candidates"] == synth.candidate_pairs(reads.off, k, d, t, o)


def _fnv1a(b: bytes) -> int:
    h = 1469598103934665603
    for c in b:
        h = ((h ^ c) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def test_cpp_stream_reader_matches_python_reader(tmp_path):
    """The LINKS_CPP host's streaming FASTA/FASTQ reader (csrc/host/seq_stream.hpp) yields the same
    records, in the same order with the same ordinals, as the Python reader -- across plain, gzip,
    multi-line, CRLF and FASTQ inputs, and across staging-buffer rollovers."""
    import subprocess
    from .conftest import ROOT, GoldenCase
    exe = str(tmp_path / "seq_stream_dump")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", os.path.join(ROOT, "tests", "cpp", "seq_stream_dump.cpp"), "-o", exe],
                   check=True)
    rng = np.random.default_rng(9)
    recs = [bytes(rng.choice(np.frombuffer(b"ACGTNacgt", np.uint8), int(n))) for n in (0, 1, 59, 60, 61, 500, 3000, 7)]
    fa = str(tmp_path / "multi.fa")
    with open(fa, "wb") as f:
        for i, r in enumerate(recs):
            f.write(b">r%d some description\r\n" % i)
            for j in range(0, len(r), 60):
                f.write(r[j:j + 60] + b"\r\n")
    fq = str(tmp_path / "x.fq.gz")
    with gzip.open(fq, "wb") as f:
        for i, r in enumerate(recs[1:]):
            f.write(b"@q%d\n" % i + r + b"\n+\n" + b"@" * len(r) + b"\n")   # quality made of '@' to bait the parser
    c = GoldenCase("adv_offset_k20")
    files = [fa, fq] + [os.path.join(c.dir, fn + ".gz") for fn in c.meta["read_files"]]
    expect = []
    for p in files:
        s = read_seqfile(p)
        expect += [s.record(i) for i in range(s.n)]
    longest = max(len(r) for r in expect)
    for cap, max_rec in ((longest + 10, 1 << 20), (longest * 3, 5), (64 << 20, 1 << 20)):
        out = subprocess.run([exe, str(cap), str(max_rec)] + files, check=True, capture_output=True, text=True).stdout.split("\n")
        rows = [ln.split("\t") for ln in out if ln]
        assert [int(r[0]) for r in rows] == list(range(len(expect)))
        assert [int(r[1]) for r in rows] == [len(r) for r in expect]
        assert [int(r[2], 16) for r in rows] == [_fnv1a(r) for r in expect]
    # a record that cannot fit the staging buffer is an error, not a truncation
    bad = subprocess.run([exe, str(longest - 1), "100"] + files, capture_output=True, text=True)
    assert bad.returncode == 1 and "larger than the staging buffer" in bad.stderr


def test_links_cpp_cli_contract_without_gpu(tmp_path):
    """Option parsing and error behaviour of the C++ host that need no device: usage on missing
    -f/-s, empty .fof, more steps than distances (reference src/InputParser.hpp:230-255), and a
    loud non-zero exit (no CPU fallback) when no GPU is present."""
    import subprocess
    from links_b200 import build
    from .conftest import have_gpu
    exe = build.HOST_BIN
    if not os.path.exists(exe):
        build.build_all()
    r = subprocess.run([exe], capture_output=True, text=True)
    assert r.returncode != 0 and "Usage:" in r.stdout
    fof = tmp_path / "empty.fof"; fof.write_text("")
    fa = tmp_path / "c.fa"; fa.write_text(">1\nACGTACGTACGTACGTACGTACGT\n")
    r = subprocess.run([exe, "-f", str(fa), "-s", str(fof)], capture_output=True, text=True)
    assert r.returncode != 0 and "cannot be empty" in r.stderr
    fof.write_text(str(fa) + "\n")
    r = subprocess.run([exe, "-f", str(fa), "-s", str(fof), "-d", "100", "-t", "2,3"], capture_output=True, text=True)
    assert r.returncode != 0 and "can't be lower than number of step sizes" in r.stderr
    if not have_gpu():
        r = subprocess.run([exe, "-f", str(fa), "-s", str(fof), "-b", str(tmp_path / "o")], capture_output=True, text=True)
        assert r.returncode == 1 and "no CPU path" in r.stderr
        assert "-d \n" in r.stdout.replace("-d  \n", "-d \n") and "-k 15" in r.stdout and "-t 2" in r.stdout
