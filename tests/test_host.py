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
