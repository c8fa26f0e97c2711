"""Host ingest: the multi-threaded FASTA reader of LINKS_CPP (links_b200/csrc/host/fasta_ingest.hpp)
must hand the C-ABI exactly the records the serial reader (seq_stream.hpp) does -- which is what btllib::SeqReader
gives the reference in LONG_MODE (src/LINKS.hpp:404-468).  CPU only; tests/host/ingest_check.cpp does the comparing."""
from __future__ import annotations

import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def ingest_check(tmp_path_factory):
    exe = str(tmp_path_factory.mktemp("ingest") / "ingest_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "links_b200", "csrc", "host"),
                    os.path.join(ROOT, "tests", "host", "ingest_check.cpp"), "-o", exe], check=True)
    return exe


def _fasta(rng, n_records, max_len, width, crlf=False, blank=0.0, empty=0.0, final_newline=True) -> bytes:
    nl = b"\r\n" if crlf else b"\n"
    out = []
    alphabet = np.frombuffer(b"ACGTacgtNn", np.uint8)
    for i in range(n_records):
        out.append(b">r%d some description" % i + nl)
        if rng.random() < empty:
            continue
        seq = alphabet[rng.integers(0, len(alphabet), int(rng.integers(1, max_len + 1)))].tobytes()
        w = width if width else len(seq)
        for j in range(0, len(seq), w):
            out.append(seq[j:j + w] + nl)
            if rng.random() < blank:
                out.append(nl)
    data = b"".join(out)
    if not final_newline and data.endswith(nl):
        data = data[:-len(nl)]
    return data


def _check(exe, path, threads, cap, window, max_records=1 << 20, min_segment=1 << 16, grow=False):
    r = subprocess.run([exe, path, str(threads), str(cap), str(window), str(max_records), str(min_segment)] + (["grow"] if grow else []),
                       capture_output=True, text=True, env=dict(os.environ, **({"INGEST_CHECK_VERBOSE": "1"} if grow else {})))
    assert r.returncode == 0, r.stdout + r.stderr
    return r.stdout.split()


@pytest.mark.parametrize("seed", range(6))
def test_parallel_ingest_equals_serial_reader(ingest_check, tmp_path, seed):
    rng = np.random.default_rng(seed)
    data = _fasta(rng, n_records=int(rng.integers(1, 400)), max_len=int(rng.choice([30, 700, 20000])),
                  width=int(rng.choice([0, 7, 60, 80])), crlf=bool(seed % 2), blank=0.05 * (seed % 3), empty=0.1 * (seed % 2),
                  final_newline=bool(seed % 3))
    p = str(tmp_path / "x.fa")
    open(p, "wb").write(data)
    expect = None
    for threads in (1, 3, 8):
        for window in (1, 977, 65536 * 3 + 5, 1 << 30):  # window 1: one record per window; odd sizes put seams mid-line
            for cap in (20001, 1 << 22):
                got = _check(ingest_check, p, threads, cap, window)
                assert got[0] == "OK", got
                expect = expect or got
                assert got == expect


def test_parallel_ingest_edges(ingest_check, tmp_path):
    def run(data: bytes, name="e.fa", **kw):
        p = str(tmp_path / name)
        open(p, "wb").write(data)
        return _check(ingest_check, p, kw.get("threads", 4), kw.get("cap", 1000), kw.get("window", 10), kw.get("max_records", 3))
    assert run(b">a\nAC\n>b\n>c\n\n\nGT\n>d") == ["OK", "3", "4"]          # empty record kept; unterminated last header is no record
    assert run(b"\n\n>a\nACGT") == ["OK", "1", "4"]                         # leading blank lines, no final newline
    assert run(b">only header\n") == ["OK", "1", "0"]
    assert run(b"@r\nACGT\n+\nIIII\n") == ["OK", "1", "4"]                  # strict four-line FASTQ is taken
    assert run(b"@r\nACGT\nAC\n+\nIIIIII\n") == ["FALLBACK"]               # wrapped FASTQ stays with the serial reader
    assert run(b"@r\nACGT\n+\nIII\n") == ["FALLBACK"]                       # so does a truncated quality line
    assert run(b"@r\n\n+\n\n@s\nA\n+\nI\n") == ["FALLBACK"]                # and an empty sequence
    assert run(b"@r\nACGT\n+\nIIII\n>x\nAC\n") == ["FALLBACK"]              # and a file that changes format
    assert run(b">a\nACGT\n", name="e.fa.gz") == ["FALLBACK"]               # so do compressed inputs (by suffix, as btllib decides)
    assert run(b"") == ["FALLBACK"]
    assert run(b">a\n" + b"A" * 2000 + b"\n", cap=1000) == ["ERROR", "both"]  # one record larger than the staging buffer
    # '>' inside a sequence line is data, only a '>' at a line start opens a record
    assert run(b">a\nAC>GT\n>b\nA\n") == ["OK", "2", "6"]
    # max_records = 3 cuts batches by count
    assert run(b"".join(b">%d\nA\n" % i for i in range(10)), cap=1000, window=1 << 20) == ["OK", "10", "10"]


def _fastq(rng, n_records, max_len, crlf=False, final_newline=True, trailing_blank=0) -> bytes:
    nl = b"\r\n" if crlf else b"\n"
    out = []
    bases = np.frombuffer(b"ACGTacgtN", np.uint8)
    quals = np.frombuffer(b"@+@+!I5>#", np.uint8)  # quality lines that start with '@' or '+' are the hard case
    for i in range(n_records):
        n = int(rng.integers(1, max_len + 1))
        out += [b"@r%d/1 @+" % i, nl, bases[rng.integers(0, len(bases), n)].tobytes(), nl, b"+" + (b"r%d" % i if i % 3 == 0 else b""), nl,
                quals[rng.integers(0, len(quals), n)].tobytes(), nl]
    data = b"".join(out)
    if not final_newline:
        data = data[:-len(nl)]
    return data + nl * trailing_blank


@pytest.mark.parametrize("seed", range(6))
def test_parallel_fastq_ingest_equals_serial_reader(ingest_check, tmp_path, seed):
    rng = np.random.default_rng(100 + seed)
    data = _fastq(rng, n_records=int(rng.integers(1, 300)), max_len=int(rng.choice([3, 40, 5000])), crlf=bool(seed % 2),
                  final_newline=bool(seed % 3), trailing_blank=2 if seed == 4 else 0)  # (CRLF blank lines are an error to the serial reader)
    p = str(tmp_path / "x.fq")
    open(p, "wb").write(data)
    expect = None
    for threads in (1, 3, 8, 16):
        for min_segment in (1, 7, 64, 1 << 16):  # 1: as many pieces as threads, cut anywhere -- also inside quality lines
            for cap in (5001, 1 << 22):
                got = _check(ingest_check, p, threads, cap, 1 << 20, min_segment=min_segment)
                assert got[0] in ("OK", "FALLBACK"), got  # FALLBACK (serial reader) is always correct; OK must agree with it
                if got[0] == "OK":
                    expect = expect or got
                    assert got == expect
    assert expect is not None  # at least the one-piece run must have taken the parallel path


def test_parallel_fastq_on_the_golden_fastq_input(ingest_check, tmp_path):
    """tests/golden/adv_offset_k20 has FASTQ reads: the file LINKS_CPP gets in the GPU test."""
    from .conftest import GoldenCase
    c = GoldenCase("adv_offset_k20")
    _cp, fof = c.materialize(str(tmp_path))
    for path in open(fof).read().split():
        if open(path, "rb").read(1) == b"@":
            for min_segment in (1 << 16, 1000):
                assert _check(ingest_check, path, 8, 1 << 22, 1 << 20, min_segment=min_segment)[0] == "OK"


@pytest.mark.parametrize("fmt", ["fasta", "fastq", "fastq-wrapped"])
def test_a_record_larger_than_the_staging_buffer_grows_it(ingest_check, tmp_path, fmt):
    """A chromosome-scale contig must not depend on LINKS_STAGE_MB: both readers ask for larger buffers (LINKS_CPP's grow_stage)
    and carry on -- records before and after the long one included, the serial reader keeping the part it had already copied."""
    rng = np.random.default_rng(7)
    alphabet = np.frombuffer(b"ACGT", np.uint8)
    lens = [50, 900, 30000, 120, 70000, 10]
    seqs = [alphabet[rng.integers(0, 4, n)].tobytes() for n in lens]
    if fmt == "fasta":
        data = b"".join(b">r%d\n" % i + b"\n".join(s[j:j + 70] for j in range(0, len(s), 70)) + b"\n" for i, s in enumerate(seqs))
    elif fmt == "fastq":
        data = b"".join(b"@r%d\n" % i + s + b"\n+\n" + b"I" * len(s) + b"\n" for i, s in enumerate(seqs))
    else:  # wrapped FASTQ: the serial reader's job
        wrap = lambda x: b"\n".join(x[j:j + 100] for j in range(0, len(x), 100))  # noqa: E731
        data = b"".join(b"@r%d\n" % i + wrap(s) + b"\n+\n" + wrap(b"I" * len(s)) + b"\n" for i, s in enumerate(seqs))
    p = str(tmp_path / ("x.fa" if fmt == "fasta" else "x.fq"))
    open(p, "wb").write(data)
    want = (["FALLBACK"] if fmt == "fastq-wrapped" else ["OK"]) + ["6", str(sum(lens))]
    assert _check(ingest_check, p, 4, 1000, 1 << 20, min_segment=1000, grow=True) == want
    if fmt != "fastq-wrapped":
        assert _check(ingest_check, p, 4, 1000, 1 << 20, min_segment=1000, grow=False) == ["ERROR", "both"]
