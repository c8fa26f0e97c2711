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


def _check(exe, path, threads, cap, window, max_records=1 << 20):
    r = subprocess.run([exe, path, str(threads), str(cap), str(window), str(max_records)], capture_output=True, text=True)
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
    assert run(b"@r\nACGT\n+\nIIII\n") == ["FALLBACK"]                      # FASTQ stays with the serial reader
    assert run(b">a\nACGT\n", name="e.fa.gz") == ["FALLBACK"]               # so do compressed inputs (by suffix, as btllib decides)
    assert run(b"") == ["FALLBACK"]
    assert run(b">a\n" + b"A" * 2000 + b"\n", cap=1000) == ["ERROR", "both"]  # one record larger than the staging buffer
    # '>' inside a sequence line is data, only a '>' at a line start opens a record
    assert run(b">a\nAC>GT\n>b\nA\n") == ["OK", "2", "6"]
    # max_records = 3 cuts batches by count
    assert run(b"".join(b">%d\nA\n" % i for i in range(10)), cap=1000, window=1 << 20) == ["OK", "10", "10"]
