#!/usr/bin/env python
"""Differential fuzz of the host's sequence readers against the reference's own reader (btllib::SeqReader, LONG_MODE).

    python tools/reader_fuzz.py [--cases 200] [--seed 1] [--keep]     (build container only: needs oracle/_ref/libbtl_probe.so)

Random files in the shapes the reference accepts: one-line and wrapped FASTA, four-line FASTQ, LF and CRLF line ends, a
missing final newline, lower case / N / IUPAC, header descriptions, records of length 1 and of tens of kilobases, records
with an empty sequence, gzip now and then.  Every record (length, FNV-1a of its bytes, ordinal) read by
links_b200/csrc/host/seq_stream.hpp (through tests/cpp/seq_stream_dump.cpp, with a small and a large staging buffer), by
the multi-threaded ingest (tests/host/ingest_check.cpp compares it with the serial reader record by record) and by the Python
reader (links_b200/seqio.py) must be the one btllib hands to the reference.  Exit code 1 if any file differs.

One difference is known and kept (--empty-records reproduces it): a FASTA header that is followed directly by the next
header.  btllib reads the line after a header as sequence whatever it starts with (seq_reader_multiline_fasta_module.hpp:
97-110, seq_reader_fasta_module.hpp:82-88), so beyond the 16 KB it sniffs the format on -- inside them it refuses the whole
file -- the record swallows the next header line (">r9" becomes four invalid bases and r9's bases are appended to r8), and
every later contig id shifts by one.  The readers here yield an empty record.  Such a file is malformed for both.
"""
from __future__ import annotations

import argparse
import ctypes as C
import gzip
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def fnv1a(b: bytes) -> int:
    h = 1469598103934665603
    for x in b:
        h = ((h ^ x) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def btllib_records_here(path: str):
    from oracle import oracle_py
    lib = C.CDLL(oracle_py.REF_PROBE)
    lib.btl_read_records.restype = C.c_uint64
    lib.btl_read_records.argtypes = [C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    cap = 1 << 14
    lens, fnv, nums = (np.zeros(cap, np.uint64) for _ in range(3))
    n = lib.btl_read_records(path.encode(), cap, lens.ctypes.data, fnv.ctypes.data, nums.ctypes.data)
    assert n <= cap
    return [(int(lens[i]), int(fnv[i]), int(nums[i])) for i in range(n)]


def btllib_records(path: str):
    """In a child process: btllib ends the process when it does not recognise the format.  None = btllib refused the file."""
    r = subprocess.run([sys.executable, os.path.abspath(__file__), "--probe", path], capture_output=True, text=True)
    if r.returncode != 0:
        return None
    return [tuple(x) for x in json.loads(r.stdout)]


def make_file(rng: np.random.Generator, empty_records: bool = False) -> tuple[bytes, str]:
    fastq = rng.random() < 0.35
    eol = b"\r\n" if rng.random() < 0.25 else b"\n"
    wrap = 0 if fastq or rng.random() < 0.4 else int(rng.choice([1, 7, 60, 80, 1000]))
    alphabet = np.frombuffer(b"ACGT" * 6 + b"acgtNnRYKM", np.uint8)
    n_rec = int(rng.integers(1, 60))
    out = []
    for i in range(n_rec):
        ln = int(rng.choice([1, 2, 15, 100, 1000, 5000, 40000], p=[0.05, 0.05, 0.1, 0.3, 0.3, 0.15, 0.05]))
        if rng.random() < 0.02 and i > 0 and not fastq and empty_records:
            ln = 0  # a header without sequence lines: the KNOWN difference (see the module docstring)
        seq = alphabet[rng.integers(0, len(alphabet), ln)].tobytes()
        name = b"r%d" % i + (b" some description %d" % int(rng.integers(0, 1000)) if rng.random() < 0.3 else b"")
        if fastq:
            qual = bytes(rng.integers(33, 74, ln, dtype=np.uint8))
            out.append(b"@" + name + eol + seq + eol + b"+" + (name if rng.random() < 0.2 else b"") + eol + qual + eol)
        else:
            out.append(b">" + name + eol)
            if wrap and ln:
                out.append(eol.join(seq[j:j + wrap] for j in range(0, ln, wrap)) + eol)
            elif ln:
                out.append(seq + eol)
    data = b"".join(out)
    if rng.random() < 0.2 and data.endswith(eol):
        data = data[:-len(eol)]  # no final newline
    return data, ("fq" if fastq else "fa")


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--cases", type=int, default=200)
    ap.add_argument("--seed", type=int, default=1)
    ap.add_argument("--keep", action="store_true")
    ap.add_argument("--empty-records", action="store_true", help="also write FASTA records without sequence lines (the known difference)")
    ap.add_argument("--probe", help=argparse.SUPPRESS)
    a = ap.parse_args()
    if a.probe:
        print(json.dumps(btllib_records_here(a.probe)))
        return 0
    from oracle import oracle_py
    if not os.path.exists(oracle_py.REF_PROBE):
        print(f"{oracle_py.REF_PROBE} is missing: make -C oracle ref")
        return 2
    from links_b200.seqio import read_seqfile
    tmp = tempfile.mkdtemp(prefix="reader_fuzz_")
    exe, chk = os.path.join(tmp, "seq_stream_dump"), os.path.join(tmp, "ingest_check")
    subprocess.run(["g++", "-std=c++17", "-O1", os.path.join(ROOT, "tests", "cpp", "seq_stream_dump.cpp"), "-o", exe], check=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "links_b200", "csrc", "host"),
                    os.path.join(ROOT, "tests", "host", "ingest_check.cpp"), "-o", chk], check=True)
    n_bad = n_rec = n_refused = 0
    for i in range(a.cases):
        rng = np.random.default_rng(a.seed * 1_000_003 + i)
        data, ext = make_file(rng, a.empty_records)
        gz = rng.random() < 0.15
        path = os.path.join(tmp, f"case{i}.{ext}" + (".gz" if gz else ""))
        with (gzip.open(path, "wb") if gz else open(path, "wb")) as f:
            f.write(data)
        want = btllib_records(path)
        if want is None:  # the reference cannot read this file either: nothing to be equal to
            n_refused += 1
            os.remove(path)
            continue
        n_rec += len(want)
        why = []
        s = read_seqfile(path)
        if [(len(s.record(j)), fnv1a(s.record(j)), j) for j in range(s.n)] != want:
            why.append("python reader")
        for cap in (1 << 16, 1 << 22):
            r = subprocess.run([exe, str(cap), "13", path], capture_output=True, text=True)
            rows = [ln.split("\t") for ln in r.stdout.split("\n") if ln]
            if r.returncode != 0 or [(int(x[1]), int(x[2], 16), int(x[0])) for x in rows] != want:
                why.append(f"serial reader (staging {cap})")
        if not gz:
            for threads, window in ((3, 977), (8, 1)):
                r = subprocess.run([chk, path, str(threads), str(1 << 20), str(window), str(1 << 20), "64"], capture_output=True, text=True)
                if r.returncode != 0:
                    why.append(f"parallel ingest ({threads} threads): {r.stdout[-200:]}{r.stderr[-200:]}")
        if why:
            n_bad += 1
            print(f"case {i} (seed {a.seed}) {os.path.basename(path)}: {len(want)} records by btllib; differs in {why}", flush=True)
        elif not a.keep:
            os.remove(path)
    print(f"{a.cases} files: {n_bad} differ ({n_rec} records compared; btllib itself refused {n_refused} files)")
    if not (a.keep and n_bad):
        shutil.rmtree(tmp, ignore_errors=True)
    else:
        print(f"kept {tmp}")
    return 1 if n_bad else 0


if __name__ == "__main__":
    sys.exit(main())
