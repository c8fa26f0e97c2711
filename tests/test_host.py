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


def _btllib_records(path: str):
    """(length, FNV-1a, ordinal) of every record as btllib::SeqReader (LONG_MODE) hands it to the reference."""
    import ctypes as C
    from oracle import oracle_py
    if not os.path.exists(oracle_py.REF_PROBE):
        pytest.skip("oracle/_ref/libbtl_probe.so not built (needs /root/reference at build time)")
    lib = C.CDLL(oracle_py.REF_PROBE)
    lib.btl_read_records.restype = C.c_uint64
    lib.btl_read_records.argtypes = [C.c_char_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]
    cap = 4096
    lens, fnv, nums = (np.zeros(cap, np.uint64) for _ in range(3))
    n = lib.btl_read_records(path.encode(), cap, lens.ctypes.data, fnv.ctypes.data, nums.ctypes.data)
    assert n <= cap
    return [(int(lens[i]), int(fnv[i]), int(nums[i])) for i in range(n)]


def test_readers_follow_btllib_on_trailing_whitespace_and_fail_on_bad_archives(tmp_path):
    """btllib rtrim()s the sequence after every line (all trailing whitespace, not just one CR); blanks INSIDE a line stay and
    become invalid bases.  The serial reader, the multi-threaded reader and the Python reader must give the records the real
    btllib::SeqReader gives.  A missing or truncated compressed input is an error, never a shorter input (btllib's DataSource
    aborts; popen() alone would not notice)."""
    import subprocess
    from .conftest import ROOT
    exe = str(tmp_path / "seq_stream_dump")
    subprocess.run(["g++", "-std=c++17", "-O1", "-Wall", os.path.join(ROOT, "tests", "cpp", "seq_stream_dump.cpp"), "-o", exe], check=True)
    chk = str(tmp_path / "ingest_check")
    subprocess.run(["g++", "-O2", "-std=c++17", "-pthread", "-I", os.path.join(ROOT, "links_b200", "csrc", "host"),
                    os.path.join(ROOT, "tests", "host", "ingest_check.cpp"), "-o", chk], check=True)
    rng = np.random.default_rng(21)
    alphabet = np.frombuffer(b"ACGTacgtN", np.uint8)
    tails = [b"", b" ", b"\t", b" \t ", b"\r", b" \r", b"\t\t\r", b"\x0b", b"\x0c "]
    out = []
    for i in range(60):
        out.append(b">r%d desc\n" % i)
        if i == 0:
            # btllib decides the format on the first 16 KB (seq_reader.hpp:168) and REJECTS a file with a blank or tab there
            # (seq_reader_multiline_fasta_module.hpp:64-75, only CR is tolerated): the odd lines start after a clean 20 KB record
            for _ in range(250):
                out.append(alphabet[rng.integers(0, len(alphabet), 80)].tobytes() + b"\r\n")
            continue
        for _ in range(int(rng.integers(1, 5))):
            line = alphabet[rng.integers(0, len(alphabet), int(rng.integers(1, 90)))].tobytes()
            if rng.random() < 0.2:  # a blank in the middle of a line is data
                cut = int(rng.integers(0, len(line) + 1))
                line = line[:cut] + b" " + line[cut:]
            if line[:1] == b" ":
                line = b"A" + line
            out.append(line + tails[int(rng.integers(0, len(tails)))] + b"\n")
    fa = str(tmp_path / "ws.fa")
    open(fa, "wb").write(b"".join(out))
    want = _btllib_records(fa)
    assert len(want) == 60 and [w[2] for w in want] == list(range(60))
    s = read_seqfile(fa)
    assert [(len(s.record(i)), _fnv1a(s.record(i)), i) for i in range(s.n)] == want
    # serial reader, small read buffers do not exist as a knob: the 8 MB buffer sees whole lines here; the staging cap is varied
    for cap in (1 << 15, 1 << 20):
        rows = [ln.split("\t") for ln in subprocess.run([exe, str(cap), "7", fa], check=True, capture_output=True, text=True).stdout.split("\n") if ln]
        assert [(int(r[1]), int(r[2], 16), int(r[0])) for r in rows] == want
    # the multi-threaded reader against the serial one (ingest_check compares record by record)
    for threads, window in ((1, 1 << 30), (3, 977), (8, 1)):
        r = subprocess.run([chk, fa, str(threads), str(1 << 20), str(window), str(1 << 20), "64"], capture_output=True, text=True)
        assert r.returncode == 0, r.stdout + r.stderr
    # compressed inputs: missing file, truncated archive
    gz = str(tmp_path / "ok.fa.gz")
    with gzip.open(gz, "wb") as f:
        f.write(b"".join(out) * 50)
    ok = subprocess.run([exe, str(1 << 20), "100", gz], capture_output=True, text=True)
    assert ok.returncode == 0 and len([ln for ln in ok.stdout.split("\n") if ln]) == 60 * 50
    missing = subprocess.run([exe, str(1 << 20), "100", str(tmp_path / "nope.fa.gz")], capture_output=True, text=True)
    assert missing.returncode != 0 and "cannot open" in missing.stderr
    data = open(gz, "rb").read()
    cut = str(tmp_path / "cut.fa.gz")
    open(cut, "wb").write(data[:len(data) // 2])
    bad = subprocess.run([exe, str(1 << 20), "100", cut], capture_output=True, text=True)
    assert bad.returncode != 0 and "decompressing" in bad.stderr


def test_readers_differential_fuzz_against_btllib():
    """tools/reader_fuzz.py: random FASTA / FASTQ files (wrapped or not, LF / CRLF, no final newline, gzip, lower case, IUPAC,
    records of 1 base to 40 kb) read by the serial reader, the multi-threaded ingest and the Python reader against the real
    btllib::SeqReader.  Needs oracle/_ref/libbtl_probe.so, so it is skipped where that is absent.  400 files (11.9 k records)
    of the same generator were compared when this was added: none differed; the one known difference (a header directly
    followed by a header) is described in the tool's docstring."""
    import subprocess
    import sys
    from oracle import oracle_py
    from .conftest import ROOT
    if not os.path.exists(oracle_py.REF_PROBE):
        pytest.skip("oracle/_ref/libbtl_probe.so not built (needs /root/reference at build time)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "reader_fuzz.py"), "--cases", "10", "--seed", "8"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "10 files: 0 differ" in r.stdout


def test_command_line_differential_fuzz_against_reference_binary():
    """tools/cli_fuzz.py: random option lines (valid and broken -d / -t lists, lost option values, unknown options, omitted
    -b, junk in numeric options) through the unmodified reference binary and through links_b200/bin/LINKS_CPP: the same
    options block, and where the reference stops in its argument handling the same exit code and message.  Needs
    oracle/_ref/LINKS_CPP; skipped where that is absent.  300 lines of the same generator: none differed (a trailing comma
    in a list used to: the reference ends in std::stoul(""), this binary had skipped the empty item)."""
    import subprocess
    import sys
    from .conftest import ROOT
    if not os.path.exists(os.path.join(ROOT, "oracle", "_ref", "LINKS_CPP")):
        pytest.skip("oracle/_ref/LINKS_CPP not built (needs /root/reference)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "cli_fuzz.py"), "--cases", "40", "--seed", "6"], capture_output=True, text=True)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "40 command lines: 0 differ" in r.stdout
