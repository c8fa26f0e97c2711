"""ctypes binding of the CPU oracle (oracle/liblinks_oracle.so) and, when present, of the shim
around the reference's own btllib (oracle/_ref/libbtl_probe.so).  TEST INFRASTRUCTURE ONLY:
importable from tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg, never from
links_b200/.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "liblinks_oracle.so")
REF_DIR = os.path.join(HERE, "_ref")
REF_BIN = os.path.join(REF_DIR, "LINKS_CPP")
REF_PROBE = os.path.join(REF_DIR, "libbtl_probe.so")


class LoParams(C.Structure):
    _fields_ = [("k", C.c_uint), ("n_dist", C.c_uint), ("dist", C.POINTER(C.c_uint32)), ("step", C.POINTER(C.c_uint16)),
                ("offset", C.c_uint64), ("insert_stdev", C.c_float)]


ROW_DTYPE = np.dtype([("bin", "<i4"), ("tig_a", "<u4"), ("tig_b", "<u4"), ("rev_a", "u1"), ("rev_b", "u1"),
                      ("pad_", "u1", (2,)), ("links", "<u8"), ("gaps", "<i8")])
ISSUE_DTYPE = np.dtype([("tig", "<u4"), ("pad0_", "u1", (4,)), ("tig_len", "<u8"), ("s1", "<u8"), ("e1", "<u8"),
                        ("s2", "<u8"), ("e2", "<u8"), ("o1", "u1"), ("o2", "u1"), ("pad1_", "u1", (6,)),
                        ("pet_size", "<u8"), ("insert_size", "<i8")])
OBS_DTYPE = np.dtype([("h_a", "<u8"), ("h_b", "<u8"), ("d_idx", "<u4"), ("read_idx", "<u4"), ("rank", "<u8")])
STAT_NAMES = ("n_candidates", "n_survivors", "n_unique_pairs", "n_mates", "n_tracked", "ct_ok_pairs",
              "ct_problem_pairs", "ct_ok_contig", "ct_iz_issues", "ct_illogical", "ct_gate_fail")


class LoStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in STAT_NAMES]


class LoResult(C.Structure):
    _fields_ = [("rows", C.c_void_p), ("n_rows", C.c_uint64), ("issues", C.c_void_p), ("n_issues", C.c_uint64),
                ("obs", C.c_void_p), ("n_obs", C.c_uint64), ("stats", LoStats)]


_lib = None


def build() -> None:
    subprocess.run(["make", "-C", HERE, "oracle"], check=True, stdout=subprocess.DEVNULL)


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB):
            build()
        lib = C.CDLL(LIB)
        lib.lo_run.restype = C.POINTER(LoResult)
        lib.lo_run.argtypes = [C.POINTER(LoParams), C.c_void_p, C.c_uint64, C.c_uint, C.c_void_p, C.c_void_p, C.c_uint64,
                               C.c_void_p, C.c_void_p, C.c_uint64, C.c_int]
        lib.lo_run_from_pairs.restype = C.POINTER(LoResult)
        lib.lo_run_from_pairs.argtypes = [C.POINTER(LoParams), C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p,
                                          C.c_void_p, C.c_uint64]
        lib.lo_free.argtypes = [C.POINTER(LoResult)]
        lib.lo_bf_size.argtypes = [C.c_uint64, C.c_float, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint)]
        lib.lo_bf_insert_seq.argtypes = [C.c_void_p, C.c_uint64, C.c_uint, C.c_uint, C.c_void_p, C.c_uint64]
        lib.lo_bf_contains.restype = C.c_int
        lib.lo_bf_contains.argtypes = [C.c_void_p, C.c_uint64, C.c_uint, C.c_uint, C.c_uint64, C.c_uint64]
        lib.lo_bf_header.restype = C.c_size_t
        lib.lo_bf_header.argtypes = [C.c_char_p, C.c_size_t, C.c_uint, C.c_uint, C.c_uint64]
        lib.lo_nthash_walk.restype = C.c_uint64
        lib.lo_nthash_walk.argtypes = [C.c_void_p, C.c_uint64, C.c_uint, C.c_uint, C.c_uint64, C.c_uint64, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p]
        lib.lo_nthash_kmer.argtypes = [C.c_char_p, C.c_uint, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64)]
        lib.lo_multihash.argtypes = [C.c_uint64, C.c_uint64, C.c_uint, C.c_uint, C.c_void_p]
        lib.lo_format_row.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        lib.lo_format_issue.argtypes = [C.c_void_p, C.c_char_p, C.c_size_t]
        _lib = lib
    return _lib


def bf_size(file_bytes: int, fpr: float) -> tuple[int, int, int]:
    m, b, h = C.c_uint64(), C.c_uint64(), C.c_uint()
    load().lo_bf_size(file_bytes, C.c_float(fpr), C.byref(m), C.byref(b), C.byref(h))
    return m.value, b.value, h.value


def bf_header(k: int, hash_num: int, nbytes: int) -> bytes:
    buf = C.create_string_buffer(512)
    n = load().lo_bf_header(buf, 512, k, hash_num, nbytes)
    return buf.raw[:n]


def bf_build(seq: np.ndarray, off: np.ndarray, nbytes: int, hash_num: int, k: int) -> np.ndarray:
    lib = load()
    bits = np.zeros(nbytes, np.uint8)
    seq = np.ascontiguousarray(seq, np.uint8)
    for i in range(len(off) - 1):
        a, b = int(off[i]), int(off[i + 1])
        lib.lo_bf_insert_seq(bits.ctypes.data, nbytes, hash_num, k, seq.ctypes.data + a, b - a)
    return bits


def bf_contains(bits: np.ndarray, hash_num: int, k: int, fwd: np.ndarray, rev: np.ndarray) -> np.ndarray:
    lib = load()
    return np.array([lib.lo_bf_contains(bits.ctypes.data, bits.size, hash_num, k, int(f), int(r))
                     for f, r in zip(fwd, rev)], dtype=np.uint8)


def nthash_kmer(kmer: bytes) -> tuple[int, int]:
    f, r = C.c_uint64(), C.c_uint64()
    load().lo_nthash_kmer(kmer, len(kmer), C.byref(f), C.byref(r))
    return f.value, r.value


def multihash(fwd: int, rev: int, k: int, m: int) -> np.ndarray:
    out = np.zeros(m, np.uint64)
    load().lo_multihash(fwd, rev, k, m, out.ctypes.data)
    return out


def _walk(fn, seq: bytes, hash_num: int, k: int, start: int):
    cap = max(len(seq), 1)
    pos = np.zeros(cap, np.uint64)
    fwd = np.zeros(cap, np.uint64)
    rev = np.zeros(cap, np.uint64)
    hs = np.zeros(cap * hash_num, np.uint64)
    buf = np.frombuffer(seq, np.uint8) if seq else np.zeros(1, np.uint8)
    n = fn(buf.ctypes.data, len(seq), hash_num, k, start, cap, pos.ctypes.data, fwd.ctypes.data, rev.ctypes.data,
           hs.ctypes.data)
    return pos[:n], fwd[:n], rev[:n], hs[:n * hash_num].reshape(n, hash_num)


def nthash_walk(seq: bytes, hash_num: int, k: int, start: int = 0):
    return _walk(load().lo_nthash_walk, seq, hash_num, k, start)


def run(k: int, dist: list[int], step: list[int], offset: int, insert_stdev: float, bf_bits: np.ndarray, hash_num: int,
        contig_seq: np.ndarray, contig_off: np.ndarray, read_seq: np.ndarray, read_off: np.ndarray, keep_obs: bool = False):
    """Stages 2-4 on the CPU.  Returns (rows, issues, stats dict, obs or None)."""
    lib = load()
    n = len(dist)
    d = (C.c_uint32 * n)(*dist)
    s = (C.c_uint16 * n)(*step)
    p = LoParams(k, n, d, s, offset, insert_stdev)
    cs = np.ascontiguousarray(contig_seq, np.uint8)
    co = np.ascontiguousarray(contig_off, np.uint64)
    rs = np.ascontiguousarray(read_seq, np.uint8)
    ro = np.ascontiguousarray(read_off, np.uint64)
    bits = np.ascontiguousarray(bf_bits, np.uint8)
    res = lib.lo_run(C.byref(p), bits.ctypes.data, bits.size, hash_num, cs.ctypes.data, co.ctypes.data, len(co) - 1,
                     rs.ctypes.data, ro.ctypes.data, len(ro) - 1, int(keep_obs))
    if not res:
        raise MemoryError("lo_run failed")
    try:
        r = res.contents
        rows = np.empty(r.n_rows, ROW_DTYPE)
        if r.n_rows:
            C.memmove(rows.ctypes.data, r.rows, r.n_rows * ROW_DTYPE.itemsize)
        issues = np.empty(r.n_issues, ISSUE_DTYPE)
        if r.n_issues:
            C.memmove(issues.ctypes.data, r.issues, r.n_issues * ISSUE_DTYPE.itemsize)
        obs = None
        if keep_obs:
            obs = np.empty(r.n_obs, OBS_DTYPE)
            if r.n_obs:
                C.memmove(obs.ctypes.data, r.obs, r.n_obs * OBS_DTYPE.itemsize)
        stats = {nm: int(getattr(r.stats, nm)) for nm in STAT_NAMES}
    finally:
        lib.lo_free(res)
    return rows, issues, stats, obs


def run_from_pairs(k: int, dist: list[int], step: list[int], offset: int, insert_stdev: float, contig_seq: np.ndarray,
                   contig_off: np.ndarray, h_a: np.ndarray, h_b: np.ndarray, d_idx: np.ndarray):
    """Stages 3-4 on a given distinct pair set.  Returns (rows, issues, stats dict)."""
    lib = load()
    n = len(dist)
    d = (C.c_uint32 * n)(*dist)
    s = (C.c_uint16 * n)(*step)
    p = LoParams(k, n, d, s, offset, insert_stdev)
    cs = np.ascontiguousarray(contig_seq, np.uint8)
    co = np.ascontiguousarray(contig_off, np.uint64)
    a = np.ascontiguousarray(h_a, np.uint64)
    b = np.ascontiguousarray(h_b, np.uint64)
    di = np.ascontiguousarray(d_idx, np.uint32)
    res = lib.lo_run_from_pairs(C.byref(p), cs.ctypes.data, co.ctypes.data, len(co) - 1, a.ctypes.data, b.ctypes.data,
                                di.ctypes.data, len(a))
    if not res:
        raise MemoryError("lo_run_from_pairs failed")
    try:
        r = res.contents
        rows = np.empty(r.n_rows, ROW_DTYPE)
        if r.n_rows:
            C.memmove(rows.ctypes.data, r.rows, r.n_rows * ROW_DTYPE.itemsize)
        issues = np.empty(r.n_issues, ISSUE_DTYPE)
        if r.n_issues:
            C.memmove(issues.ctypes.data, r.issues, r.n_issues * ISSUE_DTYPE.itemsize)
        stats = {nm: int(getattr(r.stats, nm)) for nm in STAT_NAMES}
    finally:
        lib.lo_free(res)
    return rows, issues, stats


def format_rows(rows: np.ndarray) -> bytes:
    lib = load()
    buf = C.create_string_buffer(256)
    out = []
    for i in range(len(rows)):
        rec = rows[i:i + 1].tobytes()
        n = lib.lo_format_row(rec, buf, 256)
        out.append(buf.raw[:n])
    return b"".join(out)


def format_issues(issues: np.ndarray) -> bytes:
    lib = load()
    buf = C.create_string_buffer(1024)
    out = []
    for i in range(len(issues)):
        rec = issues[i:i + 1].tobytes()
        n = lib.lo_format_issue(rec, buf, 1024)
        out.append(buf.raw[:n])
    return b"".join(out)


# ---- the reference's own btllib, when oracle/_ref was built (this container) -------------------

_probe = None


def have_ref_probe() -> bool:
    return os.path.exists(REF_PROBE)


def have_ref_bin() -> bool:
    return os.path.exists(REF_BIN)


def _load_probe() -> C.CDLL:
    global _probe
    if _probe is None:
        lib = C.CDLL(REF_PROBE)
        lib.btl_nthash_walk.restype = C.c_uint64
        lib.btl_nthash_walk.argtypes = [C.c_void_p, C.c_uint64, C.c_uint, C.c_uint, C.c_uint64, C.c_uint64, C.c_void_p,
                                        C.c_void_p, C.c_void_p, C.c_void_p]
        lib.btl_bf_build.restype = C.c_uint64
        lib.btl_bf_build.argtypes = [C.c_uint64, C.c_uint, C.c_uint, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_char_p]
        _probe = lib
    return _probe


def ref_nthash_walk(seq: bytes, hash_num: int, k: int, start: int = 0):
    return _walk(_load_probe().btl_nthash_walk, seq, hash_num, k, start)


def ref_bf_build(seq: np.ndarray, off: np.ndarray, nbytes: int, hash_num: int, k: int) -> np.ndarray:
    lib = _load_probe()
    rounded = (nbytes + 7) // 8 * 8
    out = np.zeros(rounded, np.uint8)
    seq = np.ascontiguousarray(seq, np.uint8)
    off = np.ascontiguousarray(off, np.uint64)
    lib.btl_bf_build(nbytes, hash_num, k, seq.ctypes.data, off.ctypes.data, len(off) - 1, out.ctypes.data, None)
    return out
