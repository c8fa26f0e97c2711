"""ctypes view of include/links_gpu.h (liblinks_gpu.so).  No compute happens in Python."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("LINKS_GPU_LIB") or os.path.join(HERE, "liblinks_gpu.so")  # override: kernel-variant experiments

LINKS_OK = 0
LINKS_E_INVAL, LINKS_E_CUDA, LINKS_E_NOMEM, LINKS_E_LIMIT, LINKS_E_NODEVICE = -1, -2, -3, -4, -5


class LinksParams(C.Structure):
    _fields_ = [
        ("k", C.c_uint32),
        ("n_dist", C.c_uint32),
        ("dist", C.POINTER(C.c_uint32)),
        ("step", C.POINTER(C.c_uint16)),
        ("offset", C.c_uint64),
        ("insert_stdev", C.c_float),
        ("device", C.c_int32),
    ]


class LinksRow(C.Structure):
    _fields_ = [
        ("bin", C.c_int32),
        ("tig_a", C.c_uint32),
        ("tig_b", C.c_uint32),
        ("rev_a", C.c_uint8),
        ("rev_b", C.c_uint8),
        ("pad_", C.c_uint8 * 2),
        ("links", C.c_uint64),
        ("gaps", C.c_int64),
    ]


class LinksIssue(C.Structure):
    _fields_ = [
        ("tig", C.c_uint32),
        ("o1", C.c_uint8),
        ("o2", C.c_uint8),
        ("pad_", C.c_uint8 * 2),
        ("tig_len", C.c_uint64),
        ("s1", C.c_uint64),
        ("e1", C.c_uint64),
        ("s2", C.c_uint64),
        ("e2", C.c_uint64),
        ("pet_size", C.c_uint64),
        ("insert_size", C.c_int64),
    ]


class LinksStats(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in (
        "n_candidates", "n_survivors", "n_unique_pairs", "n_bf_probes", "n_contig_kmers",
        "ct_ok_pairs", "ct_problem_pairs", "ct_ok_contig", "ct_iz_issues", "ct_illogical", "ct_gate_fail",
        "n_rows", "n_bf_loads", "n_member_sectors")]


class LinksTiming(C.Structure):
    _fields_ = [(n, C.c_float) for n in (
        "ms_h2d", "ms_pack", "ms_bf_build", "ms_table_build", "ms_extract", "ms_classify", "ms_sort_reduce",
        "ms_d2h", "ms_grow", "ms_exchange")] + [("launches", C.c_uint64), ("extract_launches", C.c_uint64)]


# name -> (restype, argtypes); this table is also what tests check against the header
SIGNATURES = {
    "links_create": (C.c_int, [C.POINTER(C.c_void_p), C.POINTER(LinksParams)]),
    "links_destroy": (None, [C.c_void_p]),
    "links_last_error": (C.c_char_p, [C.c_void_p]),
    "links_sync": (C.c_int, [C.c_void_p]),
    "links_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "links_reset": (C.c_int, [C.c_void_p]),
    "links_table_mode": (C.c_int, [C.c_void_p, C.c_int]),
    "links_contigs_push_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "links_reads_push_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64]),
    "links_host_alloc": (C.c_void_p, [C.c_size_t]),
    "links_host_free": (None, [C.c_void_p]),
    "links_bf_size": (C.c_int, [C.c_uint64, C.c_float, C.POINTER(C.c_uint64), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]),
    "links_bf_alloc": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32]),
    "links_bf_load": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]),
    "links_bf_download": (C.c_int, [C.c_void_p, C.c_void_p]),
    "links_bf_covers_contigs": (C.c_int, [C.c_void_p, C.c_int]),
    "links_bf_header": (C.c_size_t, [C.c_char_p, C.c_size_t, C.c_uint32, C.c_uint32, C.c_uint64]),
    "links_contigs_push": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint32]),
    "links_reads_push": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64]),
    "links_finalize": (C.c_int, [C.c_void_p, C.POINTER(C.POINTER(LinksRow)), C.POINTER(C.c_uint64),
                                 C.POINTER(C.POINTER(LinksIssue)), C.POINTER(C.c_uint64), C.POINTER(LinksStats)]),
    "links_free_result": (None, [C.POINTER(LinksRow), C.POINTER(LinksIssue)]),
    "links_format_row": (C.c_int, [C.POINTER(LinksRow), C.c_char_p, C.c_size_t]),
    "links_format_issue": (C.c_int, [C.POINTER(LinksIssue), C.c_char_p, C.c_size_t]),
    "links_pairs_export": (C.c_int, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "links_rows_merge": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.POINTER(LinksRow)), C.POINTER(C.c_uint64)]),
    "links_pair_owner": (C.c_uint32, [C.c_uint64, C.c_uint64, C.c_uint32]),
    "links_bf_download_dev": (C.c_int, [C.c_void_p, C.c_void_p]),
    "links_bf_load_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32]),
    "links_pairs_import": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64]),
    "links_pairs_clear": (C.c_int, [C.c_void_p]),
    "links_exchange_attach": (C.c_int, [C.c_void_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p), C.c_uint64]),
    "links_exchange_finish": (C.c_int, [C.c_void_p]),
    "links_bf_alloc_shared": (C.c_int, [C.c_void_p, C.c_uint64, C.c_uint32, C.c_uint32, C.c_uint32, C.POINTER(C.c_void_p)]),
    "links_bf_allreduce": (C.c_int, [C.c_void_p]),
    "links_kmer_hashes": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "links_kmer_hashes_static": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
    "links_bf_contains": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.c_void_p]),
    "links_pairs_download": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint64, C.POINTER(C.c_uint64)]),
    "links_get_timing": (C.c_int, [C.c_void_p, C.POINTER(LinksTiming), C.c_int]),
}

_lib = None


def load() -> C.CDLL:
    """Load liblinks_gpu.so; raise loudly if it is missing (there is no fallback path)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -m links_b200.build` "
            "(nvcc, sm_100a).  links_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
