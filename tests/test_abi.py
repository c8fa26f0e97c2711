"""The C-ABI library loads on a CPU-only box and exports every symbol include/links_gpu.h declares.
No compute is attempted here; links_create must refuse loudly without a device (no CPU fallback)."""
from __future__ import annotations

import ctypes as C
import os
import re

import numpy as np
import pytest

from .conftest import ROOT, GoldenCase, have_gpu


def _declared_symbols() -> set[str]:
    hdr = open(os.path.join(ROOT, "include", "links_gpu.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    return set(re.findall(r"\b(links_[a-z0-9_]+)\s*\(", hdr))


def test_header_symbols_are_exported(gpu_lib):
    from links_b200 import _abi
    declared = _declared_symbols()
    assert declared, "no declarations parsed"
    for name in sorted(declared):
        assert hasattr(gpu_lib, name), f"{name} declared in include/links_gpu.h but not exported"
    assert declared == set(_abi.SIGNATURES), "ctypes signature table and header disagree"


def test_no_oracle_in_product():
    """The product path must not import, link or execute anything under oracle/."""
    pkg = os.path.join(ROOT, "links_b200")
    for d, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".hpp", ".h")):
                src = open(os.path.join(d, f), errors="ignore").read()
                assert "oracle" not in src.lower() or f == "build.py", f"{f} mentions the oracle"


def test_create_fails_loudly_without_device(gpu_lib):
    if have_gpu():
        pytest.skip("a CUDA device is present")
    from links_b200 import Links, LinksError, LinksOptions
    with pytest.raises(LinksError) as e:
        Links(LinksOptions())
    assert e.value.code == -5 and "no CPU path" in str(e.value)


def test_host_helpers_bf_sizing_and_header(gpu_lib):
    """links_bf_size / links_bf_header are pure host arithmetic: check them against what the reference
    binary printed and wrote (tests/golden/*/case.json)."""
    from links_b200 import bf_header, bf_size
    for name in ("c1", "adv_ns_lower", "adv_offset_k20", "adv_tandem", "adv_k40"):
        c = GoldenCase(name)
        m, nbytes, h = bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
        assert (m, h) == (c.meta["bf_m_bits"], c.meta["bf_hash_num"])
        hdr = bf_header(c.opts["k"], h, nbytes)
        assert len(hdr) + nbytes == c.meta["bloom_file_bytes"]
        assert hdr.decode().startswith(c.meta["bloom_header"])
    m, nbytes, h = bf_size(721525, 0.001)
    assert (m, nbytes, h) == (10373792, 1296728, 9)  # SURVEY 3.3, reference stdout


def test_bf_sizing_matches_oracle_widely(gpu_lib, oracle):
    from links_b200 import bf_size
    rng = np.random.default_rng(0)
    for _ in range(300):
        n = int(rng.integers(1, 4_000_000_000))
        p = float(rng.choice([0.5, 0.1, 0.01, 0.001, 0.0001, 1e-6]))
        assert bf_size(n, p) == oracle.bf_size(n, p)


def test_row_and_issue_formatting(gpu_lib):
    from links_b200 import _abi
    r = _abi.LinksRow(bin=-1, tig_a=35, tig_b=0, rev_a=1, rev_b=0, links=5, gaps=-14472)
    buf = C.create_string_buffer(128)
    n = gpu_lib.links_format_row(C.byref(r), buf, 128)
    assert buf.raw[:n] == b"-1\tr35\tf\t5\t-14472\n"  # empty tig id, as in the golden tsv
    i = _abi.LinksIssue(tig=18, o1=0, o2=0, tig_len=34323, s1=17092, e1=17107, s2=9207, e2=9222, pet_size=7885,
                        insert_size=20000)
    buf = C.create_string_buffer(512)
    n = gpu_lib.links_format_issue(C.byref(i), buf, 512)
    assert buf.raw[:n] == (b"Pairs unsatisfied in distance within a contig.  Pairs on contig id: 18 (length: 34323)\n"
                           b"kmer start: 17092 kmer end: 17107 orientation: 0\n"
                           b"kmer start: 9207 kmer end: 9222 orientation: 0\n"
                           b"Calculated distance apart on contig: 7885 - Distance on read: 20000\n")


def test_static_entry_point_allocates_nothing_and_states_its_limit():
    """links_kmer_hashes_static (no counterpart in the reference): its body calls no allocating routine, its scratch is a
    file-scope static __device__ array whose address comes from cudaGetSymbolAddress, and header, source and test agree on
    the size limit.  Source-level checks; the behaviour is tested on the GPU in tests/test_zz_no_alloc.py."""
    src = open(os.path.join(ROOT, "links_b200", "csrc", "links_gpu.cu")).read()
    hdr = open(os.path.join(ROOT, "include", "links_gpu.h")).read()
    start = src.index("int links_kmer_hashes_static(")
    body = src[start:src.index("\n}\n", start)]
    for banned in ("cudaMalloc", "cudaFree", "cudaHostAlloc", "cudaMallocHost", ".ensure(", "DevBuf", "SeqBatchDev", "std::vector", "new ", "malloc("):
        assert banned not in body, f"links_kmer_hashes_static uses {banned}"
    assert "static_scratch_base(" in body and "kStaticMaxBases" in body and "LINKS_E_LIMIT" in body
    assert re.search(r"static\s+__device__\s+__align__\(256\)\s+unsigned char\s+g_static_scratch\[", src)
    helper = src[src.index("int static_scratch_base("):src.index("} // namespace", src.index("int static_scratch_base("))]
    assert "cudaGetSymbolAddress(&p, g_static_scratch)" in helper and "CU(c," in helper
    m = re.search(r"#define\s+LINKS_STATIC_MAX_BASES\s+\(1u\s*<<\s*(\d+)\)", hdr)
    assert m, "LINKS_STATIC_MAX_BASES not found in include/links_gpu.h"
    from . import test_zz_no_alloc
    assert test_zz_no_alloc.MAX_BASES == 1 << int(m.group(1))
