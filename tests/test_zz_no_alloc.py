"""links_kmer_hashes_static: the allocation-free variant of links_kmer_hashes (fixed scratch in a static __device__ array).

No counterpart in the reference.  Collected last on purpose:
the entry point is called through the C-ABI and compared, bit for bit, with links_kmer_hashes on the
same inputs -- twice in a row and with a shorter sequence after a longer one, since every call shares the array and finds
the previous call's contents in it."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu

MAX_BASES = 1 << 17  # LINKS_STATIC_MAX_BASES, include/links_gpu.h


def _links(k):
    from links_b200 import Links, LinksOptions
    return Links(LinksOptions(k=k, distances=[4000], step_sizes=[2]))


def _dirty(rng, n, alphabet=b"ACGTACGTACGTACGTACGTACGTacgtNnRYUu-."):
    a = np.frombuffer(alphabet, np.uint8)
    return a[rng.integers(0, len(a), n)].tobytes()


def _same(a, b):
    return all(np.array_equal(x, y) for x, y in zip(a, b))


@pytest.mark.parametrize("k", [15, 20, 33])
def test_static_variant_matches_allocating_entry_point(gpu_lib, k):
    rng = np.random.default_rng(100 + k)
    # long first, then short ones that leave the tail of the array holding the long call's bytes; MAX_BASES is the limit itself
    seqs = [_dirty(rng, MAX_BASES), _dirty(rng, 5000), _dirty(rng, k), _dirty(rng, k - 1), _dirty(rng, 70001, b"ACGT"), _dirty(rng, 257)]
    with _links(k) as L:
        for seq in seqs:
            want = L.kmer_hashes(seq)
            first = L.kmer_hashes_static(seq)
            second = L.kmer_hashes_static(seq)  # two calls in a row: the second finds the first one's scratch
            assert _same(first, want) and _same(second, want), (k, len(seq))
            assert len(want[0]) == max(len(seq) - k + 1, 0)


def test_static_variant_across_contexts(gpu_lib):
    # the array belongs to the library (one per GPU), not to a context: two contexts with different k take turns on it
    rng = np.random.default_rng(7)
    seq = _dirty(rng, 30000)
    with _links(17) as A, _links(40) as B:
        for _ in range(2):
            assert _same(A.kmer_hashes_static(seq), A.kmer_hashes(seq))
            assert _same(B.kmer_hashes_static(seq), B.kmer_hashes(seq))


def test_static_variant_rejects_sizes_beyond_the_array(gpu_lib):
    from links_b200 import LinksError
    seq = b"A" * (MAX_BASES + 1)
    with _links(20) as L:
        with pytest.raises(LinksError) as ei:
            L.kmer_hashes_static(seq)
        assert "LINKS_STATIC_MAX_BASES" in str(ei.value)
        # and the context is still usable
        assert _same(L.kmer_hashes_static(seq[:1000]), L.kmer_hashes(seq[:1000]))


def test_static_variant_known_answer(gpu_lib):
    # reference btllib/tests/nthash.cpp:10-28, through the new entry point
    with _links(20) as L:
        fwd, rev, valid = L.kmer_hashes_static(b"ACGTACACTGGACTGAGTCT")
    assert valid.tolist() == [1]
    assert min(int(fwd[0]), int(rev[0])) == 10434435546371013747
