"""links_b200 -- B200-native hot path of bcgsc/LINKS (k-mer-pair extract -> Bloom probe ->
contig k-mer map -> link aggregation) behind the C-ABI in include/links_gpu.h.

This package is the Python host-side mirror of the reference's `class LINKS` stage interface
(reference src/LINKS.hpp:32-36); all compute happens in liblinks_gpu.so (hand-written sm_100a
CUDA).  There is no CPU or PyTorch fallback: importing works anywhere, but creating a `Links`
object without the built library or without a CUDA device raises.
"""
from __future__ import annotations

from .pairing import Links, LinksError, LinksOptions, bf_size, bf_header  # noqa: F401
from .seqio import SeqSet, read_fof, read_seqfile, write_fasta  # noqa: F401

__all__ = ["Links", "LinksError", "LinksOptions", "SeqSet", "bf_size", "bf_header", "read_fof", "read_seqfile",
           "write_fasta"]
