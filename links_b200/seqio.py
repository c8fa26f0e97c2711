"""Sequence file reading for the Python host mirror: FASTA / FASTQ (optionally .gz) into one
contiguous ASCII buffer plus record offsets -- the layout links_contigs_push / links_reads_push
take.  Mirrors what btllib::SeqReader hands LINKS in LONG_MODE: record order, no case folding,
`num` = 0-based ordinal (reference src/LINKS.hpp:404-468).
"""
from __future__ import annotations

import gzip
from dataclasses import dataclass

import numpy as np


@dataclass
class SeqSet:
    seq: np.ndarray   # uint8, concatenated bases
    off: np.ndarray   # uint64, n+1 offsets
    names: list[str]

    @property
    def n(self) -> int:
        return len(self.off) - 1

    def record(self, i: int) -> bytes:
        return self.seq[int(self.off[i]):int(self.off[i + 1])].tobytes()

    @staticmethod
    def from_records(records: list[bytes], names: list[str] | None = None) -> "SeqSet":
        off = np.zeros(len(records) + 1, dtype=np.uint64)
        if records:
            off[1:] = np.cumsum([len(r) for r in records], dtype=np.uint64)
        seq = np.frombuffer(b"".join(records), dtype=np.uint8).copy() if records else np.zeros(0, np.uint8)
        return SeqSet(seq, off, names or [str(i + 1) for i in range(len(records))])

    @staticmethod
    def concat(sets: list["SeqSet"]) -> "SeqSet":
        if not sets:
            return SeqSet.from_records([])
        seq = np.concatenate([s.seq for s in sets]) if sets else np.zeros(0, np.uint8)
        offs = [np.zeros(1, np.uint64)]
        base = np.uint64(0)
        names: list[str] = []
        for s in sets:
            offs.append(s.off[1:] + base)
            base = base + s.off[-1]
            names += s.names
        return SeqSet(seq, np.concatenate(offs).astype(np.uint64), names)


def _open(path: str):
    if path.endswith(".gz"):
        return gzip.open(path, "rb")
    return open(path, "rb")


def read_seqfile(path: str) -> SeqSet:
    """FASTA (single or multi-line) or FASTQ (4-line records)."""
    records: list[bytes] = []
    names: list[str] = []
    with _open(path) as f:
        data = f.read()
    if not data:
        return SeqSet.from_records([])
    lines = data.split(b"\n")
    first = next((ln for ln in lines if ln.strip()), b"")
    if first.startswith(b"@"):
        i = 0
        n = len(lines)
        while i < n:
            if not lines[i].strip():
                i += 1
                continue
            names.append(lines[i][1:].split()[0].decode() if len(lines[i]) > 1 else "")
            records.append(lines[i + 1].rstrip() if i + 1 < n else b"")
            i += 4
    else:
        cur: list[bytes] | None = None
        for ln in lines:
            ln = ln.rstrip()
            if ln.startswith(b">"):
                if cur is not None:
                    records.append(b"".join(cur))
                names.append(ln[1:].split()[0].decode() if len(ln) > 1 else "")
                cur = []
            elif cur is not None and ln:
                cur.append(ln)
        if cur is not None:
            records.append(b"".join(cur))
    return SeqSet.from_records(records, names)


def read_fof(fof_path: str) -> list[str]:
    """File-of-filenames: whitespace separated paths (reference src/InputParser.hpp:54-64)."""
    with open(fof_path) as f:
        return f.read().split()


def write_fasta(path: str, s: SeqSet) -> None:
    with open(path, "wb") as f:
        for i in range(s.n):
            f.write(b">" + s.names[i].encode() + b"\n")
            f.write(s.record(i))
            f.write(b"\n")
