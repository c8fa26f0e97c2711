"""Python host mirror of the reference's `class LINKS` (src/LINKS.hpp:30-282) over the C-ABI.

Stage methods keep the reference's names and order (src/main.cpp:28-59):

    links = Links(LinksOptions(assembly_file=..., fof_file=..., distances=[4000], ...))
    links.init_bloom_filter()   # sizes + builds (or loads, -r) the Bloom filter, writes <b>.bloom
    links.start_read_fasta()    # reads every file of the .fof -> k-mer pairs -> Bloom probe -> pair set
    links.start_read_contig()   # contig k-mer table
    links.pair_contigs()        # classify + aggregate, writes <b>.tigpair_checkpoint.tsv / .pairing_issues

plus in-memory entry points (push_contigs / push_reads / finalize) used by tests and bench.py.
Every byte of arithmetic happens in liblinks_gpu.so on the GPU.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass, field

import numpy as np

from . import _abi
from .seqio import SeqSet, read_fof, read_seqfile

ROW_DTYPE = np.dtype([("bin", "<i4"), ("tig_a", "<u4"), ("tig_b", "<u4"), ("rev_a", "u1"), ("rev_b", "u1"),
                      ("pad_", "u1", (2,)), ("links", "<u8"), ("gaps", "<i8")])
ISSUE_DTYPE = np.dtype([("tig", "<u4"), ("o1", "u1"), ("o2", "u1"), ("pad_", "u1", (2,)), ("tig_len", "<u8"),
                        ("s1", "<u8"), ("e1", "<u8"), ("s2", "<u8"), ("e2", "<u8"), ("pet_size", "<u8"),
                        ("insert_size", "<i8")])
assert ROW_DTYPE.itemsize == C.sizeof(_abi.LinksRow)
assert ISSUE_DTYPE.itemsize == C.sizeof(_abi.LinksIssue)


class LinksError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"links_gpu error {code}: {msg}")
        self.code = code


@dataclass
class LinksOptions:
    """The InputParser fields LINKS reads (reference src/InputParser.hpp:68-91), same defaults."""
    assembly_file: str = ""
    fof_file: str = ""
    distances: list[int] = field(default_factory=lambda: [4000])
    step_sizes: list[int] = field(default_factory=lambda: [2])
    k: int = 15
    verbose: bool = False
    min_size: int = 500
    insert_stdev: float = 0.1
    base_name: str = ""
    offset: int = 0
    bf_file: str = ""
    fpr: float = 0.001
    bf_off: bool = False      # parsed but never consulted by the reference (src/LINKS.hpp has no use of it)
    thread: int = 3           # -j: CPU worker count in the reference; unused here
    device: int = 0

    def padded_steps(self) -> list[int]:
        # src/InputParser.hpp:224-229: pad -t with its last value up to len(-d)
        s = list(self.step_sizes)
        while len(s) < len(self.distances):
            s.append(s[-1])
        if len(s) > len(self.distances):
            raise ValueError("Number of provided distances can't be lower than number of step sizes provided.")
        return s


def bf_size(file_bytes: int, fpr: float) -> tuple[int, int, int]:
    """(m_bits, bytes, hash_num) exactly as LINKS::make_bf sizes the filter (src/LINKS.hpp:297-302)."""
    lib = _abi.load()
    m, b, h = C.c_uint64(), C.c_uint64(), C.c_uint32()
    rc = lib.links_bf_size(int(file_bytes), C.c_float(fpr), C.byref(m), C.byref(b), C.byref(h))
    if rc:
        raise LinksError(rc, "links_bf_size")
    return m.value, b.value, h.value


def bf_header(k: int, hash_num: int, nbytes: int) -> bytes:
    lib = _abi.load()
    buf = C.create_string_buffer(512)
    n = lib.links_bf_header(buf, 512, k, hash_num, nbytes)
    return buf.raw[:n]


def parse_bloom_file(path: str) -> tuple[int, int, int, np.ndarray]:
    """Read a BTLKmerBloomFilter_v6 file (btllib bloom_filter.hpp:697-740): (k, hash_num, bytes, bits)."""
    with open(path, "rb") as f:
        data = f.read()
    end = data.index(b"[HeaderEnd]\n") + len(b"[HeaderEnd]\n")
    hdr = data[:end].decode()
    if not hdr.startswith("[BTLKmerBloomFilter_v6]"):
        raise ValueError("Magic string does not match (likely version mismatch)")
    kv = {}
    for ln in hdr.splitlines():
        if "=" in ln:
            a, b = ln.split("=", 1)
            kv[a.strip()] = b.strip().strip('"')
    pos = end
    for _ in range(50):  # PLACEHOLDER_NEWLINES
        pos = data.index(b"\n", pos) + 1
    nbytes = int(kv["bytes"])
    bits = np.frombuffer(data, dtype=np.uint8, count=nbytes, offset=pos).copy()
    return int(kv["k"]), int(kv["hash_num"]), nbytes, bits


class Links:
    def __init__(self, opts: LinksOptions):
        self.opts = opts
        self._lib = _abi.load()
        steps = opts.padded_steps()
        n = len(opts.distances)
        self._dist = (C.c_uint32 * n)(*opts.distances)
        self._step = (C.c_uint16 * n)(*steps)
        prm = _abi.LinksParams(k=opts.k, n_dist=n, dist=self._dist, step=self._step, offset=opts.offset,
                               insert_stdev=opts.insert_stdev, device=opts.device)
        h = C.c_void_p()
        rc = self._lib.links_create(C.byref(h), C.byref(prm))
        if rc:
            raise LinksError(rc, (self._lib.links_last_error(None) or b"").decode())
        self._h = h
        self.hash_num = 0
        self.bf_bytes = 0
        self._n_contigs = 0
        self._n_reads = 0
        self._contigs_pushed = False
        self.stats: dict[str, int] = {}

    # ------------------------------------------------------------------ plumbing
    def close(self) -> None:
        if getattr(self, "_h", None):
            self._lib.links_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def _check(self, rc: int) -> None:
        if rc:
            raise LinksError(rc, (self._lib.links_last_error(self._h) or b"").decode())

    # ------------------------------------------------------------------ in-memory API
    def bf_alloc(self, nbytes: int, hash_num: int) -> None:
        self._check(self._lib.links_bf_alloc(self._h, nbytes, hash_num))
        self.bf_bytes, self.hash_num = nbytes, hash_num

    def bf_load(self, bits: np.ndarray, hash_num: int) -> None:
        bits = np.ascontiguousarray(bits, dtype=np.uint8)
        self._check(self._lib.links_bf_load(self._h, bits.ctypes.data, bits.size, hash_num))
        self.bf_bytes, self.hash_num = bits.size, hash_num

    def bf_bits(self) -> np.ndarray:
        out = np.empty(self.bf_bytes, dtype=np.uint8)
        self._check(self._lib.links_bf_download(self._h, out.ctypes.data))
        return out

    def push_contigs(self, s: SeqSet, first_tig_id: int | None = None) -> None:
        first = self._n_contigs + 1 if first_tig_id is None else first_tig_id
        seq = np.ascontiguousarray(s.seq, dtype=np.uint8)
        off = np.ascontiguousarray(s.off, dtype=np.uint64)
        self._check(self._lib.links_contigs_push(self._h, seq.ctypes.data, off.ctypes.data, s.n, first))
        self._n_contigs = max(self._n_contigs, first + s.n - 1)

    def push_reads(self, s: SeqSet, first_read_idx: int | None = None) -> None:
        first = self._n_reads if first_read_idx is None else first_read_idx
        seq = np.ascontiguousarray(s.seq, dtype=np.uint8)
        off = np.ascontiguousarray(s.off, dtype=np.uint64)
        self._check(self._lib.links_reads_push(self._h, seq.ctypes.data, off.ctypes.data, s.n, first))
        self._n_reads = max(self._n_reads, first + s.n)

    def push_reads_raw(self, seq_ptr: int, off_ptr: int, n: int, first_read_idx: int) -> None:
        """Raw-pointer variant (pinned buffers from links_host_alloc) used by bench.py."""
        self._check(self._lib.links_reads_push(self._h, seq_ptr, off_ptr, n, first_read_idx))

    def push_contigs_raw(self, seq_ptr: int, off_ptr: int, n: int, first_tig_id: int) -> None:
        self._check(self._lib.links_contigs_push(self._h, seq_ptr, off_ptr, n, first_tig_id))
        self._n_contigs = max(self._n_contigs, first_tig_id + n - 1)

    def push_reads_dev(self, dev_ptr: int, off: np.ndarray, first_read_idx: int = 0) -> None:
        """Reads whose concatenated ASCII already sits in device memory (bench.py's HBM-resident leg)."""
        off = np.ascontiguousarray(off, dtype=np.uint64)
        self._check(self._lib.links_reads_push_dev(self._h, dev_ptr, off.ctypes.data, len(off) - 1, first_read_idx))

    def push_contigs_dev(self, dev_ptr: int, off: np.ndarray, first_tig_id: int = 1) -> None:
        off = np.ascontiguousarray(off, dtype=np.uint64)
        self._check(self._lib.links_contigs_push_dev(self._h, dev_ptr, off.ctypes.data, len(off) - 1, first_tig_id))
        self._n_contigs = max(self._n_contigs, first_tig_id + len(off) - 2)

    # ------------------------------------------------------------------ multi-GPU hooks (links_b200/multi.py drives them)
    def n_unique(self) -> int:
        n = C.c_uint64()
        self._check(self._lib.links_pairs_download(self._h, None, None, None, 0, C.byref(n)))
        return int(n.value)

    def pairs_export(self, n_ranks: int, dev_dst: int | None, cap_records: int = 0) -> list[int]:
        """Distinct pairs as {h_a,h_b,order_key} records grouped by owner rank; returns counts per rank."""
        counts = (C.c_uint64 * n_ranks)()
        self._check(self._lib.links_pairs_export(self._h, n_ranks, dev_dst, cap_records, counts))
        return [int(v) for v in counts]

    def pairs_clear(self) -> None:
        self._check(self._lib.links_pairs_clear(self._h))

    def pairs_import(self, dev_records: int, n_records: int) -> None:
        self._check(self._lib.links_pairs_import(self._h, dev_records, n_records))

    def rows_merge(self, rows: np.ndarray) -> np.ndarray:
        """Sum rows with equal key on the device (partial link tables of several ranks)."""
        rows = np.ascontiguousarray(rows, dtype=ROW_DTYPE)
        out_p = C.POINTER(_abi.LinksRow)()
        n_out = C.c_uint64()
        self._check(self._lib.links_rows_merge(self._h, rows.ctypes.data, len(rows), C.byref(out_p), C.byref(n_out)))
        try:
            out = np.empty(n_out.value, dtype=ROW_DTYPE)
            if n_out.value:
                C.memmove(out.ctypes.data, out_p, n_out.value * ROW_DTYPE.itemsize)
        finally:
            self._lib.links_free_result(out_p, None)
        return out

    def exchange_attach(self, rank: int, n_ranks: int, inbox_ptrs: list[int], inbox_bytes: int) -> None:
        """Peer-memory pair exchange (links_gpu.h): inbox_ptrs[r] = rank r's inbox as addressable here."""
        arr = (C.c_void_p * n_ranks)(*inbox_ptrs)
        self._check(self._lib.links_exchange_attach(self._h, rank, n_ranks, arr, inbox_bytes))

    def bf_alloc_shared(self, nbytes: int, hash_num: int, rank: int, n_ranks: int, bloom_ptrs: list[int]) -> None:
        """Bloom filter built in shares and OR-ed over peer memory (links_gpu.h links_bf_alloc_shared)."""
        arr = (C.c_void_p * n_ranks)(*bloom_ptrs)
        self._check(self._lib.links_bf_alloc_shared(self._h, nbytes, hash_num, rank, n_ranks, arr))
        self.bf_bytes, self.hash_num = nbytes, hash_num

    def bf_allreduce(self) -> None:
        self._check(self._lib.links_bf_allreduce(self._h))

    def exchange_finish(self) -> None:
        self._check(self._lib.links_exchange_finish(self._h))

    def bf_download_dev(self, dev_dst: int) -> None:
        self._check(self._lib.links_bf_download_dev(self._h, dev_dst))

    def bf_load_dev(self, dev_bits: int, nbytes: int, hash_num: int) -> None:
        self._check(self._lib.links_bf_load_dev(self._h, dev_bits, nbytes, hash_num))
        self.bf_bytes, self.hash_num = nbytes, hash_num

    def bf_covers_contigs(self, yes: bool = True) -> None:
        """Vouch that a loaded filter was built from the contigs about to be pushed (links_gpu.h)."""
        self._check(self._lib.links_bf_covers_contigs(self._h, 1 if yes else 0))

    TABLE_MODES = {"auto": 0, "full": 1, "mates": 2}

    def table_mode(self, mode: str) -> None:
        """Contig k-mer table build (links_gpu.h LINKS_TABLE_*): 'full', 'mates' (only hashes that occur in a pair,
        reference src/LINKS.hpp:580,591) or 'auto'."""
        self._check(self._lib.links_table_mode(self._h, self.TABLE_MODES[mode]))

    def set_stream(self, cuda_stream: int) -> None:
        self._check(self._lib.links_set_stream(self._h, cuda_stream))

    def reset(self) -> None:
        self._check(self._lib.links_reset(self._h))
        self._n_contigs = 0
        self._n_reads = 0
        self._contigs_pushed = False

    def sync(self) -> None:
        self._check(self._lib.links_sync(self._h))

    def finalize(self) -> tuple[np.ndarray, np.ndarray, dict[str, int]]:
        rows_p = C.POINTER(_abi.LinksRow)()
        iss_p = C.POINTER(_abi.LinksIssue)()
        n_rows, n_iss = C.c_uint64(), C.c_uint64()
        st = _abi.LinksStats()
        self._check(self._lib.links_finalize(self._h, C.byref(rows_p), C.byref(n_rows), C.byref(iss_p), C.byref(n_iss),
                                             C.byref(st)))
        try:
            # Two row buffers kept by this object and used alternately (fresh megabyte arrays cost ~0.2 ms of page
            # faults per job): the array returned here stays intact through the NEXT finalize and is overwritten by
            # the one after -- copy it to keep it longer.
            self._rows_flip = 1 - getattr(self, "_rows_flip", 1)
            bufs = self.__dict__.setdefault("_rows_bufs", [np.empty(0, dtype=ROW_DTYPE), np.empty(0, dtype=ROW_DTYPE)])
            if len(bufs[self._rows_flip]) < n_rows.value:
                bufs[self._rows_flip] = np.empty(n_rows.value + n_rows.value // 4 + 64, dtype=ROW_DTYPE)
            rows = bufs[self._rows_flip][:n_rows.value]
            if n_rows.value:
                C.memmove(rows.ctypes.data, rows_p, n_rows.value * ROW_DTYPE.itemsize)
            issues = np.empty(n_iss.value, dtype=ISSUE_DTYPE)
            if n_iss.value:
                C.memmove(issues.ctypes.data, iss_p, n_iss.value * ISSUE_DTYPE.itemsize)
        finally:
            self._lib.links_free_result(rows_p, iss_p)
        self.stats = {n: int(getattr(st, n)) for n, _ in _abi.LinksStats._fields_}
        return rows, issues, self.stats

    def kmer_hashes(self, seq: bytes) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
        n = max(len(seq) - self.opts.k + 1, 0)
        fwd = np.zeros(n, np.uint64)
        rev = np.zeros(n, np.uint64)
        valid = np.zeros(n, np.uint8)
        buf = np.frombuffer(seq, dtype=np.uint8)
        if n:
            self._check(self._lib.links_kmer_hashes(self._h, buf.ctypes.data, len(seq), fwd.ctypes.data, rev.ctypes.data,
                                                    valid.ctypes.data))
        return fwd, rev, valid

    def kmer_hashes_static(self, seq: bytes) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
        """kmer_hashes through links_kmer_hashes_static: no device allocation, at most LINKS_STATIC_MAX_BASES (2^17) bases."""
        n = max(len(seq) - self.opts.k + 1, 0)
        fwd = np.zeros(n, np.uint64)
        rev = np.zeros(n, np.uint64)
        valid = np.zeros(n, np.uint8)
        buf = np.frombuffer(seq, dtype=np.uint8)
        self._check(self._lib.links_kmer_hashes_static(self._h, buf.ctypes.data, len(seq), fwd.ctypes.data, rev.ctypes.data,
                                                       valid.ctypes.data))
        return fwd, rev, valid

    def bf_contains(self, fwd: np.ndarray, rev: np.ndarray) -> np.ndarray:
        fwd = np.ascontiguousarray(fwd, np.uint64)
        rev = np.ascontiguousarray(rev, np.uint64)
        out = np.zeros(fwd.size, np.uint8)
        self._check(self._lib.links_bf_contains(self._h, fwd.ctypes.data, rev.ctypes.data, fwd.size, out.ctypes.data))
        return out

    def pairs(self) -> tuple[np.ndarray, np.ndarray, np.ndarray]:
        n = C.c_uint64()
        self._check(self._lib.links_pairs_download(self._h, None, None, None, 0, C.byref(n)))
        a = np.zeros(n.value, np.uint64)
        b = np.zeros(n.value, np.uint64)
        d = np.zeros(n.value, np.uint32)
        if n.value:
            self._check(self._lib.links_pairs_download(self._h, a.ctypes.data, b.ctypes.data, d.ctypes.data, n.value,
                                                       C.byref(n)))
        return a, b, d

    def timing(self, reset: bool = False) -> dict[str, float]:
        t = _abi.LinksTiming()
        self._check(self._lib.links_get_timing(self._h, C.byref(t), int(reset)))
        return {n: float(getattr(t, n)) for n, _ in _abi.LinksTiming._fields_ if n != "pad_"}

    def format_rows(self, rows: np.ndarray) -> bytes:
        buf = C.create_string_buffer(256)
        out = []
        arr = (_abi.LinksRow * max(len(rows), 1)).from_buffer_copy(rows.tobytes() if len(rows) else bytes(ROW_DTYPE.itemsize))
        for i in range(len(rows)):
            n = self._lib.links_format_row(C.byref(arr[i]), buf, 256)
            out.append(buf.raw[:n])
        return b"".join(out)

    def format_issues(self, issues: np.ndarray) -> bytes:
        buf = C.create_string_buffer(1024)
        out = []
        arr = (_abi.LinksIssue * max(len(issues), 1)).from_buffer_copy(
            issues.tobytes() if len(issues) else bytes(ISSUE_DTYPE.itemsize))
        for i in range(len(issues)):
            n = self._lib.links_format_issue(C.byref(arr[i]), buf, 1024)
            out.append(buf.raw[:n])
        return b"".join(out)

    # ------------------------------------------------------------------ the reference's four stages
    def init_bloom_filter(self) -> None:
        """LINKS::init_bloom_filter / make_bf (src/LINKS.hpp:284-352)."""
        o = self.opts
        contigs = read_seqfile(o.assembly_file)
        if o.bf_file:
            if not os.path.exists(o.bf_file):
                raise FileNotFoundError(f"Invalid file: {o.bf_file} -- fatal")
            k, hash_num, _nbytes, bits = parse_bloom_file(o.bf_file)
            self.bf_load(bits, hash_num)
        else:
            n = os.stat(o.assembly_file).st_size  # file BYTES, not k-mers (src/LINKS.hpp:349,684-692)
            _m, nbytes, hash_num = bf_size(n, o.fpr)
            self.bf_alloc(nbytes, hash_num)
        # ALL records are inserted, no -z filter (src/LINKS.hpp:318-324); they also stay resident for stage 3
        self.push_contigs(contigs, 1)
        self._contigs_pushed = True
        if not o.bf_file:
            with open(o.base_name + ".bloom", "wb") as f:
                f.write(bf_header(o.k, self.hash_num, self.bf_bytes))
                f.write(self.bf_bits().tobytes())

    def start_read_fasta(self) -> None:
        """LINKS::start_read_fasta (src/LINKS.hpp:470-486): every file of the .fof, in order."""
        for path in read_fof(self.opts.fof_file):
            self.push_reads(read_seqfile(path))

    def start_read_contig(self) -> None:
        """LINKS::start_read_contig (src/LINKS.hpp:487-513).  The contigs were uploaded (2-bit packed)
        by init_bloom_filter and are still resident; the k-mer table itself is built by the first
        kernel of pair_contigs, so this stage has nothing left to read."""
        if not self._contigs_pushed:
            self.push_contigs(read_seqfile(self.opts.assembly_file), 1)
            self._contigs_pushed = True

    def pair_contigs(self) -> tuple[np.ndarray, np.ndarray, dict[str, int]]:
        """LINKS::pair_contigs (src/LINKS.hpp:780-1041): writes .pairing_issues and .tigpair_checkpoint.tsv."""
        rows, issues, stats = self.finalize()
        b = self.opts.base_name
        with open(b + ".pairing_issues", "wb") as f:
            f.write(self.format_issues(issues))
        with open(b + ".tigpair_checkpoint.tsv", "wb") as f:
            f.write(self.format_rows(rows))
        return rows, issues, stats
