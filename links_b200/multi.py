"""Multi-GPU host logic: reads sharded over ranks, Bloom filter replicated by a
broadcast, distinct k-mer pairs routed to one owner rank by an all-to-all keyed on a hash of
(h_a, h_b), per-rank finalize, partial link rows summed on rank 0.

The reference has no counterpart (one address space, src/LINKS.hpp:354-378); what must be
preserved is mate_pair's set semantics: a pair seen on several ranks counts once, first distance
(minimum order key) wins.  Everything here is transport and bookkeeping; the bucketing, the
insert and the classify/aggregate kernels are in liblinks_gpu.so.  The functions take a
torch.distributed process group so the same code runs over NCCL (device tensors) and, in the CPU
tests, over gloo.
"""
from __future__ import annotations

import numpy as np

RECORD_WORDS = 3  # {h_a, h_b, order_key} as int64 words (LINKS_PAIR_RECORD_BYTES = 24)
_M64 = np.uint64(0xFFFFFFFFFFFFFFFF)


def pair_owner(h_a: np.ndarray, h_b: np.ndarray, n_ranks: int) -> np.ndarray:
    """Vectorised links_pair_owner(): the owner rank of each ordered pair."""
    a = np.asarray(h_a, dtype=np.uint64)
    b = np.asarray(h_b, dtype=np.uint64)
    with np.errstate(over="ignore"):
        x = a * np.uint64(0x9E3779B97F4A7C15) + b
        x ^= x >> np.uint64(32)
        x *= np.uint64(0xD6E8FEB86659FD93)
        return (((x & np.uint64(0xFFFFFFFF)) * np.uint64(n_ranks)) >> np.uint64(32)).astype(np.uint32)


def exchange_records(send, counts, dist, group=None):
    """All-to-all of variable-size record buckets.  `send` is an int64 tensor [n, 3] grouped by
    owner rank, counts[r] = records for rank r.  Returns the received [m, 3] tensor."""
    import torch
    world = dist.get_world_size(group)
    dev = send.device
    c_out = torch.tensor([int(c) for c in counts], dtype=torch.int64, device=dev)
    c_in = torch.empty(world, dtype=torch.int64, device=dev)
    dist.all_to_all_single(c_in, c_out, group=group)
    in_counts = [int(v) for v in c_in.tolist()]
    recv = torch.empty((sum(in_counts), RECORD_WORDS), dtype=torch.int64, device=dev)
    dist.all_to_all_single(recv, send, output_split_sizes=in_counts, input_split_sizes=[int(c) for c in counts], group=group)
    return recv


def exchange_pairs(L, dist, device, group=None, profile: bool = False) -> dict:
    """export -> all-to-all -> clear -> import on a Links context.  Returns byte/record counts
    (and, with profile=True, host milliseconds per phase measured with device syncs)."""
    import time

    import torch
    world = dist.get_world_size(group)
    t = [time.perf_counter()]
    ms: dict[str, float] = {}

    def mark(name):
        if profile:
            torch.cuda.synchronize()
            now = time.perf_counter()
            ms[name] = (now - t[0]) * 1e3
            t[0] = now

    n = L.n_unique()
    send = torch.empty((max(n, 1), RECORD_WORDS), dtype=torch.int64, device=device)
    mark("alloc")
    counts = L.pairs_export(world, send.data_ptr(), max(n, 1))
    mark("export")
    recv = exchange_records(send[:n], counts, dist, group)
    mark("all_to_all")
    L.pairs_clear()
    mark("clear")
    if recv.shape[0]:
        L.pairs_import(recv.data_ptr(), recv.shape[0])
    mark("import")
    out = {"sent_records": int(n), "recv_records": int(recv.shape[0]), "sent_bytes": int(n) * 8 * RECORD_WORDS}
    if profile:
        out["host_ms"] = {k: round(v, 3) for k, v in ms.items()}
    return out


def replicate_bloom(L, nbytes: int, hash_num: int, dist, device, src: int = 0, group=None) -> None:
    """Rank `src` has built the filter; everyone else adopts its bits (they then skip the insert)."""
    import torch
    buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
    if dist.get_rank(group) == src:
        L.bf_download_dev(buf.data_ptr())
    dist.broadcast(buf, src=src, group=group)
    if dist.get_rank(group) != src:
        # the context may run on its own stream (no links_set_stream): the copy below must not overtake the broadcast
        torch.cuda.current_stream().synchronize()
        L.bf_load_dev(buf.data_ptr(), nbytes, hash_num)
        L.bf_covers_contigs(True)  # rank `src` built these bits from the same contigs every rank pushes next


def attach_exchange(L, dist, device, inbox_bytes: int, group=None):
    """Peer-memory exchange (include/links_gpu.h, links_exchange_attach): allocate this rank's inbox in
    symmetric memory, exchange the peer pointers, hand them to the context.  Returns the objects that
    must stay alive as long as the context.  After this, a job is
        reset -> Bloom filter -> contigs -> reads -> L.exchange_finish() -> L.finalize()
    and rank 0's finalize returns the summed link table; there is no all-to-all and no row gather."""
    import torch
    import torch.distributed._symmetric_memory as symm
    g = group if group is not None else dist.group.WORLD
    size = torch.tensor([int(inbox_bytes)], dtype=torch.int64, device=device)
    dist.all_reduce(size, op=dist.ReduceOp.MAX, group=g)  # one size for all ranks (the layout is shared)
    inbox_bytes = (int(size.item()) + 4095) // 4096 * 4096
    buf = symm.empty(int(inbox_bytes), dtype=torch.uint8, device=device)
    hdl = symm.rendezvous(buf, g)
    ptrs = [int(p) for p in hdl.buffer_ptrs]
    L.exchange_attach(dist.get_rank(g), dist.get_world_size(g), ptrs, int(inbox_bytes))
    torch.cuda.synchronize()
    dist.barrier(group=g)  # every header is zeroed before anyone writes to a peer
    return buf, hdl


def alloc_shared_bloom(dist, device, nbytes: int, group=None):
    """Peer-addressable Bloom buffers for links_bf_alloc_shared: (keep-alive objects, [pointer of rank r's buffer as seen here]).
    Allocate once; reuse for every job with the same filter size."""
    import torch
    import torch.distributed._symmetric_memory as symm
    g = group if group is not None else dist.group.WORLD
    buf = symm.empty((int(nbytes) + 4095) // 4096 * 4096, dtype=torch.uint8, device=device)
    hdl = symm.rendezvous(buf, g)
    return (buf, hdl), [int(p) for p in hdl.buffer_ptrs]


def build_bloom_shared(L, nbytes: int, hash_num: int, ptrs: list[int], push_contigs, dist, group=None) -> None:
    """One job's filter built in shares: every rank inserts 1/N of the contig k-mers (push_contigs() pushes ALL contigs, which
    also stay resident for stage 3), then the copies are OR-ed through peer memory (exchange.cuh bloom_or_reduce_kernel)."""
    g = group if group is not None else dist.group.WORLD
    L.bf_alloc_shared(nbytes, hash_num, dist.get_rank(g), dist.get_world_size(g), ptrs)
    push_contigs()
    L.sync()
    dist.barrier(group=g)   # every share is in its rank's copy
    L.bf_allreduce()
    dist.barrier(group=g)   # every slice is stored in every copy


_BIN_CODE = {-1: 0, 10: 1, 500: 2, 5000: 3, 10000: 4}  # get_distance_bin's range (src/LINKS.hpp:764-770)


def _row_key(rows: np.ndarray) -> np.ndarray:
    """One u64 per row, monotone in (bin, tig_a, rev_a, tig_b, rev_b); tig < 2^29 (LINKS_MAX_TIGS)."""
    b = rows["bin"].astype(np.int64)
    code = np.select([b < 0, b == 10, b == 500, b == 5000], [0, 1, 2, 3], default=4).astype(np.uint64)
    a = (rows["tig_a"].astype(np.uint64) << np.uint64(1)) | rows["rev_a"].astype(np.uint64)
    c = (rows["tig_b"].astype(np.uint64) << np.uint64(1)) | rows["rev_b"].astype(np.uint64)
    return (code << np.uint64(60)) | (a << np.uint64(30)) | c


def sum_rows(parts: list[np.ndarray], row_dtype) -> np.ndarray:
    """Sum partial link tables by (bin, tig_a, rev_a, tig_b, rev_b); output sorted by that key --
    the same order links_finalize returns (add_to_pair_map is additive, src/LINKS.hpp:738-761)."""
    parts = [p for p in parts if len(p)]
    if not parts:
        return np.empty(0, dtype=row_dtype)
    rows = np.concatenate(parts)
    key = _row_key(rows)
    order = np.argsort(key, kind="stable")
    rows, key = rows[order], key[order]
    head = np.ones(len(rows), dtype=bool)
    head[1:] = key[1:] != key[:-1]
    idx = np.flatnonzero(head)
    out = rows[idx].copy()
    out["links"] = np.add.reduceat(rows["links"].astype(np.uint64), idx)
    out["gaps"] = np.add.reduceat(rows["gaps"].astype(np.int64), idx)
    return out


def gather_rows(rows: np.ndarray, dist, device, dst: int = 0, group=None, merge=None):
    """Collect every rank's partial rows on rank `dst` and sum them; other ranks get None.
    `merge` (e.g. Links.rows_merge, the device sort + segmented sum) replaces the numpy sum_rows."""
    import torch
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    item = rows.dtype.itemsize
    n = torch.tensor([len(rows)], dtype=torch.int64, device=device)
    sizes = torch.empty(world, dtype=torch.int64, device=device)
    dist.all_gather_into_tensor(sizes, n, group=group)
    sizes = [int(v) for v in sizes.tolist()]
    cap = max(max(sizes), 1)
    mine = torch.empty(cap * item, dtype=torch.uint8, device=device)
    if len(rows):
        mine[:len(rows) * item] = torch.from_numpy(rows.view(np.uint8).reshape(-1)).to(device)
    allbuf = torch.empty(world * cap * item, dtype=torch.uint8, device=device)
    dist.all_gather_into_tensor(allbuf, mine, group=group)
    if rank != dst:
        return None
    host = allbuf.cpu().numpy().reshape(world, cap * item)
    parts = [host[r, :sizes[r] * item].view(rows.dtype) for r in range(world)]
    if merge is not None:
        return merge(np.concatenate(parts))
    return sum_rows(parts, rows.dtype)
