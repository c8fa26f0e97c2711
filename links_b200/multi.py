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


def exchange_pairs(L, dist, device, group=None) -> dict:
    """export -> all-to-all -> clear -> import on a Links context.  Returns byte/record counts."""
    import torch
    world = dist.get_world_size(group)
    n = L.n_unique()
    send = torch.empty((max(n, 1), RECORD_WORDS), dtype=torch.int64, device=device)
    counts = L.pairs_export(world, send.data_ptr(), max(n, 1))
    recv = exchange_records(send[:n], counts, dist, group)
    L.pairs_clear()
    if recv.shape[0]:
        L.pairs_import(recv.data_ptr(), recv.shape[0])
    return {"sent_records": int(n), "recv_records": int(recv.shape[0]), "sent_bytes": int(n) * 8 * RECORD_WORDS}


def replicate_bloom(L, nbytes: int, hash_num: int, dist, device, src: int = 0, group=None) -> None:
    """Rank `src` has built the filter; everyone else adopts its bits (they then skip the insert)."""
    import torch
    buf = torch.empty(nbytes, dtype=torch.uint8, device=device)
    if dist.get_rank(group) == src:
        L.bf_download_dev(buf.data_ptr())
    dist.broadcast(buf, src=src, group=group)
    if dist.get_rank(group) != src:
        L.bf_load_dev(buf.data_ptr(), nbytes, hash_num)


def _row_keys(rows: np.ndarray) -> np.ndarray:
    k = np.empty((len(rows), 5), dtype=np.int64)
    k[:, 0], k[:, 1], k[:, 2], k[:, 3], k[:, 4] = rows["bin"], rows["tig_a"], rows["rev_a"], rows["tig_b"], rows["rev_b"]
    return k


def sum_rows(parts: list[np.ndarray], row_dtype) -> np.ndarray:
    """Sum partial link tables by (bin, tig_a, rev_a, tig_b, rev_b); output sorted by that key --
    the same order links_finalize returns (add_to_pair_map is additive, src/LINKS.hpp:738-761)."""
    parts = [p for p in parts if len(p)]
    if not parts:
        return np.empty(0, dtype=row_dtype)
    rows = np.concatenate(parts)
    k = _row_keys(rows)
    order = np.lexsort((k[:, 4], k[:, 3], k[:, 2], k[:, 1], k[:, 0]))
    rows, k = rows[order], k[order]
    head = np.ones(len(rows), dtype=bool)
    head[1:] = (k[1:] != k[:-1]).any(axis=1)
    idx = np.flatnonzero(head)
    out = rows[idx].copy()
    out["links"] = np.add.reduceat(rows["links"].astype(np.uint64), idx)
    out["gaps"] = np.add.reduceat(rows["gaps"].astype(np.int64), idx)
    return out


def gather_rows(rows: np.ndarray, dist, device, dst: int = 0, group=None):
    """Collect every rank's partial rows on rank `dst` and sum them; other ranks get None."""
    import torch
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    raw = np.frombuffer(rows.tobytes(), dtype=np.uint8)
    n = torch.tensor([raw.size], dtype=torch.int64, device=device)
    sizes = [torch.zeros(1, dtype=torch.int64, device=device) for _ in range(world)]
    dist.all_gather(sizes, n, group=group)
    sizes = [int(s.item()) for s in sizes]
    cap = max(max(sizes), 1)
    mine = torch.zeros(cap, dtype=torch.uint8, device=device)
    if raw.size:
        mine[:raw.size] = torch.from_numpy(raw.copy()).to(device)
    bufs = [torch.empty(cap, dtype=torch.uint8, device=device) for _ in range(world)]
    dist.all_gather(bufs, mine, group=group)
    if rank != dst:
        return None
    parts = [np.frombuffer(b[:s].cpu().numpy().tobytes(), dtype=rows.dtype) for b, s in zip(bufs, sizes)]
    return sum_rows(parts, rows.dtype)
