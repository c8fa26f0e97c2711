"""N>1 host logic on CPU: world_size-2 gloo run of the pair exchange and the row merge
(links_b200/multi.py), with the CPU oracle standing in for the per-rank kernels.  Rank r extracts
the pairs of ITS read shard, the distinct pairs are routed to their owner rank by the same hash
the export kernel uses, each owner classifies its share, rank 0 sums the partial link tables --
and the result must equal the reference binary's output for the whole job."""
from __future__ import annotations

import os
import socket

import numpy as np
import pytest

from .conftest import GoldenCase, sort_issue_records, sort_lines

OKEY_RANK_BITS, OKEY_DIST_BITS = 28, 8  # links_b200/csrc/kernels.cuh make_okey


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _dedup_min_okey(a, b, okey):
    order = np.lexsort((okey, b, a))
    a, b, okey = a[order], b[order], okey[order]
    head = np.ones(len(a), dtype=bool)
    head[1:] = (a[1:] != a[:-1]) | (b[1:] != b[:-1])
    return a[head], b[head], okey[head]


def _worker(rank: int, world: int, port: int, case_name: str, out_dir: str):
    import torch
    import torch.distributed as dist

    from links_b200 import multi
    from links_b200.seqio import SeqSet
    from oracle import oracle_py as oracle

    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        c = GoldenCase(case_name)
        contigs, reads = c.contigs(), c.reads()
        k, d, t, o, e = c.opts["k"], c.opts["d"], c.steps(), c.opts["o"], c.opts["e"]
        _m, nbytes, h = oracle.bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
        bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, k)
        # contiguous read shards, global read ordinals kept for the order key
        lo, hi = reads.n * rank // world, reads.n * (rank + 1) // world
        sub = SeqSet(reads.seq, reads.off[lo:hi + 1], reads.names[lo:hi])
        base = int(sub.off[0])
        _r, _i, stats, obs = oracle.run(k, d, t, o, e, bits, h, contigs.seq, contigs.off, sub.seq[base:int(sub.off[-1])],
                                        sub.off - np.uint64(base), keep_obs=True)
        okey = ((obs["read_idx"].astype(np.uint64) + np.uint64(lo)) << np.uint64(OKEY_RANK_BITS + OKEY_DIST_BITS)) \
            | (obs["d_idx"].astype(np.uint64) << np.uint64(OKEY_RANK_BITS)) | obs["rank"].astype(np.uint64)
        a, b, okey = _dedup_min_okey(obs["h_a"].copy(), obs["h_b"].copy(), okey)   # the rank's pair set
        owner = multi.pair_owner(a, b, world)
        order = np.argsort(owner, kind="stable")
        send = np.stack([a[order], b[order], okey[order]], axis=1).view(np.int64)
        counts = np.bincount(owner, minlength=world).tolist()
        recv = multi.exchange_records(torch.from_numpy(np.ascontiguousarray(send)), counts, dist).numpy().view(np.uint64)
        assert (multi.pair_owner(recv[:, 0], recv[:, 1], world) == rank).all()
        ra, rb, rk = _dedup_min_okey(recv[:, 0].copy(), recv[:, 1].copy(), recv[:, 2].copy())  # import: min order key wins
        d_idx = ((rk >> np.uint64(OKEY_RANK_BITS)) & np.uint64((1 << OKEY_DIST_BITS) - 1)).astype(np.uint32)
        rows, issues, st = oracle.run_from_pairs(k, d, t, o, e, contigs.seq, contigs.off, ra, rb, d_idx)
        merged = multi.gather_rows(rows, dist, torch.device("cpu"))
        all_issues = [None] * world
        dist.all_gather_object(all_issues, oracle.format_issues(issues))
        n_unique = torch.tensor([len(ra), stats["n_candidates"]], dtype=torch.int64)
        dist.all_reduce(n_unique)
        if rank == 0:
            open(os.path.join(out_dir, "rows.tsv"), "wb").write(oracle.format_rows(merged))
            open(os.path.join(out_dir, "issues.txt"), "wb").write(b"".join(all_issues))
            open(os.path.join(out_dir, "counts.txt"), "w").write(f"{int(n_unique[0])} {int(n_unique[1])}")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case_name", ["adv_tandem", "adv_offset_k20"])
def test_two_rank_exchange_and_merge_equals_reference(oracle, tmp_path, case_name):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), case_name, str(tmp_path)), nprocs=world, join=True)
    c = GoldenCase(case_name)
    assert sort_lines(open(tmp_path / "rows.tsv", "rb").read()) == c.expected_tsv()
    assert sort_issue_records(open(tmp_path / "issues.txt", "rb").read()) == c.expected_issues()
    # the owners' shares partition the whole job's distinct pair set
    contigs, reads = c.contigs(), c.reads()
    _m, nbytes, h = oracle.bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, c.opts["k"])
    _r, _i, stats, _ = oracle.run(c.opts["k"], c.opts["d"], c.steps(), c.opts["o"], c.opts["e"], bits, h, contigs.seq,
                                  contigs.off, reads.seq, reads.off)
    n_unique, n_cand = map(int, open(tmp_path / "counts.txt").read().split())
    assert n_unique == stats["n_unique_pairs"] and n_cand == stats["n_candidates"]


def test_pair_owner_matches_library(gpu_lib):
    """numpy restatement of the routing hash == links_pair_owner() (pure host code in the library)."""
    from links_b200 import multi
    rng = np.random.default_rng(11)
    a = rng.integers(0, 2**63, 2000, dtype=np.uint64) * np.uint64(2) + np.uint64(1)
    b = rng.integers(0, 2**63, 2000, dtype=np.uint64)
    for n in (1, 2, 3, 8, 255):
        got = multi.pair_owner(a, b, n)
        exp = np.array([gpu_lib.links_pair_owner(int(x), int(y), n) for x, y in zip(a, b)], dtype=np.uint32)
        assert np.array_equal(got, exp)
        assert got.max() < n
    assert np.bincount(multi.pair_owner(a, b, 8), minlength=8).min() > 150  # roughly uniform


def test_sum_rows_is_additive(oracle):
    from links_b200 import multi
    c = GoldenCase("c1")
    contigs, reads = c.contigs(), c.reads()
    _m, nbytes, h = oracle.bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, c.opts["k"])
    rows, _i, _s, obs = oracle.run(c.opts["k"], c.opts["d"], c.steps(), c.opts["o"], c.opts["e"], bits, h, contigs.seq,
                                   contigs.off, reads.seq, reads.off, keep_obs=True)
    first = {}
    for x, y, di in zip(obs["h_a"].tolist(), obs["h_b"].tolist(), obs["d_idx"].tolist()):
        first.setdefault((x, y), di)
    keys = np.array(list(first.keys()), dtype=np.uint64)
    dis = np.array(list(first.values()), dtype=np.uint32)
    owner = multi.pair_owner(keys[:, 0], keys[:, 1], 3)
    parts = []
    for r in range(3):
        m = owner == r
        parts.append(oracle.run_from_pairs(c.opts["k"], c.opts["d"], c.steps(), c.opts["o"], c.opts["e"], contigs.seq,
                                           contigs.off, keys[m, 0], keys[m, 1], dis[m])[0])
    merged = multi.sum_rows(parts, rows.dtype)
    assert oracle.format_rows(merged) == oracle.format_rows(rows)
    assert sort_lines(oracle.format_rows(merged)) == c.expected_tsv()
