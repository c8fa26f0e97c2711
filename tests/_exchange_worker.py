"""One "rank" of the peer-memory exchange test: a separate PROCESS with its own CUDA context on GPU 0,
as in deployment (one process per GPU).  The inboxes are ordinary device buffers shared between the
processes through CUDA IPC (torch.multiprocessing), which is exactly what peer memory is on an NVLink
box.  Run by tests/test_gpu_parity.py::test_peer_memory_exchange_equals_single_context."""
from __future__ import annotations

import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def rank_main(r: int, n_ranks: int, name: str, inbox_bytes: int, queues, results, barrier, n_jobs: int, dedup: str = "owner") -> None:
    os.environ.setdefault("LINKS_EXCHANGE_TIMEOUT_MS", "20000")
    # dedup = owner | local, optionally "+sharedbf" (Bloom filter built in shares and OR-ed over peer memory) and "+mates"
    # (contig table restricted to the hashes of the pairs this rank owns)
    dedup, *extras = dedup.split("+")
    os.environ["LINKS_EXCHANGE_DEDUP"] = dedup  # owner: pairs leave from inside extract_kernel; local: local set + route_kernel
    if ROOT not in sys.path:
        sys.path.insert(0, ROOT)
    try:
        import torch

        from links_b200 import Links, LinksOptions, bf_size
        from links_b200.seqio import SeqSet
        from tests.conftest import GoldenCase
        c = GoldenCase(name)
        contigs, reads = c.contigs(), c.reads()
        _m, nbytes, h = bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
        torch.cuda.set_device(0)
        mine = torch.zeros(inbox_bytes, dtype=torch.uint8, device="cuda")
        torch.cuda.synchronize()
        for d in range(n_ranks):
            if d != r:
                queues[d].put((r, mine))  # CUDA IPC handle travels, not the bytes
        peers = {r: mine}
        while len(peers) < n_ranks:
            s, t = queues[r].get(timeout=60)
            peers[s] = t
        ptrs = [int(peers[i].data_ptr()) for i in range(n_ranks)]
        bf_ptrs, bf_peers = None, None
        if "sharedbf" in extras:
            barrier.wait(timeout=60)  # the inbox handles have all been taken off the queues
            bf_mine = torch.zeros(nbytes, dtype=torch.uint8, device="cuda")
            torch.cuda.synchronize()
            for d in range(n_ranks):
                if d != r:
                    queues[d].put((r, bf_mine))
            bf_peers = {r: bf_mine}
            while len(bf_peers) < n_ranks:
                s, t = queues[r].get(timeout=60)
                bf_peers[s] = t
            bf_ptrs = [int(bf_peers[i].data_ptr()) for i in range(n_ranks)]
        bounds = [reads.n * i // n_ranks for i in range(n_ranks + 1)]
        lo, hi = bounds[r], bounds[r + 1]
        out = []
        with Links(LinksOptions(k=c.opts["k"], distances=list(c.opts["d"]), step_sizes=c.steps(), offset=c.opts["o"],
                                insert_stdev=c.opts["e"])) as L:
            L.exchange_attach(r, n_ranks, ptrs, inbox_bytes)
            if "mates" in extras:
                L.table_mode("mates")
            barrier.wait(timeout=60)
            for job in range(n_jobs):
                if job:
                    L.reset()
                if bf_ptrs:
                    L.bf_alloc_shared(nbytes, h, r, n_ranks, bf_ptrs)
                    L.push_contigs(contigs)
                    L.sync()
                    barrier.wait(timeout=60)
                    L.bf_allreduce()
                    barrier.wait(timeout=60)
                else:
                    L.bf_alloc(nbytes, h)
                    L.push_contigs(contigs)
                bloom_sha = None
                if bf_ptrs:
                    import hashlib

                    from links_b200 import bf_header
                    bloom_sha = hashlib.sha256(bf_header(c.opts["k"], h, nbytes) + L.bf_bits().tobytes()).hexdigest()
                if hi > lo:  # with fewer reads than ranks some ranks only receive
                    L.push_reads(SeqSet(reads.seq, reads.off[lo:hi + 1], []), first_read_idx=lo)
                L.exchange_finish()
                rows, issues, stats = L.finalize()
                pa, pb, pd = L.pairs()
                out.append({"tsv": L.format_rows(rows), "issues": L.format_issues(issues), "stats": dict(stats), "bloom_sha": bloom_sha,
                            "pa": pa.tobytes(), "pb": pb.tobytes(), "pd": pd.tobytes()})
            barrier.wait(timeout=60)  # nobody unmaps an inbox a peer may still write to
        results.put((r, "ok", out))
        barrier.wait(timeout=60)
        del peers, bf_peers
    except BaseException as e:  # noqa: BLE001
        import traceback
        results.put((r, "error", f"{e!r}\n{traceback.format_exc()}"))
        try:
            barrier.abort()
        except Exception:  # noqa: BLE001
            pass
