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
        bounds = [reads.n * i // n_ranks for i in range(n_ranks + 1)]
        lo, hi = bounds[r], bounds[r + 1]
        out = []
        with Links(LinksOptions(k=c.opts["k"], distances=list(c.opts["d"]), step_sizes=c.steps(), offset=c.opts["o"],
                                insert_stdev=c.opts["e"])) as L:
            L.exchange_attach(r, n_ranks, ptrs, inbox_bytes)
            barrier.wait(timeout=60)
            for job in range(n_jobs):
                if job:
                    L.reset()
                L.bf_alloc(nbytes, h)
                L.push_contigs(contigs)
                if hi > lo:  # with fewer reads than ranks some ranks only receive
                    L.push_reads(SeqSet(reads.seq, reads.off[lo:hi + 1], []), first_read_idx=lo)
                L.exchange_finish()
                rows, issues, stats = L.finalize()
                pa, pb, pd = L.pairs()
                out.append({"tsv": L.format_rows(rows), "issues": L.format_issues(issues), "stats": dict(stats),
                            "pa": pa.tobytes(), "pb": pb.tobytes(), "pd": pd.tobytes()})
            barrier.wait(timeout=60)  # nobody unmaps an inbox a peer may still write to
        results.put((r, "ok", out))
        barrier.wait(timeout=60)
        del peers
    except BaseException as e:  # noqa: BLE001
        import traceback
        results.put((r, "error", f"{e!r}\n{traceback.format_exc()}"))
        try:
            barrier.abort()
        except Exception:  # noqa: BLE001
            pass
