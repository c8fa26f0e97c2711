#!/usr/bin/env python
"""Run one workload several times under different library switches and print the counters that must not move.

    python tools/determinism_probe.py --workload c4_human_3gbp --scale 0.05 [--runs 3]

MEASUREMENT / DEBUG PROBE.  Every configuration (member set on/off, member map on/off, contig table full/mates, chunk size)
must give the same candidates, survivors, distinct pairs, reference-order probe count and row digest, run after run:
the path is integer work and first-wins is defined by an order key, not by timing.  One JSON line per run.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c4_human_3gbp")
    ap.add_argument("--scale", type=float, default=0.05)
    ap.add_argument("--runs", type=int, default=3)
    ap.add_argument("--configs", default="default,noset,nomap,map,mates,full,chunk16")
    a = ap.parse_args()
    import torch

    import bench
    from links_b200 import Links, LinksOptions, bf_size
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(0)
    job = bench.load_job(a.workload, a.scale, 0, 1, dev, strong=True)
    _m, bf_bytes, hash_num = bf_size(job.file_bytes, 0.001)
    envs = {"default": {}, "noset": {"LINKS_MEMBER_SET": "0"}, "nomap": {"LINKS_MEMBER_MAP": "0"}, "map": {"LINKS_MEMBER_MAP": "1"},
            "mates": {"LINKS_TABLE_MODE": "mates"}, "full": {"LINKS_TABLE_MODE": "full"}, "chunk16": {"LINKS_CHUNK_BASES": str(16 << 20)}}
    for name in a.configs.split(","):
        for k in ("LINKS_MEMBER_SET", "LINKS_MEMBER_MAP", "LINKS_TABLE_MODE", "LINKS_CHUNK_BASES"):
            os.environ.pop(k, None)
        os.environ.update(envs[name])
        with Links(LinksOptions(k=job.k, distances=job.d, step_sizes=job.t, device=0)) as L:
            for run in range(a.runs):
                L.reset()
                L.bf_alloc(bf_bytes, hash_num)
                L.push_contigs_dev(job.d_contigs.data_ptr(), job.contigs_off, 1)
                L.push_reads_dev(job.d_reads.data_ptr(), job.reads_off, job.first_read)
                rows, issues, st = L.finalize()
                print(json.dumps({"config": name, "run": run, "candidates": st["n_candidates"], "survivors": st["n_survivors"],
                                  "unique_pairs": st["n_unique_pairs"], "bf_probes": st["n_bf_probes"], "bf_loads": st["n_bf_loads"],
                                  "member_sectors": st["n_member_sectors"], "rows": len(rows), "issues": len(issues),
                                  "ok_pairs": st["ct_ok_pairs"], "gate_fail": st["ct_gate_fail"], "rows_digest": bench.rows_digest(rows)}), flush=True)


if __name__ == "__main__":
    main()
