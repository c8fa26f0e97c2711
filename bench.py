#!/usr/bin/env python
"""bench.py -- the LINKS hot path (k-mer-pair extract -> Bloom probe -> contig k-mer map -> link
aggregation) on B200, on the configuration BASELINE.json's metric is quoted on.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME] [--scale S]

One step = one complete job on one synthetic batch: Bloom filter build from the contigs, every read
through extract + probe + pair-set insert, contig k-mer table build, classify, sort + reduce-by-key,
rows back on the host.  Nothing is cached between steps (links_reset keeps allocations only).

  value : candidate k-mer pairs/s with the ASCII inputs already resident in HBM
  e2e   : same job through the C-ABI with pinned HOST buffers (H2D of contigs+reads and D2H of rows
          inside the timed region)
N > 1 (torchrun): weak scaling -- every rank gets its own 50x read set over the same contigs, the
ranks' pair sets are exchanged with one NCCL all-to-all keyed on a hash of (h_a,h_b), each rank
finishes its share and rank 0 sums the partial link tables.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import shutil
import statistics
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
# one hardware queue per stream (copy, pack, compute, aux, import, torch's, NCCL's): a kernel waiting for a peer
# rank must never sit in front of this rank's own work.  Read by the driver when CUDA initialises.
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")

METRIC = "kmer_pairs_per_sec_extracted_probed_mapped"
UNIT = "pairs/s"
WORKLOADS = {
    # BASELINE.json configs[1]: synthetic 5 Mbp bacterial genome, 500 contigs, 50x ONT-like 10 kb reads, -k 15 -d 4000 -t 2
    "c2_bacterial_5mbp": "c2_bacterial_5mbp",
    # configs[2]: 100 Mbp draft, 30x 15 kb reads, -d 1000,2500,5000 -t 20,10,5 (not the headline line; run with --workload)
    "c3_celegans_100mbp": "c3_celegans_100mbp",
    # configs[4], densest point of the sweep: 100 Mbp, -d 10000 -t 1 (5x reads so that it stays a bench, not a soak)
    "c5_step_sweep_100mbp": "c5_step_sweep_100mbp",
    "tiny": "tiny",
}


_JSON_FD = 1


def claim_stdout() -> None:
    """The contract is ONE JSON line on stdout.  Libraries write there too (NCCL prints its version banner from C), so
    file descriptor 1 is pointed at stderr for the run and the JSON line goes to a private copy of the real stdout."""
    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)


def emit(obj: dict) -> None:
    sys.stdout.flush()
    os.write(_JSON_FD, (json.dumps(obj) + "\n").encode())


def env_int(name: str, default: int) -> int:
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def load_workload(name: str, scale: float, seed_shift: int = 0):
    """Synthetic inputs (cached as .npz under /tmp so back-to-back runs do not regenerate)."""
    from links_b200 import synth
    from links_b200.seqio import SeqSet
    g, nc, cov, rl, k, d, t = synth.CONFIGS[name]
    cache = os.path.join(tempfile.gettempdir(), f"links_b200_{name}_{scale}_{seed_shift}.npz")
    if os.path.exists(cache):
        z = np.load(cache)
        contigs = SeqSet(z["cs"], z["co"], [])
        reads = SeqSet(z["rs"], z["ro"], [])
    else:
        genome = synth.make_genome(int(g * scale), seed=1)
        contigs = synth.make_contigs(genome, max(int(nc * scale), 2), seed=3)
        reads = synth.make_reads(genome, cov, rl, seed=2 + 1000 * seed_shift)
        tmp = cache + f".{os.getpid()}.tmp.npz"
        np.savez(tmp, cs=contigs.seq, co=contigs.off, rs=reads.seq, ro=reads.off)
        os.replace(tmp, cache)
    return contigs, reads, k, d, t


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device = device
        self.proc = None
        self.lines: list[str] = []

    def start(self):
        if not shutil.which("nvidia-smi"):
            return
        self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.device), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-lms", "100"], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
        self.thread = threading.Thread(target=self._pump, daemon=True)
        self.thread.start()

    def _pump(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx = [], []
        reasons = set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, v in zip(names, f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def pinned_copy(lib, arr: np.ndarray):
    """Copy a numpy array into pinned host memory from links_host_alloc; returns (ptr, nbytes)."""
    n = arr.nbytes
    p = lib.links_host_alloc(max(n, 1))
    if not p:
        raise MemoryError("links_host_alloc failed")
    C.memmove(p, arr.ctypes.data, n)
    return p, n


def rows_digest(rows: np.ndarray) -> str:
    import hashlib
    keep = np.empty(len(rows), dtype=[("bin", "<i4"), ("a", "<u4"), ("b", "<u4"), ("ra", "u1"), ("rb", "u1"), ("l", "<u8"), ("g", "<i8")])
    keep["bin"], keep["a"], keep["b"], keep["ra"], keep["rb"], keep["l"], keep["g"] = (
        rows["bin"], rows["tig_a"], rows["tig_b"], rows["rev_a"], rows["rev_b"], rows["links"], rows["gaps"])
    return hashlib.sha256(keep.tobytes()).hexdigest()[:16]


def job_properties(rows, issues, stats, reads, k, d, t) -> dict:
    """Checks that hold at any size (tests/test_gpu_parity.py::test_full_size_properties has the same list)."""
    from links_b200 import synth

    def table(mirrored: bool):
        ta, ra, tb, rb = (rows[n].astype(np.int64) for n in ("tig_a", "rev_a", "tig_b", "rev_b"))
        if mirrored:
            ta, ra, tb, rb = tb, 1 - rb, ta, 1 - ra
        m = np.stack([rows["bin"].astype(np.int64), ta, ra, tb, rb, rows["links"].astype(np.int64), rows["gaps"].astype(np.int64)], axis=1)
        return m[np.lexsort(m.T[::-1])]
    classes = sum(stats[n] for n in ("ct_ok_pairs", "ct_problem_pairs", "ct_ok_contig", "ct_iz_issues", "ct_illogical", "ct_gate_fail"))
    closed = synth.candidate_pairs(reads.off, k, list(d), list(t))  # exact for reads without an invalid base (0.1 % have an N)
    return {"candidates_within_0.1pct_below_closed_form": closed - closed // 1000 <= stats["n_candidates"] <= closed,
            "classes_sum_to_unique_pairs": classes == stats["n_unique_pairs"],
            "issue_records_equal_counter": len(issues) == stats["ct_iz_issues"],
            "links_sum_twice_ok_pairs": int(rows["links"].sum()) == 2 * stats["ct_ok_pairs"],
            "rows_closed_under_mirror": bool(len(rows) % 2 == 0 and np.array_equal(table(False), table(True)))}


# ---------------------------------------------------------------------------------------------- reference arm

def reference_sample(contigs, reads, k, d, t, target_cand: int):
    """First reads of the workload up to ~target_cand candidate pairs."""
    from links_b200 import synth
    from links_b200.seqio import SeqSet
    lens = np.diff(reads.off.astype(np.int64))
    per_read = np.zeros(len(lens), dtype=np.int64)
    for dd, tt in zip(d, t):
        per_read += np.maximum(lens - k + 1 - dd, 0) // tt
    cum = np.cumsum(per_read)
    n = int(min(np.searchsorted(cum, target_cand) + 1, len(lens)))
    sub = SeqSet(reads.seq[:int(reads.off[n])], reads.off[:n + 1], [f"r{i}" for i in range(n)])
    return sub, int(cum[n - 1])


def run_reference_once(workdir: str, k, d, t, threads: int) -> float:
    """One run of the UNMODIFIED reference binary (oracle/_ref/LINKS_CPP) on files in workdir; returns seconds."""
    from oracle import oracle_py
    cmd = [oracle_py.REF_BIN, "-f", "contigs.fa", "-s", "reads.fof", "-b", "ref_out", "-j", str(threads), "-k", str(k),
           "-d", ",".join(map(str, d)), "-t", ",".join(map(str, t))]
    for suffix in (".bloom", ".tigpair_checkpoint.tsv", ".pairing_issues"):
        try:
            os.remove(os.path.join(workdir, "ref_out" + suffix))
        except FileNotFoundError:
            pass
    t0 = time.perf_counter()
    subprocess.run(cmd, cwd=workdir, check=True, stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    return time.perf_counter() - t0


def prepare_reference_dir(contigs, sample) -> str:
    from links_b200.seqio import write_fasta
    wd = tempfile.mkdtemp(prefix="links_ref_")
    contigs.names = [str(i + 1) for i in range(contigs.n)]
    write_fasta(os.path.join(wd, "contigs.fa"), contigs)
    write_fasta(os.path.join(wd, "reads.fa"), sample)
    open(os.path.join(wd, "reads.fof"), "w").write("reads.fa\n")
    return wd


def cpu_reference(contigs, reads, k, d, t, target_cand: int, runs: int, warm: int):
    from oracle import oracle_py
    if not oracle_py.have_ref_bin():
        return None
    sample, cand = reference_sample(contigs, reads, k, d, t, target_cand)
    wd = prepare_reference_dir(contigs, sample)
    threads = min(4, os.cpu_count() or 1)  # the reference errors above -j 4 (README.md:45)
    try:
        for _ in range(warm):
            run_reference_once(wd, k, d, t, threads)
        secs = [run_reference_once(wd, k, d, t, threads) for _ in range(runs)]
    finally:
        shutil.rmtree(wd, ignore_errors=True)
    return {"cand": cand, "secs": secs, "threads": threads, "n_reads": sample.n, "bases": int(sample.off[-1])}


def main_reference(args, rank: int, world: int):
    if rank != 0:
        return
    contigs, reads, k, d, t = load_workload(args.workload, args.scale)
    r = cpu_reference(contigs, reads, k, d, t, args.ref_candidates, runs=args.steps, warm=min(args.warmup, 1))
    if r is None:
        emit({"impl": "reference", "unavailable": "oracle/_ref/LINKS_CPP not built (needs /root/reference at build time)"})
        return
    ms = 1000.0 * sum(r["secs"]) / len(r["secs"])
    value = r["cand"] / (ms / 1000.0)
    sample = (f"first {r['n_reads']} reads ({r['bases']} bases, {r['cand']} candidate pairs) of {args.workload} + all contigs; "
              f"unmodified reference LINKS_CPP -j {r['threads']} from FASTA files, process wall time")
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
           "data": "synthetic", "impl": "reference",
           "config": {"workload": args.workload, "scale": args.scale, "k": k, "d": d, "t": t, "sample": sample},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["threads"], "kind": "reference", "sample": sample},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "host_cores": os.cpu_count()}
    emit(out)


# ---------------------------------------------------------------------------------------------- our arm

def main_ours(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist

    from links_b200 import Links, LinksOptions, _abi, bf_size, multi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    lib = _abi.load()
    from links_b200.synth import CONFIGS as synth_configs

    contigs, reads, k, d, t = load_workload(args.workload, args.scale, seed_shift=rank)
    file_bytes = int(contigs.off[-1]) + contigs.n * (2 + len(str(contigs.n)))  # what the FASTA would weigh on disk
    _m, bf_bytes, hash_num = bf_size(file_bytes, 0.001)

    L = Links(LinksOptions(k=k, distances=d, step_sizes=t, device=local_rank))
    stream = torch.cuda.current_stream()
    L.set_stream(stream.cuda_stream)
    # N > 1: pairs travel to their owner ranks through NVLink peer memory while the reads are still being
    # extracted (links_exchange_attach); LINKS_EXCHANGE=nccl selects the stop-and-exchange all-to-all instead
    p2p = world > 1 and os.environ.get("LINKS_EXCHANGE", "p2p") != "nccl"
    bloom_bcast = os.environ.get("LINKS_BLOOM", "local") == "broadcast"
    keep_alive = None
    if p2p:
        est_pairs = int(reads.off[-1]) // 12  # ~0.06 distinct pairs per read base on this workload, x1.4
        keep_alive = multi.attach_exchange(L, dist, torch.device("cuda", local_rank), inbox_bytes=max(64 << 20, est_pairs * 24 * 2 * 2))

    # inputs: pinned host copies (e2e leg) and device-resident copies (value leg)
    pc_seq, nb_c = pinned_copy(lib, contigs.seq)
    pc_off, _ = pinned_copy(lib, contigs.off)
    pr_seq, nb_r = pinned_copy(lib, reads.seq)
    pr_off, _ = pinned_copy(lib, reads.off)
    d_contigs = torch.from_numpy(contigs.seq).cuda()
    d_reads = torch.from_numpy(reads.seq).cuda()
    h2d_bytes = nb_c + nb_r + contigs.off.nbytes + reads.off.nbytes

    state = {}

    dev = torch.device("cuda", local_rank)
    first_read = rank * (1 << 22)  # disjoint global read ordinals per rank (order key, n_dist > 1)

    phases = os.environ.get("LINKS_BENCH_PHASES") == "1"  # diagnostic: host wall time per phase (adds syncs; not a bench number)
    ph: dict[str, float] = {}
    ph_t = [0.0]

    def mark(name: str | None):
        if not phases:
            return
        torch.cuda.synchronize()
        now = time.perf_counter()
        if name:
            ph[name] = (now - ph_t[0]) * 1e3  # last step wins: steady state, not the first-call set-up costs
        ph_t[0] = now

    def step(resident: bool):
        mark(None)
        _step(resident)

    def _step(resident: bool):
        L.reset()
        mark("reset")
        if world > 1 and bloom_bcast:
            # north_star's variant: rank 0 builds the Bloom filter, NCCL broadcast replicates it, the others adopt it.
            # It serialises the ranks (everyone waits for rank 0's build) -- measured 1.7 ms of a 10 ms step at N=2.
            if rank == 0:
                L.bf_alloc(bf_bytes, hash_num)
                (L.push_contigs_dev(d_contigs.data_ptr(), contigs.off, 1) if resident
                 else L.push_contigs_raw(pc_seq, pc_off, contigs.n, 1))
            multi.replicate_bloom(L, bf_bytes, hash_num, dist, dev)
            if rank != 0:
                (L.push_contigs_dev(d_contigs.data_ptr(), contigs.off, 1) if resident
                 else L.push_contigs_raw(pc_seq, pc_off, contigs.n, 1))
        else:
            # default (SURVEY 8e's alternative): the contigs are replicated anyway, so every rank builds the same
            # filter itself (0.3 ms, deterministic, bit-identical) and nobody waits
            L.bf_alloc(bf_bytes, hash_num)
            if resident:
                L.push_contigs_dev(d_contigs.data_ptr(), contigs.off, 1)
            else:
                L.push_contigs_raw(pc_seq, pc_off, contigs.n, 1)
        mark("bloom+contigs")
        if resident:
            L.push_reads_dev(d_reads.data_ptr(), reads.off, first_read)
        else:
            L.push_reads_raw(pr_seq, pr_off, reads.n, first_read)
        mark("reads")
        if p2p:
            L.exchange_finish()
            state["exchange"] = {"transport": "peer memory (NVLink P2P), routed per extract launch"}
        elif world > 1:
            state["exchange"] = multi.exchange_pairs(L, dist, dev, profile=phases)
            mark("exchange")
        rows, issues, stats = L.finalize()
        mark("finalize")
        if world > 1 and not p2p:
            rows = multi.gather_rows(rows, dist, dev, merge=L.rows_merge)
            mark("gather_rows")
        state["rows"], state["issues"], state["stats"] = rows, issues, stats

    def timed(resident: bool, steps: int, warmup: int):
        for _ in range(warmup):
            step(resident)
        L.timing(reset=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(steps):
            step(resident)
        e1.record(stream)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            tms = torch.tensor([ms], device="cuda", dtype=torch.float64)
            dist.all_reduce(tms, op=dist.ReduceOp.MAX)
            ms = float(tms.item())
        return ms / steps, L.timing(reset=True)

    clocks = ClockSampler(local_rank)
    clocks.start()
    ms_res, tm_res = timed(True, args.steps, args.warmup)
    digest_res = rows_digest(state["rows"]) if rank == 0 else ""
    stats = dict(state["stats"])
    ms_e2e, tm_e2e = timed(False, args.steps, max(args.warmup, 1))
    clk = clocks.stop()
    if phases:
        print(f"[rank {rank}] host ms per phase, last step: " + json.dumps({k2: round(v, 2) for k2, v in ph.items()}), file=sys.stderr)
    digest_e2e = rows_digest(state["rows"]) if rank == 0 else ""

    cand = stats["n_candidates"]
    cand_total = cand
    if world > 1:
        tc = torch.tensor([cand], device="cuda", dtype=torch.int64)
        dist.all_reduce(tc)
        cand_total = int(tc.item())
    if rank == 0:
        assert digest_res == digest_e2e, "device-resident and host-buffer legs disagree"
        n_rows = len(state["rows"])
        d2h_bytes = n_rows // 2 * 24 + len(state["issues"]) * 72 + 3 * 112
        # roofline of the dominant kernel (extract_kernel): algorithmic bytes per launch / mean launch duration
        n_launch = max(int(tm_res["extract_launches"]), 1)
        steps_l = args.steps
        alg_bytes_step = cand * (t[0] / 4.0) + stats["n_bf_probes"] * 32.0 + stats["n_survivors"] * 64.0
        alg_bytes_launch = alg_bytes_step * steps_l / n_launch
        ms_launch = tm_res["ms_extract"] / n_launch
        achieved = alg_bytes_launch / (ms_launch * 1e-3) / 1e9 if ms_launch > 0 else 0.0
        peak, peak_src = 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"
        try:
            mp = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
            peak, peak_src = float(mp["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
        except Exception:
            pass
        traffic = None  # DRAM bytes per launch of the dominant kernel, from the committed ncu capture (per candidate x candidates per launch)
        try:
            tr = json.load(open(os.path.join(ROOT, "profiles", "extract_traffic.json")))
            traffic = float(tr["dram_bytes_per_candidate"]) * cand * steps_l / n_launch
        except Exception:
            pass
        out = {
            "metric": METRIC, "value": cand_total / (ms_res * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms_res, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": args.workload, "scale": args.scale, "k": k, "d": d, "t": t, "genome_bp": int(synth_configs[args.workload][0] * args.scale),
                       "contigs": contigs.n, "reads_per_gpu": reads.n, "read_bases_per_gpu": int(reads.off[-1]),
                       "candidate_pairs_per_gpu": cand, "survivors": stats["n_survivors"], "unique_pairs": stats["n_unique_pairs"],
                       "bf_probes": stats["n_bf_probes"], "rows": n_rows, "rows_digest": digest_res,
                       "l2": "inputs (250 MB reads, GB-scale pair set and contig table) exceed the 126 MB L2; no flush",
                       "sharding": ("reads by rank (each rank its own 50x read set), contigs replicated, Bloom filter "
                                    + ("NCCL-broadcast from rank 0, " if bloom_bcast else "built by every rank from the replicated contigs (LINKS_BLOOM=broadcast: NCCL broadcast from rank 0), ")
                                    + ("distinct pairs routed to their owner rank (hash of (h_a,h_b)) through NVLink peer memory by a kernel after "
                                       "every extract launch and imported there while extraction continues, partial link rows summed on rank 0 on the device"
                                       if p2p else "distinct pairs routed by one NCCL all-to-all keyed on hash(h_a,h_b), rows summed on rank 0")) if world > 1 else "single GPU",
                       "exchange_rank0": state.get("exchange")},
            "e2e": {"value": cand_total / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": int(h2d_bytes),
                    "d2h_bytes_per_step": int(d2h_bytes)},
            "gpu_launches": int(tm_res["launches"]),
            "clocks": clk,
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "traffic_source": "profiles/extract_traffic.json (ncu dram__bytes_read+write per candidate pair, c2 workload) x candidates per launch",
                         "kernel": "extract_kernel", "launches": n_launch, "ms_per_launch": ms_launch,
                         "algorithmic_bytes_per_launch": alg_bytes_launch,
                         "bytes_per_pair": alg_bytes_step / max(cand, 1), "peak_source": peak_src,
                         # what the kernel really issued: k-mers proven present by the exact member set skip their probes
                         "issued_sectors_per_pair": {"bloom": stats.get("n_bf_loads", 0) / max(cand, 1),
                                                     "member_set": stats.get("n_member_sectors", 0) / max(cand, 1),
                                                     "reference_order_bloom": stats["n_bf_probes"] / max(cand, 1)}},
            "kernel_ms_per_step": {kk: vv / args.steps for kk, vv in tm_res.items() if kk.startswith("ms_")},
            "kernel_ms_per_step_e2e": {kk: vv / args.steps for kk, vv in tm_e2e.items() if kk.startswith("ms_")},
        }
        if args.properties:
            out["config"]["properties"] = job_properties(state["rows"], state["issues"], stats, reads, k, d, t) if world == 1 else "N=1 only"
        if world == 1 and not args.no_cpu_baseline:
            r = cpu_reference(contigs, reads, k, d, t, args.ref_candidates, runs=1, warm=0)
            if r is not None:
                v = r["cand"] / r["secs"][0]
                out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": r["threads"], "kind": "reference",
                                       "sample": f"first {r['n_reads']} reads ({r['cand']} candidate pairs) + all contigs, unmodified "
                                                 f"reference LINKS_CPP -j {r['threads']} from FASTA, {r['secs'][0]:.1f} s wall; host has {os.cpu_count()} cores"}
            else:
                out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref/LINKS_CPP missing"}
        emit(out)
    L.close()
    for p in (pc_seq, pc_off, pr_seq, pr_off):
        lib.links_host_free(p)
    if world > 1:
        dist.destroy_process_group()


def check_sharding(args, rank: int, world: int, local_rank: int):
    """Strong-scaling parity: the SAME reads split over the ranks (contiguous shards, global read
    ordinals) must give, after the pair exchange and the row merge, exactly the rows one GPU
    computes for the whole job.  Prints one JSON line on rank 0; not a timing run."""
    import torch
    import torch.distributed as dist

    from links_b200 import Links, LinksOptions, bf_size, multi
    from links_b200.seqio import SeqSet
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    contigs, reads, k, d, t = load_workload(args.workload, args.scale)
    file_bytes = int(contigs.off[-1]) + contigs.n * (2 + len(str(contigs.n)))
    _m, bf_bytes, hash_num = bf_size(file_bytes, 0.001)
    lo, hi = reads.n * rank // world, reads.n * (rank + 1) // world
    with Links(LinksOptions(k=k, distances=d, step_sizes=t, device=local_rank)) as L:
        p2p = world > 1 and os.environ.get("LINKS_EXCHANGE", "p2p") != "nccl"
        keep_alive = multi.attach_exchange(L, dist, dev, inbox_bytes=max(64 << 20, int(reads.off[-1]) // 12 * 24 * 4)) if p2p else None
        if rank == 0:
            L.bf_alloc(bf_bytes, hash_num)
            L.push_contigs(contigs)
        if world > 1:
            multi.replicate_bloom(L, bf_bytes, hash_num, dist, dev)
            if rank != 0:
                L.push_contigs(contigs)
        L.push_reads(SeqSet(reads.seq, reads.off[lo:hi + 1], []), first_read_idx=lo)
        if p2p:
            L.exchange_finish()
            ex = {"transport": "peer memory"}
        else:
            ex = multi.exchange_pairs(L, dist, dev) if world > 1 else {}
        rows, issues, stats = L.finalize()
        if world > 1 and not p2p:
            rows = multi.gather_rows(rows, dist, dev)
    tot = torch.tensor([stats["n_candidates"], stats["n_unique_pairs"], len(issues)], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    if rank == 0:
        with Links(LinksOptions(k=k, distances=d, step_sizes=t, device=local_rank)) as L:
            L.bf_alloc(bf_bytes, hash_num)
            L.push_contigs(contigs)
            L.push_reads(reads)
            rows1, issues1, stats1 = L.finalize()
        ok = (rows_digest(rows) == rows_digest(rows1) and int(tot[0]) == stats1["n_candidates"]
              and int(tot[1]) == stats1["n_unique_pairs"] and int(tot[2]) == len(issues1))
        emit({"check_sharding": "ok" if ok else "MISMATCH", "n_gpus": world, "workload": args.workload, "scale": args.scale,
                          "rows": len(rows1), "rows_digest_sharded": rows_digest(rows), "rows_digest_single": rows_digest(rows1),
                          "candidates": [int(tot[0]), stats1["n_candidates"]], "unique_pairs": [int(tot[1]), stats1["n_unique_pairs"]],
                          "issues": [int(tot[2]), len(issues1)], "exchange_rank0": ex})
        if not ok:
            raise SystemExit(1)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="c2_bacterial_5mbp", choices=sorted(WORKLOADS))
    ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--ref-candidates", type=int, default=4_000_000, help="size of the bounded CPU sample (candidate pairs)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--properties", action="store_true", help="size-independent checks of the last job's result (counts balance, mirror closure)")
    ap.add_argument("--check-sharding", action="store_true", help="parity of the N-rank path against one GPU (same reads split over ranks)")
    args = ap.parse_args()
    claim_stdout()
    rank, world, local_rank = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if args.impl == "reference":
        main_reference(args, rank, world)
    elif args.check_sharding:
        check_sharding(args, rank, world, local_rank)
    else:
        main_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
