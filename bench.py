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
N > 1 (torchrun): weak scaling on the default workload -- every rank gets its own 50x read set over the same contigs;
every surviving pair is written by extract_kernel itself into its OWNER rank's inbox (hash of (h_a,h_b)) over NVLink peer
memory and de-duplicated there while extraction continues; each rank classifies what it owns and rank 0 sums the partial
link tables on the device.  No NCCL call on the data path (LINKS_EXCHANGE=nccl selects the older all-to-all).
The default line also carries other regimes under "extra" (not the headline): C3 at full size at N=1 (a Bloom filter
that lives in DRAM) and, at N>1, C4 (human scale, device-generated, ONE read set split over the ranks, genome scaled
to N/8 of 3.1 Gbp).  --workload c4_human_3gbp / --sweep / --check-sharding run those on their own.
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
    # configs[3]: 3.1 Gbp human-scale draft, 20x 20 kb reads, -k 20 -d 5000 -t 5; generated on the device (links_b200/synth_dev.py),
    # ONE read set split over the ranks (strong scaling); --scale 0.1 fits one GPU's time budget, --scale 1.0 wants 8 GPUs
    "c4_human_3gbp": "c4_human_3gbp",
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


def fasta_file_bytes(off: np.ndarray) -> int:
    """Size of the FASTA file the reference would stat() to size its Bloom filter (src/LINKS.hpp:349): records
    '>' + 1-based ordinal + newline + sequence + newline -- what tools/make_fullsize_golden.py writes."""
    n = len(off) - 1
    digits = np.floor(np.log10(np.arange(1, n + 1))).astype(np.int64) + 1
    return int(off[-1]) + int(digits.sum()) + 3 * n


def load_workload(name: str, scale: float, seed_shift: int = 0, coverage: float | None = None):
    """Synthetic inputs (cached as .npz under /tmp so back-to-back runs do not regenerate)."""
    from links_b200 import synth
    from links_b200.seqio import SeqSet
    g, nc, cov, rl, k, d, t = synth.CONFIGS[name]
    if coverage is not None:
        cov = coverage
    # keyed by what defines the bytes (several workloads share a genome and a read set)
    cache = os.path.join(tempfile.gettempdir(), f"links_b200_g{int(g * scale)}_c{max(int(nc * scale), 2)}_x{cov}_l{rl}_s{seed_shift}.npz")
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


class Job:
    """One rank's inputs of a workload: device-resident ASCII (value leg) + offsets on the host, and -- made on demand --
    pinned host copies for the end-to-end leg."""

    def __init__(self, name, scale, k, d, t, genome_bp, d_contigs, contigs_off, file_bytes, d_reads, reads_off, first_read,
                 scaling, note, host_contigs=None, host_reads=None):
        self.name, self.scale, self.k, self.d, self.t, self.genome_bp = name, scale, k, list(d), list(t), genome_bp
        self.d_contigs, self.contigs_off, self.file_bytes = d_contigs, np.ascontiguousarray(contigs_off, dtype=np.uint64), file_bytes
        self.d_reads, self.reads_off, self.first_read = d_reads, np.ascontiguousarray(reads_off, dtype=np.uint64), first_read
        self.scaling, self.note = scaling, note
        self.host_contigs, self.host_reads = host_contigs, host_reads  # numpy ASCII when the workload was made on the host
        self.n_contigs, self.n_reads = len(self.contigs_off) - 1, len(self.reads_off) - 1
        self.pinned = None

    def closed_form_candidates(self, n_reads: int | None = None) -> list[int]:
        """Per distance: sum over reads of floor((len-k+1-d)/t) (exact for reads without an invalid base)."""
        lens = np.diff(self.reads_off[:(n_reads if n_reads is not None else self.n_reads) + 1].astype(np.int64))
        return [int((np.maximum(lens - self.k + 1 - dd, 0) // tt).sum()) for dd, tt in zip(self.d, self.t)]

    def pin(self, lib, max_read_bytes: int):
        """Pinned host copies: all contigs + the first reads up to max_read_bytes (the whole set when it fits)."""
        import torch
        if self.pinned:
            return self.pinned
        n = self.n_reads
        if int(self.reads_off[-1]) > max_read_bytes:
            n = max(int(np.searchsorted(self.reads_off, max_read_bytes, side="right")) - 1, 1)
        rb = int(self.reads_off[n])
        cs = self.host_contigs if self.host_contigs is not None else self.d_contigs.cpu().numpy()
        rs = self.host_reads[:rb] if self.host_reads is not None else self.d_reads[:rb].cpu().numpy()
        pc_seq, nb_c = pinned_copy(lib, cs)
        pc_off, _ = pinned_copy(lib, self.contigs_off)
        pr_seq, nb_r = pinned_copy(lib, rs)
        roff = np.ascontiguousarray(self.reads_off[:n + 1])
        pr_off, _ = pinned_copy(lib, roff)
        self.pinned = {"pc_seq": pc_seq, "pc_off": pc_off, "pr_seq": pr_seq, "pr_off": pr_off, "n_reads": n,
                       "h2d_bytes": nb_c + nb_r + self.contigs_off.nbytes + roff.nbytes, "whole_job": n == self.n_reads}
        return self.pinned

    def unpin(self, lib):
        if self.pinned:
            for key in ("pc_seq", "pc_off", "pr_seq", "pr_off"):
                lib.links_host_free(self.pinned[key])
            self.pinned = None


def load_job(name: str, scale: float, rank: int, world: int, device, coverage: float | None = None, strong: bool = False,
             steps: list | None = None) -> Job:
    """Workloads of links_b200.synth (numpy, host) and links_b200.synth_dev (device generator, human scale).
    strong=False (numpy workloads, the driver's default): every rank gets its own read set of the configuration's size
    over the same contigs.  strong=True: the SAME read set split into contiguous shards, global read ordinals."""
    import torch
    from links_b200 import synth, synth_dev
    if name in synth_dev.CONFIGS:
        w = synth_dev.make_config_dev(name, scale, device, rank, world)  # always one job split over the ranks
        return Job(name, scale, w.k, w.d, steps or w.t, w.genome_bp, w.d_contigs, w.contigs_off, w.contigs_file_bytes, w.d_reads,
                   w.reads_off, w.first_read, "strong" if world > 1 else "weak",
                   f"device-generated (links_b200/synth_dev.py); reads {w.first_read}..{w.first_read + len(w.reads_off) - 1} of {w.n_reads_total}")
    contigs, reads, k, d, t = load_workload(name, scale, seed_shift=0 if strong else rank, coverage=coverage)
    seq, off, first = reads.seq, reads.off, rank * (1 << 22)  # weak: disjoint global ordinals per rank (order key, n_dist > 1)
    if strong and world > 1:
        lo, hi = reads.n * rank // world, reads.n * (rank + 1) // world
        base = int(reads.off[lo])
        seq, off, first = reads.seq[base:int(reads.off[hi])], reads.off[lo:hi + 1] - np.uint64(base), lo
    return Job(name, scale, k, d, steps or t, int(synth.CONFIGS[name][0] * scale), torch.from_numpy(contigs.seq).to(device), contigs.off,
               fasta_file_bytes(contigs.off), torch.from_numpy(np.ascontiguousarray(seq)).to(device), off, first,
               "strong" if strong and world > 1 else "weak", "numpy generator (links_b200/synth.py)", host_contigs=contigs.seq,
               host_reads=np.ascontiguousarray(seq))


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


def job_properties(rows, issues, stats, reads_off, k, d, t) -> dict:
    """Checks that hold at any size (tests/test_gpu_parity.py::test_full_size_properties has the same list)."""
    from links_b200 import synth

    def table(mirrored: bool):
        ta, ra, tb, rb = (rows[n].astype(np.int64) for n in ("tig_a", "rev_a", "tig_b", "rev_b"))
        if mirrored:
            ta, ra, tb, rb = tb, 1 - rb, ta, 1 - ra
        m = np.stack([rows["bin"].astype(np.int64), ta, ra, tb, rb, rows["links"].astype(np.int64), rows["gaps"].astype(np.int64)], axis=1)
        return m[np.lexsort(m.T[::-1])]
    classes = sum(stats[n] for n in ("ct_ok_pairs", "ct_problem_pairs", "ct_ok_contig", "ct_iz_issues", "ct_illogical", "ct_gate_fail"))
    closed = synth.candidate_pairs(reads_off, k, list(d), list(t))  # exact for reads without an invalid base (0.1 % have an N)
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


def _parse_stage_seconds(stdout: str) -> dict:
    """Seconds per stage from the reference's own stdout banners (src/main.cpp:36-64)."""
    import re
    out = {}
    for key, pat in (("bloom_s", r"Bloom filter generated or loaded in: (\d+)h (\d+)m (\d+)s (\d+)ms"),
                     ("reads_s", r"Read fasta stage completed in: (\d+)h (\d+)m (\d+)s (\d+)ms"),
                     ("contigs_s", r"Read contig stage completed in: (\d+)h (\d+)m (\d+)s (\d+)ms"),
                     ("pairing_s", r"Pair contig stage completed in: (\d+)h (\d+)m (\d+)s (\d+)ms")):
        m = re.search(pat, stdout)
        if m:
            h, mi, sec, ms = (int(x) for x in m.groups())
            out[key] = h * 3600 + mi * 60 + sec + ms / 1000.0
    return out


def run_reference_once(workdir: str, k, d, t, threads: int, reads_fof: str = "reads.fof"):
    """One run of the UNMODIFIED reference binary (oracle/_ref/LINKS_CPP) on files in workdir; (wall seconds, stage seconds)."""
    from oracle import oracle_py
    cmd = [oracle_py.REF_BIN, "-f", "contigs.fa", "-s", reads_fof, "-b", "ref_out", "-j", str(threads), "-k", str(k),
           "-d", ",".join(map(str, d)), "-t", ",".join(map(str, t))]
    for suffix in (".bloom", ".tigpair_checkpoint.tsv", ".pairing_issues"):
        try:
            os.remove(os.path.join(workdir, "ref_out" + suffix))
        except FileNotFoundError:
            pass
    t0 = time.perf_counter()
    r = subprocess.run(cmd, cwd=workdir, check=True, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    return time.perf_counter() - t0, _parse_stage_seconds(r.stdout)


def prepare_reference_dir(contigs, sample, full=None) -> str:
    from links_b200.seqio import write_fasta
    wd = tempfile.mkdtemp(prefix="links_ref_")
    contigs.names = [str(i + 1) for i in range(contigs.n)]
    write_fasta(os.path.join(wd, "contigs.fa"), contigs)
    write_fasta(os.path.join(wd, "reads.fa"), sample)
    open(os.path.join(wd, "reads.fof"), "w").write("reads.fa\n")
    if full is not None:
        full.names = [f"r{i}" for i in range(full.n)]
        write_fasta(os.path.join(wd, "reads_full.fa"), full)
        open(os.path.join(wd, "reads_full.fof"), "w").write("reads_full.fa\n")
    return wd


def cpu_reference(contigs, reads, k, d, t, target_cand: int, runs: int, warm: int, full_job: bool = False):
    """The unmodified reference on the host cores: `runs` timed runs of a bounded sample (the first reads up to
    ~target_cand candidate pairs + all contigs) and, with full_job, ONE run of the whole job (same config as the GPU arm)."""
    from links_b200 import synth
    from oracle import oracle_py
    if not oracle_py.have_ref_bin():
        return None
    sample, cand = reference_sample(contigs, reads, k, d, t, target_cand)
    total = synth.candidate_pairs(reads.off, k, list(d), list(t))
    wd = prepare_reference_dir(contigs, sample, reads if full_job else None)
    threads = min(4, os.cpu_count() or 1)  # the reference errors above -j 4 (README.md:45)
    out = {"cand": cand, "threads": threads, "n_reads": sample.n, "bases": int(sample.off[-1]), "total_cand": total,
           "sample_fraction": cand / max(total, 1)}
    try:
        for _ in range(warm):
            run_reference_once(wd, k, d, t, threads)
        res = [run_reference_once(wd, k, d, t, threads) for _ in range(runs)]
        out["secs"] = [x[0] for x in res]
        out["stages_s"] = res[-1][1] if res else {}
        if full_job:
            wall, st = run_reference_once(wd, k, d, t, threads, "reads_full.fof")
            s234 = sum(st.get(x, 0.0) for x in ("reads_s", "contigs_s", "pairing_s"))
            out["full"] = {"candidate_pairs": total, "threads": threads, "host_cores": os.cpu_count(), "wall_s": round(wall, 2), "stages_s": st,
                           "pairs_per_s_wall": total / wall, "pairs_per_s_stage234": total / s234 if s234 else None,
                           "note": "whole job, same config as the GPU arm, one run; stages 2 (merge thread) and 4 (pair_contigs) are "
                                   "single-threaded in the reference whatever -j is (src/LINKS.hpp:354-378,780)"}
    finally:
        shutil.rmtree(wd, ignore_errors=True)
    return out


def main_reference(args, rank: int, world: int):
    if rank != 0:
        return
    contigs, reads, k, d, t = load_workload(args.workload, args.scale)
    full_job = args.workload == "c2_bacterial_5mbp" and args.scale == 1.0 and not args.no_ref_full
    r = cpu_reference(contigs, reads, k, d, t, args.ref_candidates, runs=args.steps, warm=min(args.warmup, 1), full_job=full_job)
    if r is None:
        emit({"impl": "reference", "unavailable": "oracle/_ref/LINKS_CPP not built (needs /root/reference at build time)"})
        return
    ms = 1000.0 * sum(r["secs"]) / len(r["secs"])
    value = r["cand"] / (ms / 1000.0)
    sample = (f"each step: first {r['n_reads']} reads ({r['bases']} bases, {r['cand']} candidate pairs = {r['sample_fraction']:.3f} of the job) "
              f"of {args.workload} + all contigs; unmodified reference LINKS_CPP -j {r['threads']} from FASTA files, process wall time")
    out = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
           "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64",
           "data": "synthetic", "impl": "reference",
           "config": {"workload": args.workload, "scale": args.scale, "k": k, "d": d, "t": t, "sample": sample,
                      "sample_fraction": r["sample_fraction"]},
           "cpu_baseline": {"value": value, "unit": UNIT, "cores": r["threads"], "kind": "reference", "sample": sample,
                            "sample_fraction": r["sample_fraction"], "stages_s_last_sample": r.get("stages_s"),
                            "full_size_same_config": r.get("full")},
           "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
           "host_cores": os.cpu_count()}
    emit(out)


# ---------------------------------------------------------------------------------------------- our arm

L2_BYTES = 126 << 20
GOLDEN_OF = {("c2_bacterial_5mbp", 1.0, (2,)): "full_c2_bacterial_5mbp", ("c5_step_sweep_100mbp", 1.0, (1,)): "full_c5_dense_t1",
             ("c3_celegans_100mbp", 1.0, (20, 10, 5)): "full_c3_celegans_100mbp"}


def _json_or_none(path: str):
    try:
        return json.load(open(path))
    except Exception:
        return None


def hbm_peak():
    mp = _json_or_none(os.path.join(ROOT, "MEASURED_PEAKS.json"))
    if mp and "hbm_gbs" in mp:
        return float(mp["hbm_gbs"]), "MEASURED_PEAKS.json hbm_gbs (burst copy)"
    return 6650.0, "fallback 6.65 TB/s (B200_PROFILING.md)"


def roofline_block(job: Job, stats: dict, tm: dict, steps: int, bf_bytes: int) -> dict:
    """extract_kernel (the dominant kernel) against four yardsticks, all over its mean CUDA-event launch time:
       frac               SURVEY 8(d)'s algorithmic bytes (packed read bases + 32 B per Bloom bit test of the REFERENCE's early-exit
                          loop + 64 B per survivor) / HBM copy peak -- the contract's figure; it counts probes the kernel does not
                          make (the exact member set answers true members) and sectors L2 serves, so it is a model, not traffic
       frac_issued        the sectors the kernel really requested (Bloom words + member-set buckets + pair-set slots + read bases)
       frac_dram          measured DRAM bytes (ncu dram__bytes_read+write per candidate of THIS workload, profiles/extract_traffic.json)
       frac_random_sector issued sectors per second / the measured random-32-byte-gather ceiling of this GPU for a table of the
                          filter's size (tools/gather_peak.cu, profiles/gather_peak.json)"""
    cand = stats["n_candidates"]
    n_launch = max(int(tm["extract_launches"]), 1)
    ms_launch = tm["ms_extract"] / n_launch
    sec_step = tm["ms_extract"] / steps * 1e-3
    closed = job.closed_form_candidates()
    tot = max(sum(closed), 1)
    read_bytes = sum(cand * (c / tot) * (tt / 4.0) for c, tt in zip(closed, job.t))  # per distance: t/4 bytes of packed bases per candidate
    alg = read_bytes + stats["n_bf_probes"] * 32.0 + stats["n_survivors"] * 64.0
    issued_sectors = stats.get("n_bf_loads", 0) + stats.get("n_member_sectors", 0) + 2 * stats["n_survivors"]
    issued = read_bytes + issued_sectors * 32.0
    peak, peak_src = hbm_peak()
    gbs = lambda b: b / sec_step / 1e9 if sec_step > 0 else 0.0
    tr = _json_or_none(os.path.join(ROOT, "profiles", "extract_traffic.json")) or {}
    per_cand = (tr.get("workloads", {}).get(job.name) or {}).get("dram_bytes_per_candidate")
    traffic = per_cand * cand * steps / n_launch if per_cand else None
    gp = _json_or_none(os.path.join(ROOT, "profiles", "gather_peak.json")) or {}
    dram_resident = bf_bytes > (64 << 20)
    ceiling = gp.get("dram_gsectors_per_s" if dram_resident else "l2_gsectors_per_s")
    out = {"bound": "hbm-random-sector+latency" if dram_resident else "l2+latency", "achieved": gbs(alg), "peak": peak, "unit": "GB/s",
           "frac": gbs(alg) / peak, "traffic": traffic,
           "traffic_source": ("profiles/extract_traffic.json: ncu dram__bytes_read+write per candidate of this workload x candidates per launch"
                              if traffic else "no ncu capture of this workload committed"),
           "frac_issued": gbs(issued) / peak, "frac_dram": gbs(per_cand * cand) / peak if per_cand else None,
           "frac_random_sector": (issued_sectors / sec_step / 1e9) / ceiling if ceiling and sec_step > 0 else None,
           "random_sector_ceiling_gsectors_per_s": ceiling, "issued_gsectors_per_s": issued_sectors / sec_step / 1e9 if sec_step > 0 else None,
           "filter_resident_in": "dram" if dram_resident else "l2", "bf_bytes": bf_bytes,
           "kernel": "extract_kernel", "launches": n_launch, "ms_per_launch": ms_launch, "algorithmic_bytes_per_launch": alg * steps / n_launch,
           "bytes_per_pair": alg / max(cand, 1), "issued_bytes_per_pair": issued / max(cand, 1), "peak_source": peak_src,
           "issued_sectors_per_pair": {"bloom": stats.get("n_bf_loads", 0) / max(cand, 1), "member_set": stats.get("n_member_sectors", 0) / max(cand, 1),
                                       "pair_set": 2 * stats["n_survivors"] / max(cand, 1), "reference_order_bloom": stats["n_bf_probes"] / max(cand, 1)}}
    return out


def golden_check(job: Job, L, rows, issues, stats) -> dict | None:
    """Rows / issues / counters of the job against the digests the unmodified reference produced on the same inputs
    (tests/golden/full_*/case.json, tools/make_fullsize_golden.py).  None when no digest exists for this workload."""
    import hashlib
    name = GOLDEN_OF.get((job.name, job.scale, tuple(job.t)))
    meta = _json_or_none(os.path.join(ROOT, "tests", "golden", name, "case.json")) if name else None
    if not meta or job.scaling == "strong" or meta["contigs_file_bytes"] != job.file_bytes or meta["n_reads"] != job.n_reads:
        return None
    tsv = b"".join(sorted(L.format_rows(rows).splitlines(keepends=True)))
    txt = L.format_issues(issues).split(b"\n")
    iss = b"".join(r + b"\n" for r in sorted(b"\n".join(txt[i:i + 4]) for i in range(0, len(txt) - 1, 4)))
    ok = (hashlib.sha256(tsv).hexdigest() == meta["tsv_sha256"] and hashlib.sha256(iss).hexdigest() == meta["issues_sha256"]
          and all(stats[k2] == v for k2, v in meta["stats"].items()))
    return {"golden": name, "match": bool(ok), "rows": len(rows), "reference_rows": meta["n_rows"]}


def measure_job(args, job: Job, lib, dist, rank: int, world: int, local_rank: int, steps: int, warmup: int, e2e: bool = True,
                table: str | None = None) -> dict | None:
    """`steps` timed jobs with the inputs resident in HBM (value) and, with e2e, the same through pinned host buffers.
    Returns the result dict on rank 0 (None elsewhere)."""
    import torch

    from links_b200 import Links, LinksOptions, bf_size, multi
    dev = torch.device("cuda", local_rank)
    _m, bf_bytes, hash_num = bf_size(job.file_bytes, 0.001)
    L = Links(LinksOptions(k=job.k, distances=job.d, step_sizes=job.t, device=local_rank))
    if table:
        L.table_mode(table)
    stream = torch.cuda.current_stream()
    L.set_stream(stream.cuda_stream)
    # N > 1: pairs travel to their owner ranks through NVLink peer memory from inside extract_kernel while the reads are still
    # being extracted (links_exchange_attach); LINKS_EXCHANGE=nccl selects the stop-and-exchange all-to-all instead
    p2p = world > 1 and os.environ.get("LINKS_EXCHANGE", "p2p") != "nccl"
    bloom_mode = os.environ.get("LINKS_BLOOM", "auto")
    bloom_bcast = bloom_mode == "broadcast"
    # a filter of hundreds of MB is DRAM-resident and its build (hash_num random read-modify-writes per contig k-mer) is the
    # longest stage of a human-scale job: built in shares, OR-ed over NVLink (links_bf_alloc_shared / links_bf_allreduce)
    bloom_shared = world > 1 and (bloom_mode == "shared" or (bloom_mode == "auto" and bf_bytes >= (256 << 20)))
    keep_bloom, bloom_ptrs = multi.alloc_shared_bloom(dist, dev, bf_bytes) if bloom_shared else (None, None)
    keep_alive = None
    if p2p:
        # every survivor (before de-duplication) becomes a record in its owner's inbox: ~0.12 per read base at k 15 / 5 % errors
        # (0.21 of the candidates), two halves, 16 or 24 bytes each; sized with 1.5x slack
        rec = 24 if len(job.d) > 1 else 16
        est = int(sum(job.closed_form_candidates()) * 0.3) + (1 << 20)
        keep_alive = multi.attach_exchange(L, dist, dev, inbox_bytes=max(64 << 20, est * rec * 2 * 3 // 2))
    pin = job.pin(lib, args.e2e_read_bytes) if e2e else None
    state = {}
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

    def push_contigs(resident: bool):
        if resident:
            L.push_contigs_dev(job.d_contigs.data_ptr(), job.contigs_off, 1)
        else:
            L.push_contigs_raw(pin["pc_seq"], pin["pc_off"], job.n_contigs, 1)

    def step(resident: bool):
        mark(None)
        L.reset()
        mark("reset")
        if bloom_shared:
            multi.build_bloom_shared(L, bf_bytes, hash_num, bloom_ptrs, lambda: push_contigs(resident), dist)
        elif world > 1 and bloom_bcast:
            # north_star's variant: rank 0 builds the Bloom filter, NCCL broadcast replicates it, the others adopt it.
            # It serialises the ranks (everyone waits for rank 0's build) -- measured 1.7 ms of a 10 ms step at N=2.
            if rank == 0:
                L.bf_alloc(bf_bytes, hash_num)
                push_contigs(resident)
            multi.replicate_bloom(L, bf_bytes, hash_num, dist, dev)
            if rank != 0:
                push_contigs(resident)
        else:
            # default (SURVEY 8e's alternative): the contigs are replicated anyway, so every rank builds the same
            # filter itself (deterministic, bit-identical) and nobody waits
            L.bf_alloc(bf_bytes, hash_num)
            push_contigs(resident)
        mark("bloom+contigs")
        if resident:
            L.push_reads_dev(job.d_reads.data_ptr(), job.reads_off, job.first_read)
        else:
            L.push_reads_raw(pin["pr_seq"], pin["pr_off"], pin["n_reads"], job.first_read)
        mark("reads")
        if p2p:
            L.exchange_finish()
            state["exchange"] = {"transport": "peer memory (NVLink P2P), records written by extract_kernel"}
        elif world > 1:
            state["exchange"] = multi.exchange_pairs(L, dist, dev, profile=phases)
            mark("exchange")
        rows, issues, stats = L.finalize()
        mark("finalize")
        if world > 1 and not p2p:
            rows = multi.gather_rows(rows, dist, dev, merge=L.rows_merge)
            mark("gather_rows")
        state["rows"], state["issues"], state["stats"] = rows, issues, stats

    def timed(resident: bool, n_steps: int, n_warm: int):
        for _ in range(n_warm):
            step(resident)
        L.timing(reset=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(stream)
        for _ in range(n_steps):
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
        return ms / n_steps, L.timing(reset=True)

    def all_sum(v: int) -> int:
        if world == 1:
            return int(v)
        tc = torch.tensor([int(v)], device="cuda", dtype=torch.int64)
        dist.all_reduce(tc)
        return int(tc.item())

    ms_res, tm_res = timed(True, steps, warmup)
    stats = dict(state["stats"])
    rows_res, issues_res = (state["rows"].copy(), state["issues"].copy()) if rank == 0 else (None, None)
    digest_res = rows_digest(rows_res) if rank == 0 else ""
    cand = stats["n_candidates"]
    cand_total = all_sum(cand)
    unique_total = all_sum(stats["n_unique_pairs"])
    surv_total = all_sum(stats["n_survivors"])
    out = None
    if rank == 0:
        n_rows = len(rows_res)
        out = {
            "metric": METRIC, "value": cand_total / (ms_res * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": warmup, "ms_per_step": ms_res, "higher_is_better": True, "scaling": job.scaling, "vs_baseline": None,
            "dtype": "u64", "data": "synthetic",
            "config": {"workload": job.name, "scale": job.scale, "k": job.k, "d": job.d, "t": job.t, "genome_bp": job.genome_bp,
                       "contigs": job.n_contigs, "reads_per_gpu": job.n_reads, "read_bases_per_gpu": int(job.reads_off[-1]),
                       "candidate_pairs_per_gpu": cand, "candidate_pairs": cand_total, "survivors": surv_total, "unique_pairs": unique_total,
                       "bf_probes": stats["n_bf_probes"], "bf_bytes": bf_bytes, "rows": n_rows, "rows_digest": digest_res,
                       "generator": job.note, "contig_table": table or os.environ.get("LINKS_TABLE_MODE", "auto"),
                       "l2": "inputs (reads, pair set, contig table) exceed the 126 MB L2; no flush",
                       "sharding": ((("one read set split into contiguous shards, " if job.scaling == "strong" else "each rank its own read set of the configuration's size, ")
                                     + "contigs replicated, Bloom filter "
                                     + ("built in shares (each rank inserts 1/N of the contig k-mers) and OR-ed over NVLink peer memory, " if bloom_shared else
                                        "NCCL-broadcast from rank 0, " if bloom_bcast else "built by every rank from the replicated contigs (LINKS_BLOOM=broadcast: NCCL broadcast from rank 0), ")
                                     + ("every surviving pair written by extract_kernel itself into its owner rank's inbox (hash of (h_a,h_b)) over NVLink peer "
                                        "memory and imported there while extraction continues, partial link rows summed on rank 0 on the device"
                                        if p2p else "distinct pairs routed by one NCCL all-to-all keyed on hash(h_a,h_b), rows summed on rank 0"))
                                    if world > 1 else "single GPU"),
                       "exchange_rank0": state.get("exchange")},
            "gpu_launches": int(tm_res["launches"]),
            "roofline": roofline_block(job, stats, tm_res, steps, bf_bytes),
            "kernel_ms_per_step": {kk: vv / steps for kk, vv in tm_res.items() if kk.startswith("ms_")},
        }
        if world == 1:
            chk = golden_check(job, L, rows_res, issues_res, stats)
            if chk is not None:
                out["config"]["reference_golden"] = chk
        if args.properties:
            out["config"]["properties"] = (job_properties(rows_res, issues_res, stats, job.reads_off, job.k, job.d, job.t)
                                           if world == 1 else "N=1 only")
    if e2e:
        ms_e2e, tm_e2e = timed(False, steps, max(warmup, 1))
        cand_e2e = all_sum(state["stats"]["n_candidates"])
        if rank == 0:
            if pin["whole_job"]:
                assert digest_res == rows_digest(state["rows"]), "device-resident and host-buffer legs disagree"
            d2h_bytes = len(state["rows"]) * 32 + len(state["issues"]) * 72 + 3 * 128
            out["e2e"] = {"value": cand_e2e / (ms_e2e * 1e-3), "unit": UNIT, "ms_per_step": ms_e2e, "h2d_bytes_per_step": int(pin["h2d_bytes"]),
                          "d2h_bytes_per_step": int(d2h_bytes), "candidate_pairs": cand_e2e,
                          "reads": ("the whole job" if pin["whole_job"] else
                                    f"host-resident slice: the first {pin['n_reads']} reads of each rank's share + all contigs (the full read set is "
                                    "generated on the device and never exists on the host)")}
            out["kernel_ms_per_step_e2e"] = {kk: vv / steps for kk, vv in tm_e2e.items() if kk.startswith("ms_")}
    if phases:
        print(f"[rank {rank}] host ms per phase, last step: " + json.dumps({k2: round(v, 2) for k2, v in ph.items()}), file=sys.stderr)
    L.close()
    del keep_alive, keep_bloom
    return out


def main_ours(args, rank: int, world: int, local_rank: int):
    import torch
    import torch.distributed as dist

    from links_b200 import _abi

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    lib = _abi.load()
    sweep = [int(x) for x in args.sweep.split(",")] if args.sweep else [None]
    clocks = ClockSampler(local_rank)
    clocks.start()
    out = None
    for t_override in sweep:
        job = load_job(args.workload, args.scale, rank, world, dev, coverage=args.coverage, strong=args.strong,
                       steps=[t_override] if t_override else None)
        res = measure_job(args, job, lib, dist, rank, world, local_rank, args.steps, args.warmup, e2e=not args.no_e2e, table=args.table)
        job.unpin(lib)
        if args.sweep:  # one JSON line per point of the step-size sweep (BASELINE.json configs[4])
            if rank == 0:
                res["clocks"] = clocks.stop()
                emit(res)
            clocks = ClockSampler(local_rank)
            clocks.start()
            del job
            continue
        out = res
        # the cached numpy arrays of this job are reused by the CPU baseline below
        host = (job.host_contigs, job.contigs_off, job.host_reads, job.reads_off, job.k, job.d, job.t)
        del job
    if args.sweep:
        clocks.stop()
    else:
        clk = clocks.stop()
        if rank == 0:
            out["clocks"] = clk
        # other regimes, driver-run inside the same line (not the headline): the DRAM-resident filter (C3, N = 1) and the
        # human-scale configuration north_star names for the multi-GPU split (C4, scaled to the ranks)
        extras = [] if args.no_extra or args.workload != "c2_bacterial_5mbp" or args.scale != 1.0 else (
            [("c3_celegans_100mbp", 1.0, None)] if world == 1 else [("c4_human_3gbp", round(world / 8.0, 4), None)])
        for name, scale, tbl in extras:
            torch.cuda.empty_cache()
            try:
                job = load_job(name, scale, rank, world, dev, strong=True)
                res = measure_job(args, job, lib, dist, rank, world, local_rank, steps=2, warmup=1, e2e=False, table=tbl)
                job.unpin(lib)
                del job
            except Exception as e:  # an extra must never take the headline line down with it
                res = {"error": f"{type(e).__name__}: {e}"[:400]}
            if rank == 0:
                out.setdefault("extra", {})[name] = res
        if rank == 0:
            if world == 1 and not args.no_cpu_baseline and host[0] is not None:
                from links_b200.seqio import SeqSet
                contigs, reads = SeqSet(host[0], host[1], []), SeqSet(host[2], host[3], [])
                full_job = args.workload == "c2_bacterial_5mbp" and args.scale == 1.0 and not args.no_ref_full
                r = cpu_reference(contigs, reads, host[4], host[5], host[6], args.ref_candidates, runs=0 if full_job else 1, warm=0, full_job=full_job)
                if r is None:
                    out["cpu_baseline"] = {"value": None, "unit": UNIT, "cores": 0, "kind": "reference", "sample": "oracle/_ref/LINKS_CPP missing"}
                elif full_job:
                    f = r["full"]
                    out["cpu_baseline"] = {"value": f["pairs_per_s_wall"], "unit": UNIT, "cores": r["threads"], "kind": "reference",
                                           "sample": (f"the WHOLE job (same config as the GPU arm: {f['candidate_pairs']} candidate pairs), one run of the unmodified "
                                                      f"reference LINKS_CPP -j {r['threads']} from FASTA files, {f['wall_s']} s process wall; host has {os.cpu_count()} cores"),
                                           "sample_fraction": 1.0, "stages_s": f["stages_s"], "pairs_per_s_stage234": f["pairs_per_s_stage234"], "note": f["note"]}
                else:
                    v = r["cand"] / r["secs"][0]
                    out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": r["threads"], "kind": "reference", "sample_fraction": r["sample_fraction"],
                                           "stages_s": r.get("stages_s"),
                                           "sample": f"first {r['n_reads']} reads ({r['cand']} candidate pairs) + all contigs, unmodified "
                                                     f"reference LINKS_CPP -j {r['threads']} from FASTA, {r['secs'][0]:.1f} s wall; host has {os.cpu_count()} cores"}
            emit(out)
    if world > 1:
        dist.destroy_process_group()


def check_sharding(args, rank: int, world: int, local_rank: int):
    """Strong-scaling parity: the SAME reads split over the ranks (contiguous shards, global read
    ordinals) must give, after the pair exchange and the row merge, exactly the rows one GPU
    computes for the whole job.  Prints one JSON line on rank 0; not a timing run."""
    import torch
    import torch.distributed as dist

    from links_b200 import Links, LinksOptions, bf_size, multi
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    job = load_job(args.workload, args.scale, rank, world, dev, coverage=args.coverage, strong=True)
    _m, bf_bytes, hash_num = bf_size(job.file_bytes, 0.001)
    with Links(LinksOptions(k=job.k, distances=job.d, step_sizes=job.t, device=local_rank)) as L:
        if args.table:
            L.table_mode(args.table)
        p2p = world > 1 and os.environ.get("LINKS_EXCHANGE", "p2p") != "nccl"
        rec = 24 if len(job.d) > 1 else 16
        est = int(sum(job.closed_form_candidates()) * 0.3) + (1 << 20)
        keep_alive = multi.attach_exchange(L, dist, dev, inbox_bytes=max(64 << 20, est * rec * 2 * 3 // 2)) if p2p else None
        bloom_mode = os.environ.get("LINKS_BLOOM", "broadcast")
        keep_bloom = None
        if world > 1 and bloom_mode == "shared":  # every rank inserts 1/N of the contig k-mers, OR over peer memory
            keep_bloom, bloom_ptrs = multi.alloc_shared_bloom(dist, dev, bf_bytes)
            multi.build_bloom_shared(L, bf_bytes, hash_num, bloom_ptrs, lambda: L.push_contigs_dev(job.d_contigs.data_ptr(), job.contigs_off, 1), dist)
        else:
            if rank == 0:
                L.bf_alloc(bf_bytes, hash_num)
                L.push_contigs_dev(job.d_contigs.data_ptr(), job.contigs_off, 1)
            if world > 1:
                multi.replicate_bloom(L, bf_bytes, hash_num, dist, dev)
                if rank != 0:
                    L.push_contigs_dev(job.d_contigs.data_ptr(), job.contigs_off, 1)
        L.push_reads_dev(job.d_reads.data_ptr(), job.reads_off, job.first_read)
        if p2p:
            L.exchange_finish()
            ex = {"transport": "peer memory"}
        else:
            ex = multi.exchange_pairs(L, dist, dev) if world > 1 else {}
        rows, issues, stats = L.finalize()
        if world > 1 and not p2p:
            rows = multi.gather_rows(rows, dist, dev)
        rows = rows.copy() if rows is not None else None
        del keep_alive, keep_bloom
    tot = torch.tensor([stats["n_candidates"], stats["n_unique_pairs"], len(issues)], dtype=torch.int64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    if rank == 0:
        del job
        torch.cuda.empty_cache()
        whole = load_job(args.workload, args.scale, 0, 1, dev, coverage=args.coverage, strong=True)
        with Links(LinksOptions(k=whole.k, distances=whole.d, step_sizes=whole.t, device=local_rank)) as L:
            # the single-GPU run takes the other build of the contig table where both fit, so the check also crosses the two
            if args.table_single:
                L.table_mode(args.table_single)
            L.bf_alloc(bf_bytes, hash_num)
            L.push_contigs_dev(whole.d_contigs.data_ptr(), whole.contigs_off, 1)
            L.push_reads_dev(whole.d_reads.data_ptr(), whole.reads_off, 0)
            rows1, issues1, stats1 = L.finalize()
            rows1 = rows1.copy()
        ok = (rows_digest(rows) == rows_digest(rows1) and int(tot[0]) == stats1["n_candidates"]
              and int(tot[1]) == stats1["n_unique_pairs"] and int(tot[2]) == len(issues1))
        emit({"check_sharding": "ok" if ok else "MISMATCH", "n_gpus": world, "workload": args.workload, "scale": args.scale,
              "contig_table": [args.table or "auto", args.table_single or "auto"], "bloom": os.environ.get("LINKS_BLOOM", "broadcast"),
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
    ap.add_argument("--no-ref-full", action="store_true", help="skip the one whole-job run of the reference (C2: about 80 s of CPU)")
    ap.add_argument("--no-e2e", action="store_true", help="skip the host-buffer leg")
    ap.add_argument("--no-extra", action="store_true", help="skip the extra regimes appended to the default line (C3 at N=1, C4 at N>1)")
    ap.add_argument("--coverage", type=float, default=None, help="override the workload's read coverage (numpy workloads)")
    ap.add_argument("--strong", action="store_true", help="N>1: split ONE read set over the ranks instead of one set per rank")
    ap.add_argument("--sweep", default="", help="comma-separated step sizes: one JSON line per -t value on the same inputs (configs[4])")
    ap.add_argument("--table", default=None, choices=["auto", "full", "mates"], help="contig table build (links_gpu.h LINKS_TABLE_*)")
    ap.add_argument("--table-single", default=None, choices=["auto", "full", "mates"], help="--check-sharding: table build of the single-GPU run")
    ap.add_argument("--e2e-read-bytes", type=int, default=4 << 30, help="largest host-resident read slice of the end-to-end leg")
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
