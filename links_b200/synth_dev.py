"""Device-side synthetic inputs for workloads the numpy generator (synth.py) cannot make in reasonable time:
BASELINE.json configs[3], a 3.1 Gbp human-scale draft with 20x 20 kb reads (62 Gbases), `-k 20 -d 5000 -t 5`.

Bench / test infrastructure over tools/synthgen/libsynth_gpu.so (its own small CUDA library, not the product ABI).
Same shapes and error model as synth.py, from a counter-based generator: the layout of the job
(repeat loci, contig cuts, read starts / lengths / strands) is drawn on the host with seeded numpy and is a few MB; the
bases themselves are pure functions of (seed, index) computed on the GPU.  Any rank can therefore generate any
contiguous slice of the read set, and N ranks together see exactly the reads one rank would (strong scaling, and
`bench.py --check-sharding` at any N).
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(HERE), "tools", "synthgen", "libsynth_gpu.so")

# name: (genome bp, n contigs, coverage, mean read len, k, distances, steps)
CONFIGS = {
    "c4_human_3gbp": (3_100_000_000, 30_000, 20.0, 20_000, 20, [5000], [5]),
}

_lib = None


def load() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} is missing: build it with `python -m links_b200.build`")
        lib = C.CDLL(LIB_PATH)
        P, U64, U32 = C.c_void_p, C.c_uint64, C.c_uint32
        lib.synth_genome.argtypes = [P, U64, U64]
        lib.synth_copy_ranges.argtypes = [P, P, P, P, P, P, U32]
        lib.synth_fill_runs.argtypes = [P, P, P, U32, C.c_uint8]
        lib.synth_read_lengths.argtypes = [P, U32, U64, U64, C.c_double, P]
        lib.synth_reads.argtypes = [P, P, P, P, P, U32, U64, U64, C.c_double, C.c_double, P]
        for f in ("synth_genome", "synth_copy_ranges", "synth_fill_runs", "synth_read_lengths", "synth_reads"):
            getattr(lib, f).restype = C.c_int
        _lib = lib
    return _lib


def _ck(rc: int, what: str) -> None:
    if rc:
        raise RuntimeError(f"{what}: CUDA error {rc}")


@dataclass
class DevWorkload:
    """One rank's share of a job.  d_contigs / d_reads are torch uint8 tensors on the device; the offsets are host arrays
    (what links_contigs_push_dev / links_reads_push_dev take)."""
    name: str
    scale: float
    k: int
    d: list
    t: list
    genome_bp: int
    d_contigs: object
    contigs_off: np.ndarray
    contigs_file_bytes: int
    d_reads: object
    reads_off: np.ndarray       # this rank's reads, offsets starting at 0
    first_read: int             # global ordinal of this rank's first read
    n_reads_total: int
    read_bases_total_source: int  # sum of the source lengths of ALL reads of the job (before errors)


def _dev(a: np.ndarray, device):
    import torch
    return torch.from_numpy(np.ascontiguousarray(a)).to(device)


def job_layout(name: str, scale: float):
    """Host-side (numpy, seeded) layout of the whole job: repeats, contig cuts, read placements."""
    g, nc, cov, rl, k, d, t = CONFIGS[name]
    G = int(g * scale)
    nc = max(int(nc * scale), 2)
    rng = np.random.default_rng(41)
    # ~5 % of the genome in 2-5 kb repeats copied to random loci
    n_rep = int(G * 0.05 / 3500) if G > 12000 else 0
    rep_len = rng.integers(2000, 5001, size=n_rep).astype(np.uint64)
    rep_src = (rng.random(n_rep) * (G - rep_len.astype(np.float64))).astype(np.uint64)
    rep_dst = (rng.random(n_rep) * (G - rep_len.astype(np.float64))).astype(np.uint64)
    if n_rep:
        # destinations must not overlap: two copies landing on the same bytes would race on the device and the genome would
        # differ from one generation to the next (every rank generates its own copy and they must be identical)
        order = np.argsort(rep_dst, kind="stable")
        rep_dst, rep_src, rep_len = rep_dst[order], rep_src[order], rep_len[order]
        keep = np.ones(n_rep, dtype=bool)
        end = 0
        for i in range(n_rep):
            if int(rep_dst[i]) < end:
                keep[i] = False
            else:
                end = int(rep_dst[i]) + int(rep_len[i])
        rep_dst, rep_src, rep_len = rep_dst[keep], rep_src[keep], rep_len[keep]
    # contigs: exponential lengths around the mean, 1 % short, a 0-2 kb gap between neighbours
    rng = np.random.default_rng(43)
    mean_len = max(G / nc - 1000, 200)
    m = int(nc * 1.3) + 16
    lens = np.maximum(rng.exponential(mean_len, size=m), 200).astype(np.int64)
    short = rng.random(m) < 0.01
    lens[short] = rng.integers(20, 500, size=int(short.sum()))
    gaps = rng.integers(0, 2001, size=m)
    starts = np.concatenate([[0], np.cumsum(lens + gaps)[:-1]])
    keep = starts < G
    starts, lens = starts[keep], lens[keep]
    lens = np.minimum(lens, G - starts)
    ok = lens > 0
    starts, lens = starts[ok].astype(np.uint64), lens[ok].astype(np.uint64)
    c_rc = (rng.random(len(lens)) < 0.5).astype(np.uint8)
    has_n = (rng.random(len(lens)) < 0.01) & (lens > 400)
    n_idx = np.flatnonzero(has_n)
    n_pos_in = (100 + rng.random(len(n_idx)) * (lens[n_idx].astype(np.float64) - 200)).astype(np.uint64)
    n_len = rng.integers(1, 60, size=len(n_idx)).astype(np.uint32)
    # reads: uniform starts, log-normal lengths around the mean, random strand
    rng = np.random.default_rng(42)
    sigma = 0.35
    mu = np.log(rl) - sigma * sigma / 2
    n_reads = int(np.ceil(G * cov / rl))
    r_len = np.clip(rng.lognormal(mu, sigma, size=n_reads), 200, G).astype(np.uint64)
    r_start = (rng.random(n_reads) * (G - r_len.astype(np.float64) + 1)).astype(np.uint64)
    r_start = np.minimum(r_start, np.uint64(G) - r_len)
    r_rc = (rng.random(n_reads) < 0.5).astype(np.uint8)
    return dict(G=G, k=k, d=list(d), t=list(t), rep=(rep_dst, rep_src, rep_len), contigs=(starts, lens, c_rc, n_idx, n_pos_in, n_len),
                reads=(r_start, r_len, r_rc))


def make_config_dev(name: str, scale: float, device, rank: int = 0, world: int = 1, read_range: tuple[int, int] | None = None,
                    seed: int = 7, err: float = 0.05, n_read_frac: float = 0.001) -> DevWorkload:
    """Generate the contigs and reads [lo, hi) of the job on `device` (default: rank's contiguous 1/world share)."""
    import torch
    lib = load()
    lay = job_layout(name, scale)
    G = lay["G"]
    with torch.cuda.device(device):
        genome = torch.empty(G, dtype=torch.uint8, device=device)
        _ck(lib.synth_genome(genome.data_ptr(), G, seed), "synth_genome")
        rep_dst, rep_src, rep_len = lay["rep"]
        if len(rep_len):
            soff = np.concatenate([[0], np.cumsum(rep_len)]).astype(np.uint64)
            scratch = torch.empty(int(soff[-1]), dtype=torch.uint8, device=device)
            d_soff, d_src, d_dst, d_len = _dev(soff[:-1], device), _dev(rep_src, device), _dev(rep_dst, device), _dev(rep_len, device)
            _ck(lib.synth_copy_ranges(genome.data_ptr(), scratch.data_ptr(), d_soff.data_ptr(), d_src.data_ptr(), d_len.data_ptr(), None, len(rep_len)), "repeats gather")
            _ck(lib.synth_copy_ranges(scratch.data_ptr(), genome.data_ptr(), d_dst.data_ptr(), d_soff.data_ptr(), d_len.data_ptr(), None, len(rep_len)), "repeats scatter")
            del scratch
        c_start, c_len, c_rc, n_idx, n_pos_in, n_len = lay["contigs"]
        c_off = np.concatenate([[0], np.cumsum(c_len)]).astype(np.uint64)
        d_contigs = torch.empty(int(c_off[-1]), dtype=torch.uint8, device=device)
        # (device copies of the small arrays are held in names: a temporary would be freed, and its block reused, mid-call)
        t_off, t_start, t_len, t_rc = _dev(c_off[:-1], device), _dev(c_start, device), _dev(c_len, device), _dev(c_rc, device)
        _ck(lib.synth_copy_ranges(genome.data_ptr(), d_contigs.data_ptr(), t_off.data_ptr(), t_start.data_ptr(), t_len.data_ptr(),
                                  t_rc.data_ptr(), len(c_len)), "contigs")
        if len(n_idx):
            pos = (c_off[n_idx] + n_pos_in).astype(np.uint64)
            ln = np.minimum(n_len.astype(np.uint64), c_off[n_idx + 1] - pos).astype(np.uint32)
            t_pos, t_ln = _dev(pos, device), _dev(ln, device)
            _ck(lib.synth_fill_runs(d_contigs.data_ptr(), t_pos.data_ptr(), t_ln.data_ptr(), len(pos), ord("N")), "N runs")
        n = len(c_len)
        digits = np.floor(np.log10(np.arange(1, n + 1))).astype(np.int64) + 1
        file_bytes = int(c_off[-1]) + int(digits.sum()) + 3 * n  # ">" + ordinal + "\n" + sequence + "\n"
        r_start, r_len, r_rc = lay["reads"]
        R = len(r_len)
        lo, hi = read_range if read_range is not None else (R * rank // world, R * (rank + 1) // world)
        m = hi - lo
        d_rstart, d_rlen, d_rrc = _dev(r_start[lo:hi], device), _dev(r_len[lo:hi], device), _dev(r_rc[lo:hi], device)
        d_olen = torch.empty(max(m, 1), dtype=torch.int64, device=device)
        _ck(lib.synth_read_lengths(d_rlen.data_ptr(), m, lo, seed + 1, err, d_olen.data_ptr()), "read lengths")
        r_off = np.concatenate([[0], np.cumsum(d_olen[:m].cpu().numpy().astype(np.uint64))]).astype(np.uint64)
        d_reads = torch.empty(max(int(r_off[-1]), 1), dtype=torch.uint8, device=device)
        t_roff = _dev(r_off, device)
        _ck(lib.synth_reads(genome.data_ptr(), d_rstart.data_ptr(), d_rlen.data_ptr(), d_rrc.data_ptr(), t_roff.data_ptr(), m, lo,
                            seed + 1, err, n_read_frac, d_reads.data_ptr()), "reads")
        del genome
        torch.cuda.empty_cache()
    return DevWorkload(name=name, scale=scale, k=lay["k"], d=lay["d"], t=lay["t"], genome_bp=G, d_contigs=d_contigs, contigs_off=c_off,
                       contigs_file_bytes=file_bytes, d_reads=d_reads, reads_off=r_off, first_read=lo, n_reads_total=R,
                       read_bases_total_source=int(r_len.sum()))
