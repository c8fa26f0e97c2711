"""GPU parity: the CUDA path, called through the C-ABI, against the CPU oracle and the committed
golden fixtures.  Integer / byte / index work: every comparison is bit-exact."""
from __future__ import annotations

import os

import numpy as np
import pytest

from .conftest import GOLDEN as GOLDEN_DIR
from .conftest import GoldenCase, sha256, sort_issue_records, sort_lines

pytestmark = pytest.mark.gpu


def _links(k=15, d=(4000,), t=(2,), o=0, e=0.1):
    from links_b200 import Links, LinksOptions
    return Links(LinksOptions(k=k, distances=list(d), step_sizes=list(t), offset=o, insert_stdev=e))


def _dirty(rng, n, alphabet=b"ACGTACGTACGTACGTacgtNnRYUu-."):
    a = np.frombuffer(alphabet, np.uint8)
    return a[rng.integers(0, len(a), n)].tobytes()


def test_nthash_known_answer_on_device(gpu_lib):
    # reference btllib/tests/nthash.cpp:10-28
    kmer = b"ACGTACACTGGACTGAGTCT"
    with _links(k=20) as L:
        fwd, rev, valid = L.kmer_hashes(kmer)
    assert valid.tolist() == [1]
    assert min(int(fwd[0]), int(rev[0])) == 10434435546371013747


@pytest.mark.parametrize("k", [1, 2, 15, 16, 17, 20, 31, 32, 33, 40, 64, 100])
def test_kmer_hashes_match_oracle(gpu_lib, oracle, k):
    rng = np.random.default_rng(k)
    for n in (0, k - 1, k, k + 1, 257, 5000):
        if n < 0:
            continue
        seq = _dirty(rng, n) if n != 257 else _dirty(rng, n, b"ACGT")
        with _links(k=k) as L:
            fwd, rev, valid = L.kmer_hashes(seq)
        pos, ofwd, orev, _ = oracle.nthash_walk(seq, 1, k)
        assert np.array_equal(np.flatnonzero(valid), pos.astype(np.int64))
        assert np.array_equal(fwd[valid == 1], ofwd) and np.array_equal(rev[valid == 1], orev)


def test_bloom_build_matches_oracle_and_golden(gpu_lib, oracle, golden_case):
    c = golden_case
    from links_b200 import bf_header, bf_size
    contigs = c.contigs()
    m, nbytes, h = bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    with _links(k=c.opts["k"]) as L:
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        bits = L.bf_bits()
    assert np.array_equal(bits, oracle.bf_build(contigs.seq, contigs.off, nbytes, h, c.opts["k"]))
    assert sha256(bf_header(c.opts["k"], h, nbytes) + bits.tobytes()) == c.meta["bloom_sha256"]


def test_bloom_contains_matches_oracle(gpu_lib, oracle):
    rng = np.random.default_rng(5)
    from links_b200.seqio import SeqSet
    s = SeqSet.from_records([_dirty(rng, 4000, b"ACGT") for _ in range(3)])
    k, nbytes, h = 15, 4096, 5
    bits = oracle.bf_build(s.seq, s.off, nbytes, h, k)
    fwd = rng.integers(0, 2**63, 3000, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, 3000, dtype=np.uint64)
    rev = rng.integers(0, 2**63, 3000, dtype=np.uint64) * np.uint64(2) + rng.integers(0, 2, 3000, dtype=np.uint64)
    _p, f2, r2, _h = oracle.nthash_walk(s.record(0), 1, k)
    fwd = np.concatenate([fwd, f2[:1000]]); rev = np.concatenate([rev, r2[:1000]])
    with _links(k=k) as L:
        L.bf_load(bits, h)
        got = L.bf_contains(fwd, rev)
    exp = oracle.bf_contains(bits, h, k, fwd, rev)
    assert np.array_equal(got, exp)
    assert exp[-1000:].all() and not exp[:3000].all()


def _run_gpu(c: GoldenCase, chunk_reads: int | None = None, table: str | None = None):
    from links_b200 import bf_size
    from links_b200.seqio import SeqSet
    contigs, reads = c.contigs(), c.reads()
    m, nbytes, h = bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    with _links(k=c.opts["k"], d=c.opts["d"], t=c.opts["t"], o=c.opts["o"], e=c.opts["e"]) as L:
        if table:
            L.table_mode(table)
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        if chunk_reads is None:
            L.push_reads(reads)
        else:
            for i in range(0, reads.n, chunk_reads):
                j = min(i + chunk_reads, reads.n)
                sub = SeqSet(reads.seq, reads.off[i:j + 1], reads.names[i:j])
                L.push_reads(sub, first_read_idx=i)
        pa, pb, pd = L.pairs()
        rows, issues, stats = L.finalize()
        return L.format_rows(rows), L.format_issues(issues), stats, (pa, pb, pd)


def test_whole_path_matches_reference_goldens(gpu_lib, golden_case):
    """.tigpair_checkpoint.tsv and .pairing_issues vs the unmodified reference binary's output."""
    tsv, issues, stats, _ = _run_gpu(golden_case)
    assert sort_lines(tsv) == golden_case.expected_tsv()
    assert sort_issue_records(issues) == golden_case.expected_issues()
    assert stats["n_rows"] == golden_case.meta["n_rows"]


def test_mates_table_matches_reference_goldens(gpu_lib, golden_case):
    """The contig table restricted to the hashes that occur in a pair (LINKS_TABLE_MATES = the reference's own rule,
    src/LINKS.hpp:580,591; what makes a 3.1 Gbp assembly fit) against the same goldens, one push and many."""
    for chunk in (None, 3):
        tsv, issues, stats, _ = _run_gpu(golden_case, chunk_reads=chunk, table="mates")
        assert sort_lines(tsv) == golden_case.expected_tsv()
        assert sort_issue_records(issues) == golden_case.expected_issues()
        assert stats["n_rows"] == golden_case.meta["n_rows"]


def test_member_map_matches_reference_goldens(gpu_lib, golden_case, monkeypatch):
    """The Bloom filter asked once per sampled read position (member_map.cuh) instead of once per candidate pair: forced on
    (LINKS_MEMBER_MAP=1) and off, both against the reference's output; the count of bit tests the reference's own loop makes
    (n_bf_probes) must come out the same either way.  Reads with an invalid base keep the per-candidate path inside the
    same launch (adv_ns_lower mixes both)."""
    probes = {}
    for mode in ("1", "0"):
        monkeypatch.setenv("LINKS_MEMBER_MAP", mode)
        for chunk in (None, 2):
            tsv, issues, stats, _ = _run_gpu(golden_case, chunk_reads=chunk)
            assert sort_lines(tsv) == golden_case.expected_tsv(), (mode, chunk)
            assert sort_issue_records(issues) == golden_case.expected_issues(), (mode, chunk)
            probes[(mode, chunk)] = (stats["n_bf_probes"], stats["n_candidates"], stats["n_survivors"], stats["n_unique_pairs"])
    assert len(set(probes.values())) == 1, probes


@pytest.mark.parametrize("table", ["full", "mates"])
def test_baseline_configs_match_reference(gpu_lib, config_case, table):
    """BASELINE.json configs[2..4] (multi-distance with per-distance steps; k 20 d 5000 t 5; the step sweep
    t = 1 ... 100 at d 10000) at a size the unmodified reference ran here: .bloom, tsv and issues bit-exact,
    with either build of the contig table."""
    from links_b200 import bf_size
    c = config_case
    contigs, reads = c.inputs()
    o = c.opts
    _m, nbytes, h = bf_size(c.meta["contigs_file_bytes"], o["p"])
    with _links(k=o["k"], d=o["d"], t=o["t"], o=o["o"], e=o["e"]) as L:
        L.table_mode(table)
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        L.push_reads(reads)
        bits = L.bf_bits()
        rows, issues, stats = L.finalize()
        tsv, iss = L.format_rows(rows), L.format_issues(issues)
    assert sort_lines(tsv) == c.expected_tsv()
    assert sort_issue_records(iss) == c.expected_issues()
    assert stats["n_rows"] == c.meta["n_rows"]
    from links_b200 import bf_header
    assert sha256(bf_header(o["k"], h, nbytes) + bits.tobytes()) == c.meta["bloom_sha256"]


def test_intermediates_match_oracle(gpu_lib, oracle, golden_case):
    """Candidate / survivor counts, the distinct pair set with its first-wins distance index, and the
    stage-4 counters, against the oracle on the same inputs."""
    c = golden_case
    contigs, reads = c.contigs(), c.reads()
    _m, nbytes, h = oracle.bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, c.opts["k"])
    _rows, _iss, ostats, obs = oracle.run(c.opts["k"], c.opts["d"], c.steps(), c.opts["o"], c.opts["e"], bits, h,
                                          contigs.seq, contigs.off, reads.seq, reads.off, keep_obs=True)
    _tsv, _issues, stats, (pa, pb, pd) = _run_gpu(c)
    for key in ("n_candidates", "n_survivors", "n_unique_pairs", "ct_ok_pairs", "ct_problem_pairs", "ct_ok_contig",
                "ct_iz_issues", "ct_illogical", "ct_gate_fail"):
        assert stats[key] == ostats[key], key
    # first observation of each (h_a,h_b) in -j 1 order wins (reference src/LINKS.hpp:370-373)
    first: dict[tuple[int, int], int] = {}
    for a, b, d in zip(obs["h_a"].tolist(), obs["h_b"].tolist(), obs["d_idx"].tolist()):
        first.setdefault((a, b), d)
    got = dict(zip(zip(pa.tolist(), pb.tolist()), pd.tolist()))
    assert got == first


@pytest.mark.parametrize("name,chunk", [("adv_tandem", 7), ("adv_ns_lower", 1), ("adv_offset_k20", 13)])
def test_push_granularity_does_not_matter(gpu_lib, name, chunk):
    """Many small links_reads_push calls == one big call (first-wins uses global read ordinals)."""
    c = GoldenCase(name)
    tsv, issues, _stats, _ = _run_gpu(c, chunk_reads=chunk)
    assert sort_lines(tsv) == c.expected_tsv()
    assert sort_issue_records(issues) == c.expected_issues()


def test_bf_load_path_equals_build_path(gpu_lib, oracle):
    """-r: a filter loaded from bytes gives the same rows as one built on the device."""
    c = GoldenCase("adv_offset_k20")
    contigs, reads = c.contigs(), c.reads()
    _m, nbytes, h = oracle.bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, c.opts["k"])
    with _links(k=c.opts["k"], d=c.opts["d"], t=c.opts["t"], o=c.opts["o"], e=c.opts["e"]) as L:
        L.bf_load(bits, h)
        L.push_contigs(contigs)
        L.push_reads(reads)
        rows, issues, _ = L.finalize()
        assert sort_lines(L.format_rows(rows)) == c.expected_tsv()


@pytest.mark.parametrize("name", ["adv_offset_k20", "adv_tandem", "c1"])
def test_member_set_modes_agree(gpu_lib, oracle, name, monkeypatch):
    """The exact member set beside the filter (links_bf_covers_contigs, kernels.cuh MemberView) is a
    shortcut, not a semantic: built with the filter (bf_alloc), built alone under a loaded filter the
    caller vouches for, absent under a plain loaded filter, and regrown across many small contig
    pushes -- same rows, same survivors, same reference-order probe count; only the issued loads differ.
    (Per-candidate filter walks only: the member map, which asks once per position, has its own test and would
    make the issued loads fewer than the reference's count.)"""
    monkeypatch.setenv("LINKS_MEMBER_MAP", "0")
    from links_b200.seqio import SeqSet
    c = GoldenCase(name)
    contigs, reads = c.contigs(), c.reads()
    _m, nbytes, h = oracle.bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    bits = oracle.bf_build(contigs.seq, contigs.off, nbytes, h, c.opts["k"])

    def run(mode):
        with _links(k=c.opts["k"], d=c.opts["d"], t=c.steps(), o=c.opts["o"], e=c.opts["e"]) as L:
            if mode in ("build", "build_batches"):
                L.bf_alloc(nbytes, h)
            else:
                L.bf_load(bits, h)
                if mode == "load_vouched":
                    L.bf_covers_contigs(True)
            if mode == "build_batches":  # one contig per push: the set is re-sized and re-filled on the way
                for i in range(contigs.n):
                    L.push_contigs(SeqSet(contigs.seq, contigs.off[i:i + 2], []), first_tig_id=i + 1)
            else:
                L.push_contigs(contigs)
            L.push_reads(reads)
            got_bits = L.bf_bits()
            rows, issues, stats = L.finalize()
            return L.format_rows(rows), L.format_issues(issues), stats, got_bits

    out = {m: run(m) for m in ("build", "load_plain", "load_vouched", "build_batches")}
    for m, (tsv, iss, stats, got_bits) in out.items():
        assert sort_lines(tsv) == c.expected_tsv(), m
        assert sort_issue_records(iss) == c.expected_issues(), m
        assert np.array_equal(got_bits, bits), m
        for key in ("n_candidates", "n_survivors", "n_unique_pairs", "n_bf_probes"):
            assert stats[key] == out["build"][2][key], (m, key)
    assert out["load_plain"][2]["n_member_sectors"] == 0
    assert out["load_plain"][2]["n_bf_loads"] >= out["load_plain"][2]["n_bf_probes"]
    for m in ("build", "load_vouched", "build_batches"):
        assert out[m][2]["n_member_sectors"] > 0, m
        assert out[m][2]["n_bf_loads"] < out["load_plain"][2]["n_bf_loads"], m


def test_stage_api_writes_reference_files(gpu_lib, tmp_path):
    """The four stage methods of the reference's class LINKS, file in -> file out."""
    from links_b200 import Links, LinksOptions
    c = GoldenCase("c1")
    cpath, fof = c.materialize(str(tmp_path))
    base = str(tmp_path / "installation_test")
    L = Links(LinksOptions(assembly_file=cpath, fof_file=fof, distances=c.opts["d"], step_sizes=c.opts["t"], k=15,
                           base_name=base))
    L.init_bloom_filter()
    L.start_read_fasta()
    L.start_read_contig()
    L.pair_contigs()
    L.close()
    assert sha256(open(base + ".bloom", "rb").read()) == c.meta["bloom_sha256"]
    assert sort_lines(open(base + ".tigpair_checkpoint.tsv", "rb").read()) == c.expected_tsv()
    assert sort_issue_records(open(base + ".pairing_issues", "rb").read()) == c.expected_issues()
    # -r: reuse the .bloom just written (reference src/LINKS.hpp:286-295)
    base2 = str(tmp_path / "again")
    L = Links(LinksOptions(assembly_file=cpath, fof_file=fof, distances=c.opts["d"], step_sizes=c.opts["t"], k=15,
                           base_name=base2, bf_file=base + ".bloom"))
    L.init_bloom_filter(); L.start_read_fasta(); L.start_read_contig(); L.pair_contigs(); L.close()
    assert sort_lines(open(base2 + ".tigpair_checkpoint.tsv", "rb").read()) == c.expected_tsv()


def test_synthetic_tiny_matches_oracle(gpu_lib, oracle):
    """ONT-like synthetic data of the bench's shape at a size the oracle finishes in seconds."""
    from links_b200 import bf_size, synth
    _g, contigs, reads, k, d, t = synth.make_config("tiny")
    _m, nbytes, h = bf_size(len(contigs.seq) + 10 * contigs.n, 0.001)
    with _links(k=k, d=d, t=t) as L:
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        L.push_reads(reads)
        bits = L.bf_bits()
        rows, issues, stats = L.finalize()
        tsv, iss = L.format_rows(rows), L.format_issues(issues)
    assert np.array_equal(bits, oracle.bf_build(contigs.seq, contigs.off, nbytes, h, k))
    orows, oiss, ostats, _ = oracle.run(k, d, t, 0, 0.1, bits, h, contigs.seq, contigs.off, reads.seq, reads.off)
    assert tsv == oracle.format_rows(orows)
    assert iss == oracle.format_issues(oiss)
    assert stats["n_candidates"] == ostats["n_candidates"]
    assert stats["n_survivors"] == ostats["n_survivors"] and stats["n_unique_pairs"] == ostats["n_unique_pairs"]
    assert len(rows) > 0


@pytest.mark.parametrize("table", ["auto", "mates"])
def test_fuzz_jobs_match_oracle(gpu_lib, oracle, table):
    """Random small jobs from tools/oracle_fuzz.py's generator (dirty contigs and reads, k 8..40, one to four distances in any
    order with their own steps, offsets, both tolerances): the oracle agreed with the unmodified reference on 420 such jobs
    in the build container; here the CUDA path must agree with the oracle on a seeded two dozen -- filter bits, rows, issues
    and counters."""
    import sys
    from .conftest import ROOT
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    from oracle_fuzz import make_case
    from links_b200 import bf_size
    n_rows = 0
    for seed in range(900, 912) if table == "auto" else range(912, 924):
        contigs, reads, o = make_case(seed)
        file_bytes = sum(len(contigs.names[i]) + 2 + int(contigs.off[i + 1] - contigs.off[i]) + 1 for i in range(contigs.n))
        _m, nbytes, h = bf_size(file_bytes, o["p"])
        with _links(k=o["k"], d=o["d"], t=o["t"], o=o["o"], e=o["e"]) as L:
            L.table_mode(table)
            L.bf_alloc(nbytes, h)
            L.push_contigs(contigs)
            L.push_reads(reads)
            bits = L.bf_bits()
            rows, issues, stats = L.finalize()
            tsv, iss = L.format_rows(rows), L.format_issues(issues)
        assert np.array_equal(bits, oracle.bf_build(contigs.seq, contigs.off, nbytes, h, o["k"])), (seed, o)
        orows, oiss, ostats, _ = oracle.run(o["k"], o["d"], o["t"], o["o"], o["e"], bits, h, contigs.seq, contigs.off, reads.seq, reads.off)
        assert sort_lines(tsv) == sort_lines(oracle.format_rows(orows)), (seed, o)
        assert sort_issue_records(iss) == sort_issue_records(oracle.format_issues(oiss)), (seed, o)
        for key in ("n_candidates", "n_survivors", "n_unique_pairs"):
            assert stats[key] == ostats[key], (seed, o, key)
        n_rows += len(rows)
    assert n_rows > 0


def test_edge_inputs(gpu_lib, oracle):
    """Empty pushes, reads shorter than k or than k+d, all-N reads, contigs shorter than k."""
    from links_b200.seqio import SeqSet
    contigs = SeqSet.from_records([b"ACGT", b"", b"N" * 50, b"ACGTTGCATGCATGCAAGGCTTAACCGGTTAC" * 20])
    reads = SeqSet.from_records([b"", b"ACG", b"N" * 3000, b"ACGTTGCATGCATGCAAGGCTTAACCGGTTAC" * 5, b"acgtn" * 400])
    k, d, t = 15, [100, 5000], [1, 3]
    nbytes, h = 1024, 4
    with _links(k=k, d=d, t=t) as L:
        L.bf_alloc(nbytes, h)
        L.push_contigs(SeqSet.from_records([]))
        L.push_contigs(contigs)
        L.push_reads(SeqSet.from_records([]))
        L.push_reads(reads)
        bits = L.bf_bits()
        rows, issues, stats = L.finalize()
        tsv, iss = L.format_rows(rows), L.format_issues(issues)
    assert np.array_equal(bits, oracle.bf_build(contigs.seq, contigs.off, nbytes, h, k))
    orows, oiss, ostats, _ = oracle.run(k, d, t, 0, 0.1, bits, h, contigs.seq, contigs.off, reads.seq, reads.off)
    assert tsv == oracle.format_rows(orows) and iss == oracle.format_issues(oiss)
    assert stats["n_candidates"] == ostats["n_candidates"]
    # nothing pushed at all
    with _links() as L:
        L.bf_alloc(64, 2)
        rows, issues, stats = L.finalize()
    assert len(rows) == 0 and len(issues) == 0 and stats["n_candidates"] == 0


def test_argument_errors(gpu_lib):
    from links_b200 import Links, LinksError, LinksOptions
    from links_b200.seqio import SeqSet
    with pytest.raises(LinksError):
        Links(LinksOptions(k=0))
    with pytest.raises(LinksError):
        Links(LinksOptions(k=100000))
    with _links() as L:
        with pytest.raises(LinksError):
            L.push_reads(SeqSet.from_records([b"ACGT" * 100]))  # no Bloom filter yet
        with pytest.raises(LinksError):
            L.bf_alloc(12, 3)  # not a multiple of 8


@pytest.mark.parametrize("name,n_ranks", [("adv_tandem", 2), ("adv_offset_k20", 3), ("c1", 2)])
def test_pair_exchange_between_contexts_equals_single_context(gpu_lib, name, n_ranks):
    """The multi-GPU path on one device: n contexts each take a contiguous shard of the reads
    (global read ordinals), export their distinct pairs bucketed by owner, every owner imports its
    buckets, finalizes, and the summed partial rows must equal the reference's rows for the job."""
    import torch

    from links_b200 import Links, LinksOptions, bf_size, multi
    from links_b200.pairing import ROW_DTYPE
    from links_b200.seqio import SeqSet
    c = GoldenCase(name)
    contigs, reads = c.contigs(), c.reads()
    _m, nbytes, h = bf_size(c.meta["contigs_file_bytes"], c.opts["p"])
    opts = dict(k=c.opts["k"], distances=c.opts["d"], step_sizes=c.opts["t"], offset=c.opts["o"], insert_stdev=c.opts["e"])
    ctxs = [Links(LinksOptions(**opts)) for _ in range(n_ranks)]
    try:
        # rank 0 builds the filter, the others adopt its bits device-to-device
        ctxs[0].bf_alloc(nbytes, h)
        ctxs[0].push_contigs(contigs)
        bf = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
        ctxs[0].bf_download_dev(bf.data_ptr())
        for L in ctxs[1:]:
            L.bf_load_dev(bf.data_ptr(), nbytes, h)
            L.push_contigs(contigs)
        sends, counts = [], []
        n_cand = 0
        for r, L in enumerate(ctxs):
            lo, hi = reads.n * r // n_ranks, reads.n * (r + 1) // n_ranks
            L.push_reads(SeqSet(reads.seq, reads.off[lo:hi + 1], reads.names[lo:hi]), first_read_idx=lo)
            n = L.n_unique()
            buf = torch.empty((max(n, 1), 3), dtype=torch.int64, device="cuda")
            only_counts = L.pairs_export(n_ranks, None, 0)
            cnt = L.pairs_export(n_ranks, buf.data_ptr(), max(n, 1))
            assert cnt == only_counts and sum(cnt) == n
            rec = buf[:n].cpu().numpy().view(np.uint64)
            owner = multi.pair_owner(rec[:, 0], rec[:, 1], n_ranks)
            assert np.array_equal(owner, np.repeat(np.arange(n_ranks), cnt).astype(np.uint32))  # grouped by owner
            a, b, _d = L.pairs()
            assert set(zip(a.tolist(), b.tolist())) == set(zip(rec[:, 0].tolist(), rec[:, 1].tolist()))
            sends.append(buf); counts.append(cnt)
        parts, issue_txt = [], b""
        for r, L in enumerate(ctxs):
            L.pairs_clear()
            assert L.n_unique() == 0
            for s in range(n_ranks):
                off = sum(counts[s][:r])
                if counts[s][r]:
                    L.pairs_import(sends[s][off:off + counts[s][r]].data_ptr(), counts[s][r])
            rows, issues, stats = L.finalize()
            n_cand += stats["n_candidates"]
            parts.append(rows)
            issue_txt += L.format_issues(issues)
        merged = multi.sum_rows(parts, ROW_DTYPE)
        assert sort_lines(ctxs[0].format_rows(merged)) == c.expected_tsv()
        on_device = ctxs[0].rows_merge(np.concatenate(parts))  # links_rows_merge: same sum, device sort + segmented reduce
        assert on_device.tobytes() == merged.tobytes()
        assert len(ctxs[0].rows_merge(np.empty(0, ROW_DTYPE))) == 0
        assert sort_issue_records(issue_txt) == c.expected_issues()
    finally:
        for L in ctxs:
            L.close()


@pytest.mark.timeout(280)
@pytest.mark.parametrize("name,n_ranks,inbox_mb,dedup", [("adv_tandem", 2, 1024, "owner"), ("adv_tandem", 2, 64, "local"),
                                                         ("adv_offset_k20", 3, 1024, "owner"), ("c1", 2, 1024, "owner"),
                                                         ("c1", 2, 512, "local"),
                                                         # one distance: stage 4 runs beside the imports (classify_increment_owned)
                                                         ("adv_k40", 2, 1024, "owner"), ("adv_k40", 3, 64, "local"),
                                                         # Bloom filter built in shares + OR over peer memory; table restricted to owned mates
                                                         ("c1", 2, 1024, "owner+sharedbf+mates"), ("adv_ns_lower", 3, 256, "owner+sharedbf"),
                                                         ("adv_tandem", 2, 256, "local+mates")])
def test_peer_memory_exchange_equals_single_context(gpu_lib, name, n_ranks, inbox_mb, dedup):
    """links_exchange_attach/finish as deployed: one PROCESS per rank (tests/_exchange_worker.py), here all
    on GPU 0, inboxes shared through CUDA IPC like peer memory on an NVLink box.  Reads are sharded;
    routing, import, the wait for the other ranks and the row merge on rank 0 all run inside the
    library.  Two jobs back to back cover the epoch/half switch.  Rank 0's rows == the reference golden;
    the owned pair sets partition the single-context pair set (same winning distance index).  Both
    exchange modes: de-duplication at the owner only (records leave from inside extract_kernel) and local
    de-duplication first (route_kernel after every extract launch)."""
    import queue

    import torch.multiprocessing as mp

    from tests._exchange_worker import rank_main
    c = GoldenCase(name)
    _tsv1, _iss1, stats1, (pa1, pb1, pd1) = _run_gpu(c)
    single = set(zip(pa1.tolist(), pb1.tolist(), pd1.tolist()))
    ctx = mp.get_context("spawn")
    queues = [ctx.Queue() for _ in range(n_ranks)]
    results = ctx.Queue()
    barrier = ctx.Barrier(n_ranks)
    n_jobs = 2
    procs = [ctx.Process(target=rank_main, args=(r, n_ranks, name, inbox_mb << 20, queues, results, barrier, n_jobs, dedup), daemon=True)
             for r in range(n_ranks)]
    for p in procs:
        p.start()
    out: dict = {}
    try:
        for _ in range(n_ranks):
            r, status, payload = results.get(timeout=200)
            assert status == "ok", f"rank {r}: {payload}"
            out[r] = payload
    except queue.Empty:
        pytest.fail(f"ranks {sorted(set(range(n_ranks)) - set(out))} did not report (deadlock?)")
    finally:
        for p in procs:
            p.join(timeout=20)
        for p in procs:
            if p.is_alive():
                p.kill()
    for job in range(n_jobs):
        assert sort_lines(out[0][job]["tsv"]) == c.expected_tsv(), f"job {job}"
        owned = []
        for r in range(n_ranks):
            o = out[r][job]
            if r:
                assert o["stats"]["n_rows"] == 0
            owned.append(set(zip(np.frombuffer(o["pa"], np.uint64).tolist(), np.frombuffer(o["pb"], np.uint64).tolist(),
                                 np.frombuffer(o["pd"], np.uint32).tolist())))
        assert sum(len(o) for o in owned) == len(single)
        assert set().union(*owned) == single
        iss = b"".join(out[r][job]["issues"] for r in range(n_ranks))
        assert sort_issue_records(iss) == c.expected_issues()
        assert sum(out[r][job]["stats"]["n_candidates"] for r in range(n_ranks)) == stats1["n_candidates"]
        for r in range(n_ranks):  # a filter built in shares is, after the OR, the reference's filter on every rank
            assert out[r][job]["bloom_sha"] in (None, c.meta["bloom_sha256"]), f"rank {r} job {job}"


@pytest.mark.parametrize("name,host_threads", [("c1", None), ("adv_offset_k20", None), ("adv_ns_lower", None), ("adv_ns_lower", "1"),
                                               ("adv_tandem", "5"), ("c1", "bf-blocks"), ("c1", "bf-blocks-batches"), ("c1", "bf-passes"), ("adv_k40", "layout-in-process"),
                                               ("c1", "stage-grows"), ("c1", "stage-grows-serial")])
def test_links_cpp_binary_writes_reference_files(gpu_lib, tmp_path, name, host_threads):
    """The C++ host LINKS_CPP, run as the reference binary is run (bin/LINKS-make:121-122): same
    options in, byte-identical .bloom, and .tigpair_checkpoint.tsv / .pairing_issues equal to the
    unmodified reference's after the sort the reference's CI applies."""
    import subprocess
    from links_b200 import build
    c = GoldenCase(name)
    cpath, fof = c.materialize(str(tmp_path))
    base = str(tmp_path / "out")
    cmd = [build.HOST_BIN, "-f", cpath, "-s", fof, "-b", base, "-k", str(c.opts["k"]), "-d", ",".join(map(str, c.opts["d"])),
           "-t", ",".join(map(str, c.opts["t"])), "-o", str(c.opts["o"]), "-e", str(c.opts["e"]), "-p", str(c.opts["p"]), "-j", "3"]
    env = dict(os.environ, LINKS_STAGE_MB="2")  # small staging buffer: several pushes per file
    if name == "c1":
        env["LINKS_STAGE_MB"] = "8"  # its single 1 Mbp read must fit
    lay_meta = None
    if host_threads in ("stage-grows", "stage-grows-serial"):  # the 1 Mbp read does not fit 0.25 MB: the staging buffers are replaced by larger ones
        env["LINKS_STAGE_MB"] = "0.25"
        host_threads = "1" if host_threads.endswith("serial") else None
    if host_threads == "layout-in-process":  # LINKS_CPP_LAYOUT=1: the layout stage runs inside LINKS_CPP on the rows still in memory
        import json
        lay_meta = json.load(open(os.path.join(GOLDEN_DIR, name, "layout", "layout.json")))
        env["LINKS_CPP_LAYOUT"] = "1"
        cmd += ["-l", str(lay_meta["opts"]["l"]), "-a", str(lay_meta["opts"]["a"]), "-z", str(lay_meta["opts"]["z"])]
        host_threads = None
    if host_threads in ("bf-blocks", "bf-blocks-batches", "bf-passes"):  # the Bloom filter built as for filters beyond L2, with 0.25 MB "L2-sized" blocks:
        env["LINKS_BF_BLOCK_MB"] = "0.25"
        if host_threads == "bf-blocks-batches":  # ... partition + apply (bloom_partition_kernel), the updates staged in many small batches
            env["LINKS_BF_STAGE_MB"] = "0.5"
        if host_threads == "bf-passes":          # ... the older way, one pass over the contigs per block (contig_kernel's BloomPass)
            env["LINKS_BF_PARTITION"] = "0"
            env["LINKS_BF_MAX_PASSES"] = "64"
        host_threads = None
    if host_threads:  # "1": the serial reader (seq_stream.hpp); default: the multi-threaded FASTA ingest (fasta_ingest.hpp)
        env["LINKS_HOST_THREADS"] = host_threads
    r = subprocess.run(cmd, capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr
    for banner in ("STAGE: Bloom filter construction/read", "STAGE: Reading sequence reads", "STAGE: Reading contigs",
                   "STAGE: Pairing contigs", "Pair contig stage completed in: "):
        assert banner in r.stdout
    assert sha256(open(base + ".bloom", "rb").read()) == c.meta["bloom_sha256"]
    assert sort_lines(open(base + ".tigpair_checkpoint.tsv", "rb").read()) == c.expected_tsv()
    assert sort_issue_records(open(base + ".pairing_issues", "rb").read()) == c.expected_issues()
    if lay_meta:
        import hashlib
        assert "STAGE: Scaffold layout (in process)" in r.stdout
        assert open(base + ".scaffolds", "rb").read() == open(os.path.join(GOLDEN_DIR, name, "layout", "expected.scaffolds"), "rb").read()
        assert hashlib.sha256(open(base + ".scaffolds.fa", "rb").read()).hexdigest() == lay_meta["scaffolds_fa_sha256"]
        for suffix in (".scaffolds", ".scaffolds.fa", ".gv", ".assembly_correspondence.tsv"):
            os.remove(base + suffix)  # the separate layout run below must recreate them
    # -r: reuse the filter just written
    r = subprocess.run(cmd[:6] + [base + "2"] + cmd[7:] + ["-r", base + ".bloom"], capture_output=True, text=True, env=env)
    assert r.returncode == 0, r.stderr
    assert sort_lines(open(base + "2.tigpair_checkpoint.tsv", "rb").read()) == c.expected_tsv()
    assert not os.path.exists(base + "2.bloom")
    # the whole pipeline as LINKS-make chains it (bin/LINKS-make:121-125): LINKS_CPP on the GPU, then the layout stage
    # (our bin/LINKS.pl = links_layout) -> scaffolds identical to the reference's
    lay = os.path.join(GOLDEN_DIR, name, "layout")
    if os.path.exists(os.path.join(lay, "layout.json")):
        import hashlib
        import json
        meta = json.load(open(os.path.join(lay, "layout.json")))
        build.build_layout()
        open(str(tmp_path / "empty.fof"), "w").close()
        r = subprocess.run([os.path.join(os.path.dirname(build.LAYOUT_BIN), "LINKS.pl"), "-f", cpath, "-s", str(tmp_path / "empty.fof"), "-b", base,
                            "-l", str(meta["opts"]["l"]), "-a", str(meta["opts"]["a"]), "-z", str(meta["opts"]["z"])], capture_output=True, text=True)
        assert r.returncode == 0, r.stderr
        assert open(base + ".scaffolds", "rb").read() == open(os.path.join(lay, "expected.scaffolds"), "rb").read()
        assert hashlib.sha256(open(base + ".scaffolds.fa", "rb").read()).hexdigest() == meta["scaffolds_fa_sha256"]


def _rows_table(rows, mirrored=False):
    """Rows as a lexicographically sorted int64 table; mirrored: every row replaced by its reverse-strand twin
    (B', A') -- the reference stores both under the same counts (src/LINKS.hpp:699-762)."""
    ta, ra, tb, rb = (rows[n].astype(np.int64) for n in ("tig_a", "rev_a", "tig_b", "rev_b"))
    if mirrored:
        ta, ra, tb, rb = tb, 1 - rb, ta, 1 - ra
    m = np.stack([rows["bin"].astype(np.int64), ta, ra, tb, rb, rows["links"].astype(np.int64), rows["gaps"].astype(np.int64)], axis=1)
    return m[np.lexsort(m.T[::-1])]


def _check_job_properties(rows, issues, stats, reads, k, d, t):
    from links_b200 import synth
    # every (A,B) position the reference visits is counted once; an invalid base removes the k-mers over it
    closed_form = synth.candidate_pairs(reads.off, k, list(d), list(t))
    assert closed_form - closed_form // 1000 <= stats["n_candidates"] <= closed_form  # equal but for the 0.1 % of reads with an N
    assert stats["n_unique_pairs"] <= stats["n_survivors"] <= stats["n_candidates"]
    # each distinct pair lands in exactly one class of pair_contigs (src/LINKS.hpp:824-968)
    assert (stats["ct_ok_pairs"] + stats["ct_problem_pairs"] + stats["ct_ok_contig"] + stats["ct_iz_issues"]
            + stats["ct_illogical"] + stats["ct_gate_fail"]) == stats["n_unique_pairs"]
    assert len(issues) == stats["ct_iz_issues"]
    # every accepted pair is added to [A][B] and to its mirror: the link column sums to twice the accepted pairs
    assert int(rows["links"].sum()) == 2 * stats["ct_ok_pairs"]
    assert len(rows) % 2 == 0 and np.array_equal(_rows_table(rows), _rows_table(rows, mirrored=True))
    assert (rows["tig_a"] != rows["tig_b"]).all() and (rows["links"] > 0).all()


@pytest.mark.parametrize("name,scale,sample_reads", [("c2_bacterial_5mbp", 1.0, 2000), ("c3_celegans_100mbp", 0.1, 500)])
def test_full_size_properties(gpu_lib, oracle, name, scale, sample_reads):
    """BASELINE.json configs[1] at its full size (5 Mbp, 500 contigs, 50x 10 kb reads, 75 M candidate pairs) and configs[2]
    at a tenth (multi-distance): beyond what the oracle runs in seconds, so checked through properties that hold at any
    size -- counts that must balance, mirror closure, the same rows from one push, from many pushes, and from a second job
    on the recycled context -- plus the oracle bit for bit on the full-size contigs with the first reads only."""
    from links_b200 import bf_size, synth
    from links_b200.seqio import SeqSet
    _g, contigs, reads, k, d, t = synth.make_config(name, scale)
    _m, nbytes, h = bf_size(len(contigs.seq) + 10 * contigs.n, 0.001)
    with _links(k=k, d=d, t=t) as L:
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        L.push_reads(reads)
        bits = L.bf_bits().copy()
        rows, issues, stats = L.finalize()
        rows, issues = rows.copy(), issues.copy()
        _check_job_properties(rows, issues, stats, reads, k, d, t)
        assert len(rows) > 0 and stats["ct_ok_pairs"] > 0
        # second job on the same context, reads arriving in 7 uneven pushes
        L.reset()
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        cuts = sorted({0, reads.n, *(reads.n * f // 100 for f in (3, 11, 40, 41, 77, 99))})
        for lo, hi in zip(cuts[:-1], cuts[1:]):
            base = int(reads.off[lo])
            L.push_reads(SeqSet(reads.seq[base:int(reads.off[hi])], reads.off[lo:hi + 1] - np.uint64(base), []), first_read_idx=lo)
        rows2, issues2, stats2 = L.finalize()
        assert np.array_equal(_rows_table(rows2), _rows_table(rows)) and len(issues2) == len(issues)
        assert {n: stats2[n] for n in stats if n.startswith(("n_c", "n_s", "n_u", "ct_"))} == \
               {n: stats[n] for n in stats if n.startswith(("n_c", "n_s", "n_u", "ct_"))}
        # full-size filter and contig table, bounded read sample: the oracle, bit for bit
        L.reset()
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        sub = SeqSet(reads.seq[:int(reads.off[sample_reads])], reads.off[:sample_reads + 1], [])
        L.push_reads(sub)
        rows3, issues3, stats3 = L.finalize()
        tsv3, iss3 = L.format_rows(rows3), L.format_issues(issues3)
    assert np.array_equal(bits, oracle.bf_build(contigs.seq, contigs.off, nbytes, h, k))
    orows, oiss, ostats, _ = oracle.run(k, d, t, 0, 0.1, bits, h, contigs.seq, contigs.off, sub.seq, sub.off)
    assert tsv3 == oracle.format_rows(orows) and iss3 == oracle.format_issues(oiss)
    for key in ("n_candidates", "n_survivors", "n_unique_pairs", "ct_ok_pairs", "ct_problem_pairs", "ct_ok_contig",
                "ct_iz_issues", "ct_illogical", "ct_gate_fail"):
        assert stats3[key] == ostats[key], key


FULL_CASES = sorted(d[5:] for d in os.listdir(GOLDEN_DIR) if d.startswith("full_"))


@pytest.mark.parametrize("name", FULL_CASES)
def test_full_size_digests_match_reference(gpu_lib, name):
    """BASELINE.json configs at their FULL size against the unmodified reference, bit for bit: tools/make_fullsize_golden.py
    ran oracle/_ref/LINKS_CPP -j 1 on the same seeded inputs (C2: 75 M candidate pairs; C5's dense point -t 1: 175 M; C3:
    789 M over three distances) and stored sha256 digests of the byte-sorted .tigpair_checkpoint.tsv, of the
    record-sorted .pairing_issues and of the .bloom file, plus the counters of its PAIRED K-MER STATS block."""
    import hashlib
    import json

    from links_b200 import bf_header, bf_size, synth
    meta = json.load(open(os.path.join(GOLDEN_DIR, "full_" + name, "case.json")))
    _g, contigs, reads, k, d, t = synth.make_config(meta["workload"], meta["scale"])
    h_in = hashlib.sha256()
    for a in (contigs.seq, contigs.off, reads.seq, reads.off):
        h_in.update(memoryview(np.ascontiguousarray(a)))
    assert h_in.hexdigest() == meta["inputs_sha256"], "links_b200.synth no longer reproduces the golden inputs"
    o = meta["opts"]
    assert (k, list(d), list(t)) == (o["k"], o["d"], o["t"])
    _m, nbytes, h = bf_size(meta["contigs_file_bytes"], o["p"])
    assert (_m, h) == (meta["bf_m_bits"], meta["bf_hash_num"])
    with _links(k=k, d=d, t=t, o=o["o"], e=o["e"]) as L:
        L.bf_alloc(nbytes, h)
        L.push_contigs(contigs)
        L.push_reads(reads)
        hb = hashlib.sha256(bf_header(k, h, nbytes))
        hb.update(memoryview(L.bf_bits()))
        rows, issues, stats = L.finalize()
        tsv, iss = L.format_rows(rows), L.format_issues(issues)
    assert len(rows) == meta["n_rows"] and len(issues) == meta["n_issues"]
    assert sha256(sort_lines(tsv)) == meta["tsv_sha256"]
    assert sha256(sort_issue_records(iss)) == meta["issues_sha256"]
    assert hb.hexdigest() == meta["bloom_sha256"]
    for key, want in meta["stats"].items():  # the order-independent counters the reference prints (src/LINKS.hpp:985-1012)
        assert stats[key] == want, key


def test_bloom_filter_of_2_pow_31_bits_and_more(gpu_lib, oracle):
    """Filters below 2^31 bits take a 32-bit remainder path in bloom_mod (nthash.cuh); this one is just above and takes the
    64-bit one.  Same bytes as the oracle (= btllib's h % array_bits, bloom_filter.hpp:617-646) either way."""
    c = GoldenCase("adv_ns_lower")
    contigs = c.contigs()
    for nbytes in ((1 << 28) - 8, (1 << 28) + 8):
        with _links(k=c.opts["k"], d=c.opts["d"], t=c.steps(), o=c.opts["o"], e=c.opts["e"]) as L:
            L.bf_alloc(nbytes, 9)
            L.push_contigs(contigs)
            bits = L.bf_bits()
        want = oracle.bf_build(contigs.seq, contigs.off, nbytes, 9, c.opts["k"])
        assert sha256(bits.tobytes()) == sha256(want.tobytes()), nbytes


C1_DISTANCES = "1000,2000,3000,4000,5000,6000,7000,8000,9000,10000,12000,14000,16000,18000,20000"


@pytest.mark.parametrize("layout", ["ours", "perl"])
def test_unmodified_wrapper_chain_over_gpu_binary(gpu_lib, tmp_path, layout):
    """The reference's own command line, unmodified: `LINKS` (sh, bin/LINKS:42-124) -> `LINKS-make` (bin/LINKS-make:121-128)
    -> whatever `LINKS_CPP` and `LINKS.pl` are first on PATH.  Exactly tests/run_me.sh:1-2 (the installation test: 15
    distances, defaults otherwise; LINKS-make derives -b, passes -r/-x/-v/-j in its own order), then the two sorted diffs of
    azure-pipelines.yml:43-47.  `ours`: this repo's GPU LINKS_CPP + C++ layout; `perl`: GPU LINKS_CPP + the reference's Perl.
    The wrappers are the reference's files installed by oracle/Makefile into oracle/_ref (git-ignored); skipped without them."""
    import hashlib
    import json
    import shutil
    import subprocess
    from links_b200 import build
    ref = os.path.join(os.path.dirname(GOLDEN_DIR), "..", "oracle", "_ref")
    ref = os.path.abspath(ref)
    if not os.path.exists(os.path.join(ref, "bin", "LINKS-make")) or not shutil.which("make"):
        pytest.skip("oracle/_ref/bin/LINKS-make (the reference's wrappers) or make not available on this box")
    if layout == "perl" and (not shutil.which("perl") or not os.path.exists(os.path.join(ref, "perl", "LINKS.pl"))):
        pytest.skip("perl or the reference's LINKS.pl not available on this box")
    build.build_all()
    ours = os.path.dirname(build.HOST_BIN)
    only_cpp = str(tmp_path / "only_cpp")  # a directory that holds our LINKS_CPP and nothing else
    os.makedirs(only_cpp)
    os.symlink(build.HOST_BIN, os.path.join(only_cpp, "LINKS_CPP"))
    path = ([ours] if layout == "ours" else [only_cpp, os.path.join(ref, "perl")]) + [os.path.join(ref, "bin"), os.environ.get("PATH", "")]
    c = GoldenCase("c1")
    wd = tmp_path / "run"
    wd.mkdir()
    open(wd / "simulated_contigs.fa", "wb").write(c.raw("contigs.fa"))
    open(wd / "simulated_genome.fa", "wb").write(c.raw("simulated_genome.fa"))
    open(wd / "reads.fof", "w").write("simulated_genome.fa\n")
    r = subprocess.run(["LINKS", "-f", "simulated_contigs.fa", "-s", "reads.fof", "-d", C1_DISTANCES, "-b", "installation_test"],
                       cwd=wd, capture_output=True, text=True, env=dict(os.environ, PATH=os.pathsep.join(path), LINKS_STAGE_MB="8"))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Running Command: LINKS-make LINKS" in r.stdout
    assert sort_lines(open(wd / "installation_test.tigpair_checkpoint.tsv", "rb").read()) == c.expected_tsv()
    assert sha256(open(wd / "installation_test.bloom", "rb").read()) == c.meta["bloom_sha256"]
    meta = json.load(open(os.path.join(GOLDEN_DIR, "c1", "layout", "layout.json")))
    assert hashlib.sha256(open(wd / "installation_test.scaffolds.fa", "rb").read()).hexdigest() == meta["scaffolds_fa_sha256"]
    assert open(wd / "installation_test.scaffolds", "rb").read() == open(os.path.join(GOLDEN_DIR, "c1", "layout", "expected.scaffolds"), "rb").read()


def test_device_generator_is_deterministic_and_shards_concatenate(gpu_lib):
    """links_b200/synth_dev.py (the human-scale workload of BASELINE.json configs[3], generated on the device): the same
    bytes from every generation, and the ranks' read shards concatenate to exactly the reads one rank generates -- the
    premise of the strong-scaling runs and of `bench.py --check-sharding` on that workload."""
    import torch

    from links_b200 import synth_dev
    dev = torch.device("cuda", 0)
    a = synth_dev.make_config_dev("c4_human_3gbp", 0.002, dev)
    b = synth_dev.make_config_dev("c4_human_3gbp", 0.002, dev)
    assert torch.equal(a.d_contigs, b.d_contigs) and torch.equal(a.d_reads, b.d_reads)
    assert np.array_equal(a.contigs_off, b.contigs_off) and np.array_equal(a.reads_off, b.reads_off)
    parts = [synth_dev.make_config_dev("c4_human_3gbp", 0.002, dev, rank=r, world=3) for r in range(3)]
    assert torch.equal(torch.cat([p.d_reads for p in parts]), a.d_reads)
    assert [p.first_read for p in parts] == [0, len(parts[0].reads_off) - 1, len(parts[0].reads_off) + len(parts[1].reads_off) - 2]
    assert sum(len(p.reads_off) - 1 for p in parts) == a.n_reads_total == len(a.reads_off) - 1
    for p in parts:
        assert torch.equal(p.d_contigs, a.d_contigs)
    reads = a.d_reads.cpu().numpy()
    assert set(np.unique(reads)) <= set(b"ACGTN") and set(np.unique(a.d_contigs.cpu().numpy())) <= set(b"ACGTN")
    cov = len(reads) / a.genome_bp
    assert 18.0 < cov < 22.0  # 20x: insertions and deletions cancel on average
    lens = np.diff(a.reads_off.astype(np.int64))
    assert 15_000 < lens.mean() < 25_000 and lens.min() >= 150
