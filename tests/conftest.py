"""Shared fixtures.  `-m "not gpu"` runs everywhere; `-m gpu` needs a B200 and calls the C-ABI."""
from __future__ import annotations

import gzip
import hashlib
import json
import os

# peer waits of the multi-GPU exchange give up quickly in tests: a logic error becomes a failure, not a hung GPU
os.environ.setdefault("LINKS_EXCHANGE_TIMEOUT_MS", "8000")
# several contexts share this process in the multi-rank emulation: give every stream its own hardware queue, so a
# kernel that waits for a peer never sits in front of that peer's kernels (must be set before CUDA initialises)
os.environ.setdefault("CUDA_DEVICE_MAX_CONNECTIONS", "32")
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")
GOLDEN_CASES = sorted(d for d in os.listdir(GOLDEN) if not d.startswith(("cfg_", "full_")) and os.path.exists(os.path.join(GOLDEN, d, "case.json")))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); calls liblinks_gpu.so")


def sha256(b: bytes) -> str:
    return hashlib.sha256(b).hexdigest()


class GoldenCase:
    """One tests/golden/<name>/ directory, inputs decompressed into tmp files on demand."""

    def __init__(self, name: str):
        self.name = name
        self.dir = os.path.join(GOLDEN, name)
        self.meta = json.load(open(os.path.join(self.dir, "case.json")))
        self.opts = self.meta["opts"]

    def raw(self, fn: str) -> bytes:
        with gzip.open(os.path.join(self.dir, fn + ".gz"), "rb") as f:
            return f.read()

    def contigs(self):
        from links_b200.seqio import read_seqfile
        return read_seqfile(os.path.join(self.dir, "contigs.fa.gz"))

    def reads(self):
        from links_b200.seqio import SeqSet, read_seqfile
        return SeqSet.concat([read_seqfile(os.path.join(self.dir, fn + ".gz")) for fn in self.meta["read_files"]])

    def steps(self) -> list[int]:
        t = list(self.opts["t"])
        while len(t) < len(self.opts["d"]):
            t.append(t[-1])
        return t

    def expected_tsv(self) -> bytes:
        return open(os.path.join(self.dir, "expected.tigpair_checkpoint.tsv"), "rb").read()

    def expected_issues(self) -> bytes:
        return open(os.path.join(self.dir, "expected.pairing_issues"), "rb").read()

    def materialize(self, tmpdir: str) -> tuple[str, str]:
        """Write contigs.fa + read files + reads.fof uncompressed; returns (contigs path, fof path)."""
        cpath = os.path.join(tmpdir, "contigs.fa")
        open(cpath, "wb").write(self.raw("contigs.fa"))
        paths = []
        for fn in self.meta["read_files"]:
            p = os.path.join(tmpdir, fn)
            open(p, "wb").write(self.raw(fn))
            paths.append(p)
        fof = os.path.join(tmpdir, "reads.fof")
        open(fof, "w").write("\n".join(paths) + "\n")
        return cpath, fof


def sort_lines(b: bytes) -> bytes:
    return b"".join(sorted(b.splitlines(keepends=True)))


def sort_issue_records(text: bytes) -> bytes:
    lines = text.split(b"\n")
    recs = [b"\n".join(lines[i:i + 4]) for i in range(0, len(lines) - 1, 4)]
    return b"".join(r + b"\n" for r in sorted(recs))


@pytest.fixture(params=GOLDEN_CASES)
def golden_case(request) -> GoldenCase:
    return GoldenCase(request.param)


class ConfigCase:
    """A BASELINE.json configuration at reference-runnable size (tools/make_config_golden.py): inputs are
    regenerated from links_b200.synth (seeded), expected outputs come from the unmodified reference."""

    def __init__(self, name: str):
        self.name = name
        self.dir = os.path.join(GOLDEN, "cfg_" + name)
        self.meta = json.load(open(os.path.join(self.dir, "case.json")))
        self.opts = self.meta["opts"]

    def inputs(self):
        from links_b200 import synth
        g, nc, cov, rl, _k, _d, _t = self.meta["synth"]
        genome = synth.make_genome(g, seed=11)
        contigs = synth.make_contigs(genome, nc, seed=13)
        reads = synth.make_reads(genome, cov, rl, seed=12)
        h = hashlib.sha256()
        for a in (contigs.seq, contigs.off, reads.seq, reads.off):
            h.update(a.tobytes())
        assert h.hexdigest() == self.meta["inputs_sha256"], "links_b200.synth no longer reproduces the golden inputs"
        return contigs, reads

    def expected_tsv(self) -> bytes:
        return open(os.path.join(self.dir, "expected.tigpair_checkpoint.tsv"), "rb").read()

    def expected_issues(self) -> bytes:
        return open(os.path.join(self.dir, "expected.pairing_issues"), "rb").read()


CONFIG_CASES = sorted(d[4:] for d in os.listdir(GOLDEN) if d.startswith("cfg_"))


@pytest.fixture(params=CONFIG_CASES)
def config_case(request) -> ConfigCase:
    return ConfigCase(request.param)


@pytest.fixture(scope="session")
def oracle():
    from oracle import oracle_py
    oracle_py.load()
    return oracle_py


def have_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def gpu_lib():
    """The built CUDA library; GPU tests must fail loudly if it is missing (no fallback)."""
    from links_b200 import _abi
    return _abi.load()
