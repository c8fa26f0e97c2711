"""bench.py's output contract on the arm that runs without a GPU: `--impl reference` times the unmodified reference
(oracle/_ref/LINKS_CPP) and prints exactly ONE JSON line on stdout with the keys the driver reads."""
from __future__ import annotations

import json
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    from oracle import oracle_py
    if not oracle_py.have_ref_bin():
        pytest.skip("oracle/_ref/LINKS_CPP not built (needs /root/reference at build time)")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "tiny", "--steps", "1",
                        "--warmup", "0", "--ref-candidates", "100000"], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, r.stdout
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["metric"] == "kmer_pairs_per_sec_extracted_probed_mapped" and j["unit"] == "pairs/s"
    assert j["higher_is_better"] is True and j["value"] > 0 and j["steps"] == 1 and j["n_gpus"] == 1
    assert j["cpu_baseline"]["kind"] == "reference" and 1 <= j["cpu_baseline"]["cores"] <= 4 and j["cpu_baseline"]["value"] == j["value"]
    assert j["e2e"] == {"value": j["value"], "unit": "pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert j["config"]["workload"] == "tiny" and "sample" in j["config"]


def test_other_ranks_of_the_reference_arm_do_nothing():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "tiny", "--gpus", "2"],
                       capture_output=True, text=True, timeout=120, env=dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1"))
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_our_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--workload", "tiny", "--steps", "1", "--warmup", "0"],
                       capture_output=True, text=True, timeout=300)
    assert r.returncode != 0 and "no CPU fallback" in (r.stderr + r.stdout)
