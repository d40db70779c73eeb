"""Host-side logic of bench.py that needs no GPU: rank -> device placement, the workload string shared by the two arms,
the single-rank host gather, the reference arm's JSON contract on a small sample."""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_spread_device_is_a_permutation_over_both_halves():
    for n in (1, 2, 3, 4, 6, 8):
        devs = [bench.spread_device(r, n) for r in range(n)]
        assert sorted(devs) == list(range(n)), (n, devs)          # one rank per GPU, every GPU used at N = n
    assert [bench.spread_device(r, 8) for r in range(8)] == [0, 4, 1, 5, 2, 6, 3, 7]
    assert [bench.spread_device(r, 8) for r in range(4)] == [0, 4, 1, 5]   # a 4-rank job straddles both groups of four
    assert [bench.spread_device(r, 8) for r in range(2)] == [0, 4]
    assert bench.spread_device(0, 1) == 0 and bench.spread_device(1, 2) == 1


def test_workload_strings_name_the_batch():
    assert "batch512" in bench.workload_name("lk", 512) and "640x480" in bench.workload_name("lk", 512)
    assert "batch4096" in bench.workload_name("sad", 4096)
    assert bench.workload_name("sad", 8192) != bench.workload_name("sad", 4096)


def test_single_rank_needs_no_gather(pkg):
    res = np.zeros(5, pkg.capi.flow_out_dtype)
    assert bench.host_gather_open(res, 0, 1) is None


def test_reference_arm_contract_small_sample():
    """python bench.py --impl reference on 8 LK streams: ONE JSON line, impl / cpu_baseline / e2e keys, the workload string
    of the b200 arm for the same batch"""
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "3",
                          "--lk-streams", "8"], capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [l for l in out.stdout.splitlines() if l.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == "frame-pairs/s" and d["higher_is_better"] is True
    assert d["config"]["workload"] == bench.workload_name("lk", 8)
    assert d["e2e"] == {"value": d["value"], "unit": "frame-pairs/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["gpu_launches"] == 0 and d["value"] > 0
