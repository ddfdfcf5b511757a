"""bench.py's contract pieces that need no GPU: the reference arm (oracle port on the host cores) prints
one JSON line with the keys the driver reads, and the byte accounting follows SURVEY.md section 8(d)."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import bench  # noqa: E402


def test_reference_arm_prints_one_json_line():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--grid", "24",
                          "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=ROOT)
    assert out.returncode == 0, out.stderr[-2000:]
    lines = [ln for ln in out.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["metric"] == bench.METRIC and d["unit"] == "GB/s"
    assert d["higher_is_better"] is True and d["steps"] == 2 and d["warmup"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == 1
    assert d["cpu_baseline"]["value"] == d["value"] == d["e2e"]["value"]
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0


def test_reference_arm_other_ranks_stay_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                          "--grid", "16", "--steps", "1", "--warmup", "1"], capture_output=True, text=True,
                         timeout=300, cwd=ROOT, env=env)
    assert out.returncode == 0 and out.stdout.strip() == ""


def test_algorithmic_bytes_follow_the_survey():
    import numpy as np
    # cfg 1 and cfg 5 figures of SURVEY.md section 8(d)
    assert bench.algorithmic_bytes("spmv", 1048576, 1048576, 5238784, 1, np.float64) == 83836932
    assert bench.algorithmic_bytes("spmv", 8000000, 8000000, 213847192, 1, np.float64) == 2726166308
    assert bench.algorithmic_bytes("spmm", 1048576, 1048576, 5238784, 64, np.complex128) == 2256453636


def test_stored_bytes_note_only_for_compressed_spmv():
    plan = "sell32 rows=8 cols=8 nnz=1000 max_rowlen=27 sell_padding=0.35% chunks=0 unit_stride=50.00% hot_cols=0"
    note = bench.stored_bytes_note(plan, 1, 12000, 1000, 1.0)
    # half of the entries trade 4 B for 4/32 B
    assert note["stored_bytes_per_step"] == int(12000 - 500 * 4 + 500 * 4 / 32)
    assert "index compression" in note["note"]
    assert bench.stored_bytes_note(plan, 64, 12000, 1000, 1.0) == {}
    assert bench.stored_bytes_note(plan.replace("50.00%", "0.00%"), 1, 12000, 1000, 1.0) == {}
    assert bench.stored_bytes_note("csr-balanced rows=8", 1, 12000, 1000, 1.0) == {}
