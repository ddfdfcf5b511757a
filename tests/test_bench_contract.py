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


def test_reference_arm_honours_the_workload():
    """--impl reference --workload cfg3 times the oracle's mm (not a silent fall-back to cfg 5)."""
    for wl, metric, unit in (("cfg3", "spgemm_output_nnz_per_s", "nnz/s"), ("cfg3t", "transpose_nnz_per_s", "nnz/s"),
                             ("cfg3a", "assembly_triplets_per_s", "triplets/s"), ("cfg2", "spmm_effective_hbm_gbps", "GB/s")):
        out = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", wl,
                              "--grid", "16", "--steps", "2", "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=ROOT)
        assert out.returncode == 0, out.stderr[-2000:]
        d = json.loads(out.stdout.strip())
        assert d["impl"] == "reference" and d["metric"] == metric and d["unit"] == unit and d["value"] > 0


def test_algorithmic_bytes_follow_the_survey():
    from sparse_linear_b200 import generators as g
    # cfg 1, cfg 5 and cfg 2 figures of SURVEY.md section 8(d)
    assert g.spmv_bytes(1048576, 1048576, 5238784) == 83836932
    assert g.spmv_bytes(8000000, 8000000, 213847192) == 2726166308
    assert g.spmm_bytes(1048576, 1048576, 5238784, 64, 16) == 2256453636


def test_stored_bytes_note_only_for_compressed_spmv():
    plan = "sell32 rows=8 cols=8 nnz=1000 max_rowlen=27 sell_padding=0.35% chunks=0 unit_stride=50.00% hot_cols=0"
    note = bench.stored_bytes_note(plan, 12000, 1000, 1.0)
    # half of the entries trade 4 B for 4/32 B
    assert note["stored_bytes_per_step"] == int(12000 - 500 * 4 + 500 * 4 / 32)
    assert "index compression" in note["note"]
    assert bench.stored_bytes_note(plan.replace("50.00%", "0.00%"), 12000, 1000, 1.0) == {}
    assert bench.stored_bytes_note("csr-balanced rows=8", 12000, 1000, 1.0) == {}


def test_package_does_not_import_the_benchmark():
    """The product must not depend on its own benchmark (round-1 review): no `import bench` under the package."""
    pkg = os.path.join(ROOT, "sparse-linear_b200")
    for name in os.listdir(pkg):
        if name.endswith(".py"):
            src = open(os.path.join(pkg, name)).read()
            assert "import bench" not in src and "oracle" not in [ln.split()[1] for ln in src.splitlines()
                                                                   if ln.strip().startswith("import ") and len(ln.split()) > 1], name
