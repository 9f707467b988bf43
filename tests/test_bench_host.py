"""bench.py host logic that runs without a GPU: workload helpers, the reference arm (the reference's CPU path timed on the
host cores) and its behaviour under a multi-rank launch (rank 0 alone works and prints)."""
import json
import os
import subprocess
import sys

from cases import ROOT

sys.path.insert(0, ROOT)
import bench  # noqa: E402


def test_hierarchy_and_algorithmic_bytes_match_the_survey():
    P = bench.interior_points(511)
    assert len(P) == 9 and P[0] == 511 ** 3 and P[-1] == 1 and sum(P) == 152346173
    b = bench.vcycle_algorithmic_bytes(511, 2, 2)
    assert abs(b - 23.0e9) < 0.05e9
    assert bench.vcycle_algorithmic_bytes(511, 2, 2, mode=2) > b
    assert bench.interior_points((9, 12, 17)) == [9 * 12 * 17, 4 * 6 * 8, 2 * 3 * 4, 1 * 1 * 2]


def test_default_grids_are_weak_scaling_and_end_at_1025_cubed():
    assert bench.default_grid(1) == (511, 511, 511)
    assert bench.default_grid(2) == (511, 511, 1023)
    assert bench.default_grid(4) == (511, 1023, 1023)
    assert bench.default_grid(8) == (1023, 1023, 1023)
    assert bench.work_ratio((511, 511, 511)) == 1.0
    assert 7.9 < bench.work_ratio((1023, 1023, 1023)) < 8.1


def test_cpu_sample_is_bounded():
    n = bench.cpu_sample_grid((511, 511, 511), steps_total=3, cores=1, budget_s=10.0)
    assert min(n) >= 63 and n[0] < 511
    assert bench.cpu_sample_grid((127, 127, 127), 1, 8, 100.0) == (127, 127, 127)


def _run(args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([sys.executable, os.path.join(ROOT, "bench.py")] + args, capture_output=True, text=True, timeout=300, env=e)


def test_reference_arm_prints_one_json_line_with_the_contract_keys():
    r = _run(["--impl", "reference", "--steps", "1", "--warmup", "0", "--grid", "63"])
    assert r.returncode == 0, r.stderr
    lines = [l for l in r.stdout.splitlines() if l.strip()]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["unit"] == "V-cycles/s" and d["higher_is_better"] is True
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["e2e"]["h2d_bytes_per_step"] == 0 and d["e2e"]["d2h_bytes_per_step"] == 0
    assert d["value"] > 0 and d["config"]["grid"] == [63, 63, 63] and d["dtype"] == "f64"


def test_reference_arm_other_ranks_exit_quietly():
    r = _run(["--impl", "reference", "--gpus", "2", "--steps", "1", "--warmup", "0"], env={"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert r.returncode == 0 and r.stdout.strip() == ""
