"""bench.py's reference arm (the CPU oracle timed through the bench contract):
runs without a GPU, rank 0 alone prints the line, the line pairs with the B200
arm's by ``n_gpus`` / ``metric`` / ``unit``."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(rank, world):
    env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
               OMP_NUM_THREADS="1")  # what torchrun exports; the arm must not inherit it
    return subprocess.run(
        [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--workload", "C1", "--n", "32",
         "--steps", "1", "--warmup", "0", "--gpus", str(world)],
        env=env, capture_output=True, text=True, timeout=300)


def test_reference_arm_line_on_rank0():
    r = _run(0, 2)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [ln for ln in r.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2
    assert d["metric"] == "rays_per_s" and d["unit"] == "rays/s" and d["higher_is_better"] is True
    assert d["value"] > 0 and d["gpu_launches"] == 0
    assert d["cpu_baseline"]["kind"] == "port" and d["cpu_baseline"]["cores"] == (os.cpu_count() or 1)
    assert d["e2e"] == {"value": d["value"], "unit": "rays/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["config"]["gpus_used"] == 0 and d["config"]["workload"].startswith("C1")


def test_reference_arm_other_ranks_exit_quietly():
    r = _run(1, 2)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""
