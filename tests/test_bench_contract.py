"""bench.py's reference arm runs on the host cores, so its JSON line can be checked here: the
keys the driver reads, the bounded CPU sample, and that ranks other than 0 stay silent."""
import json
import os
import subprocess
import sys

from conftest import ROOT


def _run(extra_env=None):
    env = dict(os.environ)
    env.update(extra_env or {})
    cmd = [sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--config", "c1",
           "--steps", "2", "--warmup", "1", "--gpus", "1"]
    return subprocess.run(cmd, capture_output=True, text=True, timeout=300, env=env, cwd=ROOT)


def test_reference_arm_json_line():
    p = _run()
    assert p.returncode == 0, p.stderr[-2000:]
    lines = [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
    assert len(lines) == 1
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["higher_is_better"] is True
    assert d["metric"] == "ntxent_fwd_bwd_tflops" and d["unit"] == "TFLOP/s"
    assert d["steps"] == 2 and d["warmup"] == 1 and d["n_gpus"] == 1
    assert d["value"] > 0 and d["ms_per_step"] > 0
    assert d["config"]["workload"].startswith("ntxent_fwd_bwd_M512_D128")
    cb = d["cpu_baseline"]
    # C1 is small: the CPU arm times whole steps, and says so
    assert cb["kind"] == "port" and cb["cores"] >= 1 and cb["value"] == d["value"] and "whole step" in cb["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["gpu_launches"] == 0 and d["vs_baseline"] is None


def test_reference_arm_uses_every_core_under_a_launcher():
    """torchrun exports OMP_NUM_THREADS=1 to its ranks; the CPU arm must not inherit that."""
    p = _run({"OMP_NUM_THREADS": "1", "RANK": "0", "WORLD_SIZE": "2", "LOCAL_RANK": "0"})
    assert p.returncode == 0, p.stderr[-2000:]
    d = json.loads([ln for ln in p.stdout.splitlines() if ln.startswith("{")][0])
    assert d["cpu_baseline"]["cores"] == (os.cpu_count() or 1)


def test_reference_arm_other_ranks_do_no_work():
    p = _run({"RANK": "1", "WORLD_SIZE": "2", "LOCAL_RANK": "1"})
    assert p.returncode == 0
    assert not [ln for ln in p.stdout.splitlines() if ln.startswith("{")]
