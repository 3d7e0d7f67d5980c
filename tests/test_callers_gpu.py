"""GPU tests of the callers either side of the hot path (SURVEY.md section 8f, section 2's IN
rows): the C++ operator tests, the C++ benchmark and the Python operator sweep -- this library's
restatements of the reference's tests/*.cpp, src/benchmark.cpp and python/test.py."""
import json
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _tools():
    import ntxent_b200  # noqa: F401
    from ntxent_b200 import _build
    return _build.build_cpp_tools()


def test_cpp_operator_tests():
    exe = _tools()[0]
    r = subprocess.run([exe], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    assert "7 of 7 tests passed" in r.stdout


def test_cpp_benchmark_quick(tmp_path):
    bench = _tools()[1]
    out = tmp_path / "bench.json"
    r = subprocess.run([bench, "--quick", "--runs", "10", "--json", str(out)], capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]
    rows = json.loads(out.read_text())
    assert len(rows) >= 6 and all(row["fwd_ms_mean"] > 0 and row["tflops"] > 0 for row in rows)
    big = [row for row in rows if row["rows"] == 32768][0]
    assert big["tflops"] > 100  # the tcgen05 path, not a fallback


def test_cpp_dist_benchmark_one_gpu():
    """The multi-GPU C++ benchmark (C-ABI communicator, no torch) at --gpus 1: every exchange form,
    directly and through `ntxent_benchmark --gpus`."""
    tools = _tools()
    for exe, extra in ((tools[2], []), (tools[1], []), (tools[2], ["--exchange", "nccl", "--bwd", "contract"])):
        r = subprocess.run([exe, "--gpus", "1", "--M", "4096", "--steps", "5", "--warmup", "2", *extra],
                           capture_output=True, text=True, timeout=300)
        assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
        d = json.loads([ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1])
        assert d["gpus"] == 1 and d["ms_mean"] > 0 and d["loss_same_on_all_ranks"] is True
        assert abs(d["loss"] - 8.7) < 1.0  # ~ ln(4096) + O(1)


def test_python_operator_sweep(tmp_path):
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "op_sweep.py"), "--quick", "--runs", "10",
                        "--out", str(tmp_path)], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    files = [f for f in os.listdir(tmp_path) if f.startswith("results_")]
    assert files and os.path.exists(tmp_path / "memory_profile.json")
    data = json.loads((tmp_path / files[0]).read_text())
    assert len(data["results"]) == 2
    # O(M*D) memory: far below the reference's three [M,M] fp32 temporaries at 2B = 1024 rows
    assert data["results"][-1]["mixed"]["peak_extra_memory_mb"] < data["results"][-1]["mixed"]["reference_temporaries_mb"]
