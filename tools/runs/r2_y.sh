#!/bin/bash
# GPU run Y (1 x B200): the backward epilogue inside the tile kernel -- all GPU tests, A/B bench, in-situ kernel times.
mkdir -p gpurun_out
timeout 1700 python -m pytest tests -m gpu -x -q > gpurun_out/r2_y_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r2_y_tests.log
B="--no-cpu-baseline --no-sustained"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py $B > gpurun_out/r2_y_bench_$name.json 2>> gpurun_out/r2_y_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_y_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f} launches {d.get('gpu_launches')} loss {d.get('loss')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run tail A=1
run notail NTXENT_NO_BWD_TAIL=1
run tail2 A=1
run notail2 NTXENT_NO_BWD_TAIL=1
tail -3 gpurun_out/r2_y_bench.err
NTXENT_NO_PDL=1 python tools/kernel_times.py 2>&1 | tail -6
NTXENT_NO_PDL=1 python tools/kernel_times.py --M 4096 2>&1 | tail -6
