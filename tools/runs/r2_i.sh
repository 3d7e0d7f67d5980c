#!/bin/bash
# GPU run I (1 x B200): all GPU tests; bench lines with and without the upper-triangle forward.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_i_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/r2_i_tests.log
B="--no-cpu-baseline --no-sustained"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py $B > gpurun_out/r2_i_bench_$name.json 2>> gpurun_out/r2_i_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_i_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run sym A=1
run nosym NTXENT_NO_SYM_FORWARD=1
grep -v "^\*\|^$\|OMP_NUM" gpurun_out/r2_i_bench.err | tail -5
