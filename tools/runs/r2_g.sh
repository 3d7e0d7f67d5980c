#!/bin/bash
# GPU run G (1 x B200): all GPU tests, then bench A/B lines (backward variants; the full-matrix
# forward three times over to see its run-to-run spread).
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_g_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/r2_g_tests.log
B="--no-cpu-baseline --no-sustained"
P=cuda-nt-xent-mpi-nccl-simclr_b200
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py $B > gpurun_out/r2_g_bench_$name.json 2>> gpurun_out/r2_g_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_g_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run sym A=1
run bwdunpacked NTXENT_LIBRARY=$PWD/$P/libntxent_b200_bwdunpacked.so
run bp1 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_bp1.so
run nosym_a NTXENT_NO_SYM_FORWARD=1
run nosym_b NTXENT_NO_SYM_FORWARD=1
run nosym_nopos NTXENT_NO_SYM_FORWARD=1 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_nopos.so
run nosym_c NTXENT_NO_SYM_FORWARD=1
run sym2 A=1
grep -v "^\*\|^$\|OMP_NUM" gpurun_out/r2_g_bench.err | tail -5
