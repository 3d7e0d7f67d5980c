#!/bin/bash
# GPU run D (1 x B200): bench A/B after moving the positive capture out of the exponential loops,
# then ncu --set full of the three tile kernels (one launch each).
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_parity_gpu.py -m gpu -x -q > gpurun_out/r2_d_tests.log 2>&1; echo "tests rc=$?"
tail -3 gpurun_out/r2_d_tests.log
B="--no-cpu-baseline --no-sustained"
P=cuda-nt-xent-mpi-nccl-simclr_b200
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py $B > gpurun_out/r2_d_bench_$name.json 2>> gpurun_out/r2_d_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_d_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run sym A=1
run nosym NTXENT_NO_SYM_FORWARD=1
run sym_p0 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_p0.so
run nosym_p0 NTXENT_NO_SYM_FORWARD=1 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_p0.so
run nosym_p2 NTXENT_NO_SYM_FORWARD=1 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_p2.so
tail -5 gpurun_out/r2_d_bench.err
A="--steps 2 --warmup 3 --no-cpu-baseline --no-sustained --no-graph"
timeout 300 python bench.py $A > gpurun_out/r2_d_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_fwd_sym_kernel -s 3 -c 1 -f \
    -o gpurun_out/r2_d_fwd_sym python bench.py $A > gpurun_out/r2_d_ncu1.log 2>&1
echo "ncu sym rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_bwd_kernel -s 3 -c 1 -f \
    -o gpurun_out/r2_d_bwd python bench.py $A > gpurun_out/r2_d_ncu2.log 2>&1
echo "ncu bwd rc=$?"
NTXENT_NO_SYM_FORWARD=1 timeout 300 python bench.py $A > gpurun_out/r2_d_plain2.log 2>&1 &&
NTXENT_NO_SYM_FORWARD=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_fwd_kernel -s 3 -c 1 -f \
    -o gpurun_out/r2_d_fwd_full python bench.py $A > gpurun_out/r2_d_ncu3.log 2>&1
echo "ncu full rc=$?"
