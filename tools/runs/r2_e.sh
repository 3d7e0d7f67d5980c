#!/bin/bash
# GPU run E (1 x B200): all GPU tests (incl. the one-rank communicator), bench lines, ncu of the
# reworked upper-triangle forward.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_e_tests.log 2>&1; echo "tests rc=$?"
tail -8 gpurun_out/r2_e_tests.log
B="--no-cpu-baseline --no-sustained"
P=cuda-nt-xent-mpi-nccl-simclr_b200
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py $B > gpurun_out/r2_e_bench_$name.json 2>> gpurun_out/r2_e_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_e_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run sym A=1
run sym_p2 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_p2.so
run nosym NTXENT_NO_SYM_FORWARD=1
tail -5 gpurun_out/r2_e_bench.err
timeout 120 $P/ntxent_dist_benchmark --gpus 1 --steps 20 --warmup 5 > gpurun_out/r2_e_dist1.json 2>&1; cat gpurun_out/r2_e_dist1.json
timeout 120 $P/ntxent_benchmark --gpus 1 --steps 20 --warmup 5 --exchange nccl --bwd contract > gpurun_out/r2_e_dist1c.json 2>&1; cat gpurun_out/r2_e_dist1c.json
timeout 300 python tools/drop_tile_proof.py > gpurun_out/r2_e_droptile.log 2>&1; cat gpurun_out/r2_e_droptile.log
A="--steps 2 --warmup 3 --no-cpu-baseline --no-sustained --no-graph"
timeout 300 python bench.py $A > gpurun_out/r2_e_plain.log 2>&1 &&
timeout 600 ncu --set full --clock-control none --import-source on -k regex:tc_fwd_sym_kernel -s 3 -c 1 -f \
    -o gpurun_out/r2_e_fwd_sym python bench.py $A > gpurun_out/r2_e_ncu1.log 2>&1
echo "ncu sym rc=$?"
