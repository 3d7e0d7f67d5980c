#!/bin/bash
# GPU run N (1 x B200): CTA-pair MMA probe; phase-skew experiment on the backward's weight warps.
mkdir -p gpurun_out
timeout 120 tools/mma2_probe > gpurun_out/r2_n_mma2.log 2>&1; echo "mma2 rc=$?"; cat gpurun_out/r2_n_mma2.log
timeout 120 tools/mma_rate_probe > gpurun_out/r2_n_mma1.log 2>&1; echo "mma1 rc=$?"; cat gpurun_out/r2_n_mma1.log
B="--no-cpu-baseline --no-sustained"
P=cuda-nt-xent-mpi-nccl-simclr_b200
run() { # name, env...
  name=$1; shift
  env NTXENT_LIBRARY=$PWD/$P/libntxent_b200_skew.so "$@" timeout 300 python bench.py $B > gpurun_out/r2_n_bench_$name.json 2>> gpurun_out/r2_n_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_n_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} loss {d.get('loss')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run s0 NTXENT_BWD_SKEW=0
for s in 200 400 600 800 1100; do run p0_$s NTXENT_BWD_SKEW=$s NTXENT_BWD_SKEW_PATTERN=0; done
for s in 150 250 350 500; do run p1_$s NTXENT_BWD_SKEW=$s NTXENT_BWD_SKEW_PATTERN=1; done
for s in 400 700; do run p2_$s NTXENT_BWD_SKEW=$s NTXENT_BWD_SKEW_PATTERN=2; done
run s0b NTXENT_BWD_SKEW=0
tail -3 gpurun_out/r2_n_bench.err
