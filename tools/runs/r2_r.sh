#!/bin/bash
# GPU run Q (1 x B200): two S buffers + start-up phase skew of the weight warps.
mkdir -p gpurun_out
P=cuda-nt-xent-mpi-nccl-simclr_b200
B="--no-cpu-baseline --no-sustained"
run() { # name, env...
  name=$1; shift
  env NTXENT_LIBRARY=$PWD/$P/libntxent_b200_s3.so "$@" timeout 300 python bench.py $B > gpurun_out/r2_r_bench_$name.json 2>> gpurun_out/r2_r_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_r_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} loss {d.get('loss')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run s0 NTXENT_BWD_SKEW=0
run s0b NTXENT_BWD_SKEW=0
run s2 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_s2.so
run base NTXENT_LIBRARY=$PWD/$P/libntxent_b200.so
tail -3 gpurun_out/r2_r_bench.err
timeout 200 python tools/trace_bwd.py > gpurun_out/r2_r_trace.log 2>&1; echo "trace rc=$?"
grep -v "^tile" gpurun_out/r2_r_trace.log | tail -8
NTXENT_LIBRARY=$PWD/$P/libntxent_b200_s3.so timeout 900 python -m pytest tests/test_parity_gpu.py -x -q 2>&1 | tail -3
