#!/bin/bash
# GPU run P (1 x B200): backward with two S buffers (variant s2) against the current library.
mkdir -p gpurun_out
P=cuda-nt-xent-mpi-nccl-simclr_b200
NTXENT_LIBRARY=$PWD/$P/libntxent_b200_s2.so timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands_gpu.py -x -q > gpurun_out/r2_p_tests.log 2>&1; echo "tests(s2) rc=$?"
tail -4 gpurun_out/r2_p_tests.log
B="--no-cpu-baseline --no-sustained"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py $B > gpurun_out/r2_p_bench_$name.json 2>> gpurun_out/r2_p_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_p_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} loss {d.get('loss')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run base A=1
run s2 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_s2.so
run base2 A=1
run s2b NTXENT_LIBRARY=$PWD/$P/libntxent_b200_s2.so
tail -3 gpurun_out/r2_p_bench.err
timeout 200 python tools/trace_bwd.py > gpurun_out/r2_p_trace.log 2>&1; echo "trace rc=$?"
grep -v "^tile  [0-5]" gpurun_out/r2_p_trace.log | tail -28
