#!/bin/bash
# GPU run Z (1 x B200): in-kernel backward epilogue v2 (warp-per-rows), and occupancy variants of the epilogue kernel.
mkdir -p gpurun_out
P=cuda-nt-xent-mpi-nccl-simclr_b200
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_guard_bands_gpu.py tests/test_comm_gpu.py -x -q > gpurun_out/r2_z_tests.log 2>&1; echo "tests rc=$?"
tail -4 gpurun_out/r2_z_tests.log
B="--no-cpu-baseline --no-sustained"
run() { # name, env...
  name=$1; shift
  env "$@" timeout 300 python bench.py $B > gpurun_out/r2_z_bench_$name.json 2>> gpurun_out/r2_z_bench.err
  python - "$name" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_z_bench_{n}.json")); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f} loss {d.get('loss')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
run tail A=1
run notail NTXENT_NO_BWD_TAIL=1
for v in e43 e44 e24 e26; do run $v NTXENT_NO_BWD_TAIL=1 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_$v.so; done
run tail2 A=1
run notail2 NTXENT_NO_BWD_TAIL=1
tail -3 gpurun_out/r2_z_bench.err
NTXENT_NO_PDL=1 python tools/kernel_times.py 2>&1 | tail -6
NTXENT_NO_PDL=1 python tools/kernel_times.py --M 4096 2>&1 | tail -6
for v in e44 e26; do NTXENT_NO_BWD_TAIL=1 NTXENT_NO_PDL=1 NTXENT_LIBRARY=$PWD/$P/libntxent_b200_$v.so python tools/kernel_times.py 2>&1 | tail -2; done
