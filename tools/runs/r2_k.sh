#!/bin/bash
# GPU run K (2 x B200): everything after the fused finalize: all GPU tests (verbose log kept),
# N = 2 bench lines (C4, and 2N = 8192 as a stand-in for the kernel sizes of an 8-GPU shard), PDL on/off.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -v > gpurun_out/r2_k_tests.log 2>&1; echo "tests rc=$?"
tail -5 gpurun_out/r2_k_tests.log
summ() { python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    txt=open(f"gpurun_out/r2_k_{n}.json").read()
    d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1]); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f} parity {d.get('parity',{}).get('dz_vs_1gpu_maxrel')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
tr() { # name, args...
  name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 \
      bench.py --gpus 2 --no-sustained "$@" > gpurun_out/r2_k_$name.json 2>> gpurun_out/r2_k_bench.err
  summ $name
}
tr n2_c4
tr n2_m8192 --M 8192
NTXENT_NO_PDL=1 tr n2_m8192_nopdl --M 8192
tr n2_m8192_eager --M 8192 --no-graph
grep -v "^\*\|^$\|OMP_NUM\|NCCL version" gpurun_out/r2_k_bench.err | tail -8
timeout 300 python bench.py --no-cpu-baseline --no-sustained > gpurun_out/r2_k_n1.json 2>> gpurun_out/r2_k_bench.err; summ n1
NTXENT_NO_SYM_FORWARD=1 timeout 300 python bench.py --no-cpu-baseline --no-sustained > gpurun_out/r2_k_n1_nosym.json 2>> gpurun_out/r2_k_bench.err; summ n1_nosym
