#!/bin/bash
# GPU run X (N x B200, N = $1): all GPU tests on N GPUs, then the bench lines at N (C4 push, C4 nccl,
# C5 push) and the C++ multi-GPU benchmark.
N=${1:-2}
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu_gpu.py tests/test_comm_gpu.py -m gpu -x -q -v > gpurun_out/r2_x_tests_${N}gpu.log 2>&1; echo "tests rc=$?"
tail -4 gpurun_out/r2_x_tests_${N}gpu.log
summ() { python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    txt=open(f"gpurun_out/{n}.json").read()
    d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1]); r=d["roofline"]
    print(f"{n:28s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f} parity {d.get('parity')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
trun() { # name, args...
  name=$1; shift
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29561 \
      bench.py --gpus $N --no-sustained "$@" > gpurun_out/$name.json 2>> gpurun_out/r2_x_bench.err
  summ $name
}
trun r2_final_n${N}_push
trun r2_final_n${N}_nccl --comm nccl
trun r2_final_n${N}_c5 --M 131072 --D 256
P=cuda-nt-xent-mpi-nccl-simclr_b200
timeout 200 $P/ntxent_benchmark --gpus $N --steps 50 --warmup 10 > gpurun_out/r2_final_cpp_dist_${N}gpu.jsonl 2>> gpurun_out/r2_x_bench.err
timeout 200 $P/ntxent_benchmark --gpus $N --steps 50 --warmup 10 --exchange nccl --bwd contract >> gpurun_out/r2_final_cpp_dist_${N}gpu.jsonl 2>> gpurun_out/r2_x_bench.err
cat gpurun_out/r2_final_cpp_dist_${N}gpu.jsonl | cut -c1-330
grep -v "^\*\|^$\|OMP_NUM\|NCCL version" gpurun_out/r2_x_bench.err | tail -5
