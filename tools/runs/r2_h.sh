#!/bin/bash
# GPU run H (8 x B200): multi-GPU tests on all GPUs, then the scaling lines (C4 at 8 / 4 GPUs in
# the three exchange forms, C5 at 8 / 4), then the C++ multi-GPU benchmark.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_h_topo.log 2>&1
timeout 900 python -m pytest tests/test_comm_gpu.py tests/test_multigpu_gpu.py -m gpu -x -q > gpurun_out/r2_h_tests.log 2>&1; echo "tests rc=$?"
tail -6 gpurun_out/r2_h_tests.log
summ() { python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    txt=open(f"gpurun_out/r2_h_{n}.json").read()
    d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1]); r=d["roofline"]
    print(f"{n:14s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f} parity {d.get('parity',{}).get('dz_vs_1gpu_maxrel')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
tr() { # nproc, name, args...
  n=$1; name=$2; shift; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29541 \
      bench.py --gpus $n --no-sustained "$@" > gpurun_out/r2_h_$name.json 2>> gpurun_out/r2_h_bench.err
  summ $name
}
tr 8 n8_push
tr 8 n8_push_eager --no-graph
tr 8 n8_nccl --comm nccl
tr 8 n8_contract --comm nccl --bwd-mode contract
tr 4 n4_push
tr 8 n8_c5 --config c5 --steps 5 --warmup 3
tr 4 n4_c5 --config c5 --steps 5 --warmup 3
grep -v "^\*\|^$\|OMP_NUM" gpurun_out/r2_h_bench.err | tail -8
P=cuda-nt-xent-mpi-nccl-simclr_b200
for ex in "push" "nccl" "nccl --bwd contract"; do
  timeout 120 $P/ntxent_dist_benchmark --gpus 8 --exchange $ex >> gpurun_out/r2_h_cpp_dist.jsonl 2>&1
done
timeout 120 $P/ntxent_dist_benchmark --gpus 8 --M 131072 --D 256 --steps 10 --warmup 3 >> gpurun_out/r2_h_cpp_dist.jsonl 2>&1
cat gpurun_out/r2_h_cpp_dist.jsonl
