#!/bin/bash
# GPU run F (2 x B200): multi-GPU tests (communicator, engine), bench lines at N = 2 for the three
# exchange forms and for C5, the C++ multi-GPU benchmark; then single-GPU A/B of the packed backward.
mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/r2_f_topo.log 2>&1
timeout 900 python -m pytest tests/test_comm_gpu.py tests/test_multigpu_gpu.py tests/test_callers_gpu.py -m gpu -x -q > gpurun_out/r2_f_tests.log 2>&1; echo "tests rc=$?"
tail -15 gpurun_out/r2_f_tests.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533"
B="--gpus 2 --no-sustained"
summ() { python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    d=json.load(open(f"gpurun_out/r2_f_{n}.json")); r=d["roofline"]
    print(f"{n:18s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} e2e {d['e2e']['ms_per_step']:.4f} parity {d.get('parity')}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
timeout 300 $TR bench.py $B > gpurun_out/r2_f_n2_push.json 2> gpurun_out/r2_f_bench.err; summ n2_push
timeout 300 $TR bench.py $B --comm nccl > gpurun_out/r2_f_n2_nccl.json 2>> gpurun_out/r2_f_bench.err; summ n2_nccl
timeout 300 $TR bench.py $B --comm nccl --bwd-mode contract > gpurun_out/r2_f_n2_contract.json 2>> gpurun_out/r2_f_bench.err; summ n2_contract
timeout 300 $TR bench.py $B --comm torch > gpurun_out/r2_f_n2_torch.json 2>> gpurun_out/r2_f_bench.err; summ n2_torch
timeout 300 $TR bench.py $B --config c5 --steps 5 --warmup 3 > gpurun_out/r2_f_n2_c5.json 2>> gpurun_out/r2_f_bench.err; summ n2_c5
tail -5 gpurun_out/r2_f_bench.err
P=cuda-nt-xent-mpi-nccl-simclr_b200
for ex in "push" "nccl" "nccl --bwd contract"; do
  timeout 120 $P/ntxent_dist_benchmark --gpus 2 --exchange $ex >> gpurun_out/r2_f_cpp_dist.jsonl 2>&1
done
cat gpurun_out/r2_f_cpp_dist.jsonl
S="--no-cpu-baseline --no-sustained"
timeout 300 python bench.py $S > gpurun_out/r2_f_n1.json 2>> gpurun_out/r2_f_bench.err; summ n1
NTXENT_LIBRARY=$PWD/$P/libntxent_b200_bwdunpacked.so timeout 300 python bench.py $S > gpurun_out/r2_f_n1_bwdunpacked.json 2>> gpurun_out/r2_f_bench.err; summ n1_bwdunpacked
