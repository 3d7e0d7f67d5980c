#!/bin/bash
# GPU run L (2 x B200): A/B of the launch-saving variants at N = 2 (C4 and a 4096-row-shard stand-in).
mkdir -p gpurun_out
summ() { python - "$1" <<'PY'
import json,sys
n=sys.argv[1]
try:
    txt=open(f"gpurun_out/r2_l_{n}.json").read()
    d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1]); r=d["roofline"]
    print(f"{n:22s} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f}")
except Exception as e:
    print(n, "FAILED", e)
PY
}
trun() { # name, args...
  name=$1; shift
  timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 \
      bench.py --gpus 2 --no-sustained "$@" > gpurun_out/r2_l_$name.json 2>> gpurun_out/r2_l_bench.err
  summ $name
}
for m in 8192 32768; do
  trun M${m}_default --M $m
  NTXENT_NO_NORM_TICKET=1 trun M${m}_signalkernel --M $m
  NTXENT_FWD_TAIL=1 trun M${m}_fwdtail --M $m
  trun M${m}_default2 --M $m
done
grep -v "^\*\|^$\|OMP_NUM\|NCCL version" gpurun_out/r2_l_bench.err | tail -5
