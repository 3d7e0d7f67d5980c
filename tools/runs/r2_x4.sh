#!/bin/bash
# GPU run X4 (4 x B200): the C4 push and C5 push lines with the final kernels.
mkdir -p gpurun_out
for cfg in "push" "c5 --M 131072 --D 256"; do
  set -- $cfg; name=$1; shift
  timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29561 \
      bench.py --gpus 4 --no-sustained "$@" > gpurun_out/r2_final_n4_$name.json 2>> gpurun_out/r2_x4.err
  python - $name <<'PY'
import json,sys
txt=open(f"gpurun_out/r2_final_n4_{sys.argv[1]}.json").read()
d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1]); r=d["roofline"]
print(f"{sys.argv[1]} step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f} parity {d['parity']['dz_vs_1gpu_maxrel']:.2e}")
PY
done
