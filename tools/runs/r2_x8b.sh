#!/bin/bash
# GPU run X8b (8 x B200): the C4 push line again (the 1-GPU reference its parity key compares with was
# wrong in run X8: in-kernel epilogue without a barrier, fixed since).
mkdir -p gpurun_out
timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 \
    bench.py --gpus 8 --no-sustained > gpurun_out/r2_final_n8_push.json 2> gpurun_out/r2_x8b.err
python - <<'PY'
import json
txt=open("gpurun_out/r2_final_n8_push.json").read()
d=json.loads([l for l in txt.splitlines() if l.startswith("{")][-1]); r=d["roofline"]
print(f"step {d['ms_per_step']:.4f} fwd {r['fwd_kernel']['ms_per_launch']:.4f} bwd {r['ms_per_launch']:.4f} side {r['side_kernels_ms']:.4f} e2e {d['e2e']['ms_per_step']:.4f} parity {d.get('parity')}")
PY
