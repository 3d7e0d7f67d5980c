"""In-situ duration of every kernel of the fused step (warm L2, back-to-back launches), from the
CUPTI activity records torch.profiler collects: what `ncu` cannot show, since it serialises the
launches and flushes the caches between them.

    python tools/kernel_times.py [--M 32768] [--D 128] [--tau 0.1] [--steps 20] [--dtype bf16]
    NTXENT_NO_PDL=1 python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 \
        tools/kernel_times.py --dist [--comm push|nccl]      # the row-partitioned step, rank 0 prints

(With programmatic dependent launch a kernel is resident while it still waits for its predecessor,
and that wait counts into its duration: NTXENT_NO_PDL=1 gives the kernels' own times.  Under --dist a
kernel that waits for a peer's flag counts that wait as well: the table is a critical-path view.)
"""
import argparse
import collections
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from torch.profiler import ProfilerActivity, profile

import ntxent_b200  # noqa: F401
from ntxent_b200 import capi

ap = argparse.ArgumentParser()
ap.add_argument("--M", type=int, default=32768)
ap.add_argument("--D", type=int, default=128)
ap.add_argument("--tau", type=float, default=0.1)
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--dtype", default="bf16")
ap.add_argument("--flags", type=int, default=0)
ap.add_argument("--dist", action="store_true")
ap.add_argument("--comm", default="push")
a = ap.parse_args()
dt = torch.bfloat16 if a.dtype == "bf16" else torch.float32
rank, world, engine = 0, 1, None
if a.dist:
    import torch.distributed as dist
    from ntxent_b200.distributed import RowPartitionedNTXent
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank)))
    dist.init_process_group("nccl")
    engine = RowPartitionedNTXent(a.M, a.D, a.tau, dt, torch.device("cuda", torch.cuda.current_device()), comm=a.comm)
z = torch.from_numpy(np.random.default_rng(2).standard_normal((a.M, a.D)).astype(np.float32)).cuda().to(dt)
if engine is None:
    ws = capi.workspace(a.M, a.D, z, a.flags)
    out = (torch.empty(1, device="cuda"), torch.empty(a.M, device="cuda"), torch.empty_like(z))

    def step():
        capi.forward_backward(z, a.tau, None, False, a.flags, ws=ws, out=out)
else:
    shard = z[rank * (a.M // world):(rank + 1) * (a.M // world)].contiguous()

    def step():
        engine.forward_backward(shard)
for _ in range(5):
    step()
torch.cuda.synchronize()
if engine is not None:
    dist.barrier()
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(a.steps):
        step()
    torch.cuda.synchronize()
if engine is not None:
    dist.barrier()
    engine.release()
    dist.destroy_process_group()
    if rank != 0:
        sys.exit(0)
    print(f"rank 0 of {world}, comm={a.comm}, {a.M // world} rows x {a.M} columns")
ev = [e for e in prof.events() if e.device_type == torch.autograd.DeviceType.CUDA]
ev.sort(key=lambda e: e.time_range.start)
by = collections.OrderedDict()
for e in ev:
    by.setdefault(e.name, []).append(e.time_range.end - e.time_range.start)
total = 0.0
for k, v in by.items():
    m = sum(v) / len(v)
    total += m * len(v) / a.steps
    print(f"{m:9.2f} us x{len(v) / a.steps:4.1f}/step  {k[:110]}")
span = (ev[-1].time_range.end - ev[0].time_range.start) / a.steps
print(f"sum of kernels {total:.1f} us/step; wall per step (first start to last end) {span:.1f} us; "
      f"gaps {span - total:.1f} us/step")
