"""One GPU, the work of every rank of an N-way row partition in turn: the row-block entry points
(normalize -> forward_rows -> backward_rows_dz) against the whole-batch call, shard by shard, and the
whole-batch call with and without the in-kernel backward epilogue.  Used to localise a parity
difference seen at N = 8 without spending 8 GPUs on it.

    python tools/shard_check.py [--M 32768] [--D 128] [--world 8] [--repeat 3]
"""
import argparse
import os
import subprocess
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch

ap = argparse.ArgumentParser()
ap.add_argument("--M", type=int, default=32768)
ap.add_argument("--D", type=int, default=128)
ap.add_argument("--tau", type=float, default=0.1)
ap.add_argument("--world", type=int, default=8)
ap.add_argument("--repeat", type=int, default=3)
ap.add_argument("--dump", default="")
a = ap.parse_args()

import ntxent_b200  # noqa: E402,F401
from ntxent_b200 import capi  # noqa: E402

z = torch.from_numpy(np.random.default_rng(7).standard_normal((a.M, a.D)).astype(np.float32)).cuda().to(torch.bfloat16)
loss, lse, dz = capi.forward_backward(z, a.tau)
torch.cuda.synchronize()
if a.dump:  # child process (other environment): just leave the whole-batch result behind
    torch.save({"loss": loss.cpu(), "dz": dz.cpu()}, a.dump)
    sys.exit(0)
scale = dz.float().abs().max()
for rep in range(a.repeat):  # run-to-run: the whole-batch call must repeat itself bit for bit
    l2, _, dz2 = capi.forward_backward(z, a.tau)
    torch.cuda.synchronize()
    print(f"whole batch, repeat {rep}: bitwise equal = {torch.equal(dz, dz2)}, "
          f"max diff / max|dz| = {((dz2.float() - dz.float()).abs().max() / scale).item():.3e}")
# the other epilogue form in a child process
tmp = "/tmp/shard_check_other.pt"
env = dict(os.environ)
other = "no tail" if env.get("NTXENT_BWD_TAIL") == "1" else "tail"
if other == "no tail":
    env.pop("NTXENT_BWD_TAIL")
else:
    env["NTXENT_BWD_TAIL"] = "1"
subprocess.run([sys.executable, __file__, "--M", str(a.M), "--D", str(a.D), "--tau", str(a.tau), "--dump", tmp],
               env=env, check=True)
o = torch.load(tmp)
d = (o["dz"].cuda().float() - dz.float()).abs()
print(f"whole batch vs the same with '{other}': max diff / max|dz| = {(d.max() / scale).item():.3e}, "
      f"rows differing by > 1e-3: {(d.max(dim=1).values > 1e-3 * scale).sum().item()}")
# the row partition
u, inv = capi.normalize(z)
lse_all, _ = capi.forward_rows(u, 0, a.M, a.tau)
rows = a.M // a.world
for r in range(a.world):
    zr = z[r * rows:(r + 1) * rows].contiguous()
    dzr = capi.backward_rows_dz(u, lse_all, r * rows, rows, a.tau, zr, inv[r * rows:(r + 1) * rows].contiguous())
    torch.cuda.synchronize()
    d = (dzr.float() - dz[r * rows:(r + 1) * rows].float()).abs()
    bad = (d.max(dim=1).values > 1e-3 * scale).nonzero().flatten()
    print(f"shard {r}: max diff / max|dz| = {(d.max() / scale).item():.3e}, rows > 1e-3: {bad.numel()}"
          + (f" (first {bad[:6].tolist()}, row tiles {sorted(set((bad // 128).tolist()))[:12]})" if bad.numel() else ""))
