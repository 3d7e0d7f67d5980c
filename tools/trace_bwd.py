"""Prints the pipeline timeline of CTA 0 of the tcgen05 backward kernel (clock64 deltas).

Needs the trace build of the library; build it where nvcc is cheap (not on the GPU box):
    python -c "import ntxent_b200; from ntxent_b200 import _build; _build.build_library(trace=True)"
"""
import sys, os, ctypes
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
os.environ["NTXENT_TRACE"] = "1"  # load libntxent_b200_trace.so (build it first: see above)
import ntxent_b200
from ntxent_b200 import capi
M, D, tau = 32768, 128, 0.1
z = torch.from_numpy(np.random.default_rng(2).standard_normal((M, D)).astype(np.float32)).cuda().to(torch.bfloat16)
ws = capi.workspace(M, D, z, 0)
for _ in range(2):
    capi.forward_backward(z, tau, ws=ws)
torch.cuda.synchronize()
trace = torch.zeros(5 * 64 * 8, dtype=torch.int64, device="cuda")
capi.lib().ntxent_b200_debug_set_trace(ctypes.c_void_p(trace.data_ptr()))
capi.forward_backward(z, tau, ws=ws)
torch.cuda.synchronize()
capi.lib().ntxent_b200_debug_set_trace(None)
t = trace.cpu().numpy().reshape(5, 64, 8)
t0 = t[0, 0, 0]
names = ["MMA", "W0 ", "W8 "]
ev = {0: ["S:start", "kfull", "sfree", "S:issued", "PV:start", "pfull", "PV:issued"],
      1: ["start", "sfull", "ld", "computed", "pfree", "arrived"], 2: ["start", "sfull", "ld", "computed", "pfree", "arrived"]}
for it in list(range(0, 6)) + list(range(20, 26)):
    for r in range(3):
        vals = [(ev[r][e], int(t[r, it, e] - t0)) for e in range(len(ev[r])) if t[r, it, e]]
        print(f"tile {it:2d} {names[r]}: " + "  ".join(f"{n}={v}" for n, v in vals))
for r in (1, 2):  # mean phase lengths of the weight warps over tiles 10..40
    d = t[r, 10:41].astype(np.int64)
    ph = [("wait s_full", d[:, 1] - d[:, 0]), ("tmem ld + s_free", d[:, 2] - d[:, 1]), ("weights", d[:, 3] - d[:, 2]),
          ("fence", d[:, 4] - d[:, 3]), ("st + p_full", d[:, 5] - d[:, 4]), ("loop back", d[1:, 0] - d[:-1, 5])]
    print(names[r], "phases:", "  ".join(f"{n} {v.mean():.0f}" for n, v in ph))
print("W8 - W0 at 'sfull' (tiles 10..40):", (t[2, 10:41, 1] - t[1, 10:41, 1]).tolist())
per_tile = (t[1, 40, 5] - t[1, 10, 5]) / 30
print("steady-state clk per tile (W0 arrivals, tiles 10..40):", per_tile)

# unit-level timeline of CTA 0 (weight warp 0)
k = t[4, 63]
cyc = int(k[2] - k[0]); ns = int(k[7] - k[6])
print(f"kernel (CTA 0): {cyc} clk in {ns} ns -> {cyc / max(ns, 1):.3f} GHz; set-up {int(k[1] - k[0])} clk")
for un in range(8):
    r = t[4, un]
    if not r[0]:
        break
    print(f"unit {un}: tiles={int(r[7])} start={int(r[0] - k[0])} first_S={int(r[1] - r[0])} tiles_done={int(r[2] - r[0])} "
          f"q_parked=+{int(r[3] - r[2])} acc_full=+{int(r[4] - r[3])} drained=+{int(r[5] - r[4])} "
          f"clk/tile={(int(r[2] - r[1])) / max(int(r[7]), 1):.1f}")
