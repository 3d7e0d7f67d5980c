"""Debug aid: tf32 tcgen05 path (fp32 operands, M >= 1024) vs the exact fp32-FMA path."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ntxent_b200
from ntxent_b200 import capi

def run(M, D, tau, seed=0):
    z = torch.from_numpy(np.random.default_rng(seed).standard_normal((M, D)).astype(np.float32)).cuda()
    print(f"--- M={M} D={D} tau={tau} plan={capi.plan_info(M, D, M, capi.F32, 0)}", flush=True)
    ls, lses = capi.forward(z, tau, flags=capi.FLAG_FORCE_SIMT)
    lt, lset = capi.forward(z, tau)
    torch.cuda.synchronize()
    print(f"fwd loss simt {ls.item():.7f} tc {lt.item():.7f} rel {abs(ls.item()-lt.item())/ls.item():.2e} lse maxdiff {(lses-lset).abs().max().item():.3e}", flush=True)
    dzs, _ = capi.backward(z, lses, None, tau, flags=capi.FLAG_FORCE_SIMT)
    dzt, _ = capi.backward(z, lses, None, tau)
    torch.cuda.synchronize()
    d = (dzs - dzt).abs().max().item() / dzs.abs().max().item()
    print(f"bwd dz rel maxdiff {d:.3e}  (nan: {torch.isnan(dzt).any().item()})", flush=True)

for c in [(1024, 128, 0.5), (1024, 64, 0.5), (2048, 128, 0.1), (1030, 128, 0.07), (8192, 128, 0.5), (4096, 64, 0.02)]:
    run(*c)
u = capi.normalize(torch.randn(2048, 128, device="cuda"))[0]
lse_all = capi.forward_rows(u, 0, 2048, 0.2, flags=capi.FLAG_FORCE_SIMT)[0]
for mode in (capi.BWD_SYMMETRIC, capi.BWD_CONTRACT):
    a = capi.backward_rows(u, lse_all, 512, 512, 0.2, mode)
    b = capi.backward_rows(u, lse_all, 512, 512, 0.2, mode, flags=capi.FLAG_FORCE_SIMT)
    print("rows bwd mode", mode, ((a[0] - b[0]).abs().max() / b[0].abs().max()).item(),
          None if a[1] is None else ((a[1] - b[1]).abs().max() / b[1].abs().max()).item())
M, D, tau = 8192, 128, 0.5
z = torch.randn(M, D, device="cuda")
ws = capi.workspace(M, D, z, 0)
for _ in range(3): capi.forward_backward(z, tau, ws=ws)
torch.cuda.synchronize(); capi.profile_begin(); t0 = time.perf_counter()
for _ in range(10): capi.forward_backward(z, tau, ws=ws)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10; p = capi.profile_end()
print(f"C3 fwd+bwd {dt*1e3:.3f} ms  fwd {p['fwd_ms']/10:.3f} ms bwd {p['bwd_ms']/10:.3f} ms -> {8*M*M*D/dt/1e12:.1f} eff TFLOP/s")
