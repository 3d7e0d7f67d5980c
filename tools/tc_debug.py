"""Debug aid: tcgen05 path vs fp32-FMA path on the same bf16 operands (run on the GPU box)."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import ntxent_b200
from ntxent_b200 import capi

def run(M, D, tau, seed=0, flags=0):
    z = torch.from_numpy(np.random.default_rng(seed).standard_normal((M, D)).astype(np.float32)).cuda().to(torch.bfloat16)
    print(f"--- M={M} D={D} tau={tau} flags={flags} plan={capi.plan_info(M, D, M, capi.BF16, flags)}", flush=True)
    ls, lses = capi.forward(z, tau, flags=capi.FLAG_FORCE_SIMT)
    torch.cuda.synchronize()
    lt, lset = capi.forward(z, tau, flags=flags)
    torch.cuda.synchronize()
    print(f"fwd loss simt {ls.item():.6f} tc {lt.item():.6f}  lse maxdiff {(lses-lset).abs().max().item():.3e}", flush=True)
    dzs, _ = capi.backward(z, lses, None, tau, flags=capi.FLAG_FORCE_SIMT)
    torch.cuda.synchronize()
    dzt, _ = capi.backward(z, lses, None, tau, flags=flags)
    torch.cuda.synchronize()
    d = (dzs.float() - dzt.float()).abs().max().item() / dzs.float().abs().max().item()
    print(f"bwd dz rel maxdiff {d:.3e}  (nan: {torch.isnan(dzt).any().item()})", flush=True)
    return d

if __name__ == "__main__":
    cases = [(128, 64, 0.5), (128, 128, 0.5), (256, 128, 0.1), (1024, 128, 0.1), (130, 128, 0.1), (16, 64, 0.07),
             (384, 256, 0.07), (512, 192, 0.2), (4096, 128, 0.1), (1000, 64, 0.02)]
    if len(sys.argv) > 1:
        cases = cases[:int(sys.argv[1])]
    for c in cases:
        run(*c)
    # contract-mode pieces
    M, D, tau = 1024, 128, 0.1
    z = torch.from_numpy(np.random.default_rng(1).standard_normal((M, D)).astype(np.float32)).cuda().to(torch.bfloat16)
    u, inv = capi.normalize(z)
    lse, lp = capi.forward_rows(u, 256, 256, tau)
    lse_s, lp_s = capi.forward_rows(u, 256, 256, tau, flags=capi.FLAG_FORCE_SIMT)
    print("rows fwd", (lse - lse_s).abs().max().item(), lp.item(), lp_s.item())
    lse_all = capi.forward_rows(u, 0, M, tau, flags=capi.FLAG_FORCE_SIMT)[0]
    for mode in (capi.BWD_SYMMETRIC, capi.BWD_CONTRACT):
        a = capi.backward_rows(u, lse_all, 256, 256, tau, mode)
        b = capi.backward_rows(u, lse_all, 256, 256, tau, mode, flags=capi.FLAG_FORCE_SIMT)
        print("rows bwd mode", mode, ((a[0] - b[0]).abs().max() / b[0].abs().max()).item(),
              None if a[1] is None else ((a[1] - b[1]).abs().max() / b[1].abs().max()).item())
    # timing at C4
    M, D, tau = 32768, 128, 0.1
    z = torch.from_numpy(np.random.default_rng(2).standard_normal((M, D)).astype(np.float32)).cuda().to(torch.bfloat16)
    ws = capi.workspace(M, D, z, 0)
    for _ in range(3):
        capi.forward_backward(z, tau, ws=ws)
    torch.cuda.synchronize()
    capi.profile_begin()
    t0 = time.perf_counter()
    for _ in range(10):
        capi.forward_backward(z, tau, ws=ws)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / 10
    p = capi.profile_end()
    print(f"C4 fwd+bwd {dt*1e3:.3f} ms  fwd {p['fwd_ms']/10:.3f} ms bwd {p['bwd_ms']/10:.3f} ms  -> {8*M*M*D/dt/1e12:.1f} eff TFLOP/s", flush=True)
