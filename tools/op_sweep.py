#!/usr/bin/env python3
"""Operator sweep through `torch.ops.ntxent_cuda` -- the role of the reference's python/test.py.

Same stages as python/test.py (stability sweep :57-79, forward timing sweep over
B in {32..512} x D in {64..512} with and without mixed precision :81-130,141-158, memory log
:25-40,178, JSON report :196-205), written for this library:
  * the op is reached exactly the way the reference script reaches it (`torch.ops.ntxent_cuda`),
    after loading the built extension;
  * "AMP" means use_mixed_precision=True (bf16 operands) -- the reference wraps the call in
    autocast, which does nothing for a custom op;
  * timing uses CUDA events, forward+backward is timed as well, and peak memory shows the O(M*D)
    footprint instead of the reference's three [M,M] fp32 temporaries;
  * the report is valid JSON (the reference dumps a dict with tuple keys and crashes).

    python tools/op_sweep.py [--runs 100] [--out benchmark_results] [--quick]
"""
import argparse
import json
import logging
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

logging.basicConfig(level=logging.INFO, format="%(asctime)s %(levelname)s %(message)s")
log = logging.getLogger("ntxent.sweep")


def load_op():
    import ntxent_b200  # noqa: F401  (import shim)
    from ntxent_b200 import _build
    torch.ops.load_library(_build.build_binding())
    return torch.ops.ntxent_cuda


class MemoryLog:
    def __init__(self):
        self.rows = []

    def mark(self, step):
        self.rows.append({"step": step, "allocated_mb": torch.cuda.memory_allocated() / 2**20,
                          "reserved_mb": torch.cuda.memory_reserved() / 2**20,
                          "peak_allocated_mb": torch.cuda.max_memory_allocated() / 2**20})
        log.info("memory at %s: %.1f MB allocated, %.1f MB peak", step, self.rows[-1]["allocated_mb"],
                 self.rows[-1]["peak_allocated_mb"])


def stability_sweep(op, rows=128, dim=256):
    """Unit rows times {1e-5, 1, 1e5} at T in {0.01, 0.07, 1}: loss and gradient stay finite."""
    g = torch.Generator(device="cuda").manual_seed(0)
    z = torch.nn.functional.normalize(torch.randn(rows, dim, device="cuda", generator=g))
    ok = True
    for scale in (1e-5, 1.0, 1e5):
        for temp in (0.01, 0.07, 1.0):
            for mp in (False, True):
                zz = (z * scale).requires_grad_(True)
                loss = op.forward(zz, temp, mp)
                loss.backward()
                fine = bool(torch.isfinite(loss).all() and torch.isfinite(zz.grad).all())
                if not fine:
                    log.error("not finite: scale=%g T=%g mixed=%s", scale, temp, mp)
                ok &= fine
    return ok


def timed(fn, warmup, runs):
    for _ in range(warmup):
        fn()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(runs)]
    torch.cuda.synchronize()
    for a, b in ev:
        a.record()
        fn()
        b.record()
    torch.cuda.synchronize()
    t = np.array([a.elapsed_time(b) for a, b in ev])
    return {"mean_ms": float(t.mean()), "std_ms": float(t.std()), "min_ms": float(t.min()), "max_ms": float(t.max())}


def perf_case(op, rows, dim, mixed, runs):
    z = torch.nn.functional.normalize(torch.randn(rows, dim, device="cuda"))
    one = torch.ones(1, device="cuda")
    torch.cuda.reset_peak_memory_stats()
    base = torch.cuda.memory_allocated()
    fwd = timed(lambda: op.forward(z, 0.07, mixed), 10, runs)

    def both():
        loss, state = op.forward_with_state(z, 0.07, mixed)
        op.backward(z, state, one, 0.07, mixed)
    fb = timed(both, 10, runs)
    peak = (torch.cuda.max_memory_allocated() - base) / 2**20
    return {"forward": fwd, "forward_backward": fb, "peak_extra_memory_mb": peak,
            "tflops_fwd_bwd": 8.0 * rows * rows * dim / (fb["mean_ms"] * 1e-3) / 1e12,
            "reference_temporaries_mb": 3 * 4 * (2 * rows) ** 2 / 2**20}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--runs", type=int, default=100)
    ap.add_argument("--out", default="benchmark_results")
    ap.add_argument("--quick", action="store_true")
    args = ap.parse_args()
    if not torch.cuda.is_available():
        raise SystemExit("CUDA is not available")
    cap = torch.cuda.get_device_capability()
    log.info("GPU %s (sm_%d%d), CUDA %s, PyTorch %s", torch.cuda.get_device_name(), cap[0], cap[1],
             torch.version.cuda, torch.__version__)
    op = load_op()
    if not op.check_tensor_core_support():
        raise SystemExit("this device cannot run the tcgen05 kernels (needs sm_100)")
    mem = MemoryLog()
    mem.mark("extension loaded")
    if not stability_sweep(op):
        raise SystemExit("numerical stability sweep failed")
    mem.mark("after stability sweep")
    batch_sizes = [32, 512] if args.quick else [32, 64, 128, 256, 512]
    dims = [128] if args.quick else [64, 128, 256, 512]
    results = []
    for b in batch_sizes:
        for d in dims:
            row = {"rows": b, "dim": d, "fp32": perf_case(op, b, d, False, args.runs),
                   "mixed": perf_case(op, b, d, True, args.runs)}
            results.append(row)
            log.info("rows=%d D=%d  fp32 fwd %.3f ms  mixed fwd %.3f ms  fwd+bwd %.3f / %.3f ms  peak +%.2f MB "
                     "(reference would hold %.1f MB of [M,M] temporaries)", b, d, row["fp32"]["forward"]["mean_ms"],
                     row["mixed"]["forward"]["mean_ms"], row["fp32"]["forward_backward"]["mean_ms"],
                     row["mixed"]["forward_backward"]["mean_ms"], row["mixed"]["peak_extra_memory_mb"],
                     row["mixed"]["reference_temporaries_mb"])
            mem.mark(f"after rows={b} D={d}")
    os.makedirs(args.out, exist_ok=True)
    stamp = time.strftime("%Y%m%d_%H%M%S")
    path = os.path.join(args.out, f"results_{stamp}.json")
    with open(path, "w") as f:
        json.dump({"device": torch.cuda.get_device_name(), "results": results, "memory": mem.rows}, f, indent=2)
    with open(os.path.join(args.out, "memory_profile.json"), "w") as f:
        json.dump(mem.rows, f, indent=2)
    log.info("results saved to %s", path)
    return 0


if __name__ == "__main__":
    sys.exit(main())
