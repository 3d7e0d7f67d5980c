#!/usr/bin/env python
"""bench.py -- NT-Xent fused forward+backward throughput on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl native|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

A "step" is one fused forward+backward of the NT-Xent loss over one synthetic batch
z ~ N(0,1), [M, D], generated once on the host with seed 1234.  Default workload is
BASELINE.json configs[3] (2N = 32768, D = 128, bf16, tau = 0.1); it fits one GPU, and for
N > 1 the same global batch is row-partitioned across the ranks (strong scaling).

Reported numbers (one JSON line on rank 0):
  value      effective TFLOP/s = 8 M^2 D / step time (BASELINE.md section 3's work model: forward
             2M^2D + recompute 2M^2D + the two backward GEMMs 4M^2D), inputs resident in HBM.
  e2e        same metric through the public op with HOST input: pinned z -> H2D -> forward+
             backward -> loss D2H, all inside the timed region.
  roofline   dominant kernel (the backward tile pass) against the measured dense bf16 peak,
             from EXECUTED tensor FLOPs: the symmetric backward executes 4 M^2 D (S recompute +
             W.U), so achieved = 4 M^2 D / its CUDA-event time.
  cpu_baseline  the oracle's fp32/OpenMP row-block form (oracle/ntxent_oracle.c) on the host
             cores for a bounded row sample of the same problem.

--impl reference times that CPU path alone (the reference has no CPU NT-Xent and its CUDA
code does not compile -- SURVEY.md sections 0.1, 0.2 -- so the reference arm is the oracle port).
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

CONFIGS = {
    # name: (M, D, tau, dtype)  -- BASELINE.json configs[0..4]
    "c1": (512, 128, 0.5, "fp32"),
    "c2": (4096, 128, 0.1, "bf16"),
    "c3": (8192, 128, 0.5, "fp32"),
    "c4": (32768, 128, 0.1, "bf16"),
    "c5": (131072, 256, 0.1, "bf16"),
}
METRIC = "ntxent_fwd_bwd_tflops"
UNIT = "TFLOP/s"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--config", default="c4", choices=sorted(CONFIGS))
    ap.add_argument("--M", type=int, default=0)
    ap.add_argument("--D", type=int, default=0)
    ap.add_argument("--tau", type=float, default=0.0)
    ap.add_argument("--dtype", default="", choices=["", "bf16", "fp32"])
    ap.add_argument("--simt", action="store_true", help="force the fp32-FMA tile kernels")
    ap.add_argument("--bwd-mode", default="symmetric", choices=["symmetric", "contract"],
                    help="multi-GPU backward formulation")
    ap.add_argument("--comm", default="symm", choices=["symm", "symm-barrier", "nccl"],
                    help="multi-GPU exchange: kernels push into the peers' symmetric memory, or NCCL collectives")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--graph", action="store_true", help="single GPU: replay the step as a CUDA graph")
    ap.add_argument("--no-graph", action="store_true",
                    help="multi-GPU: issue every launch from the host instead of replaying a CUDA graph")
    return ap.parse_args()


def workload(args):
    M, D, tau, dt = CONFIGS[args.config]
    M = args.M or M
    D = args.D or D
    tau = args.tau or tau
    dt = args.dtype or dt
    return M, D, tau, dt


def make_input(M, D):
    import numpy as np
    return np.random.default_rng(1234).standard_normal((M, D), dtype=np.float32)


def ncu_dram_traffic(kernel_tag):
    """dram__bytes_read.sum + dram__bytes_write.sum of one launch of the dominant kernel at the C4
    workload, from the committed `ncu --set full` summary (profiles/); None if it is not there."""
    path = os.path.join(ROOT, "profiles", f"r1_ncu_{kernel_tag}_summary.csv")
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    total, seen = 0.0, 0
    try:
        for line in open(path):
            f = line.strip().split(",")
            if len(f) == 3 and f[0] in ("dram__bytes_read.sum", "dram__bytes_write.sum"):
                total += float(f[2]) * scale.get(f[1], 1.0)
                seen += 1
    except OSError:
        return None
    return total if seen == 2 else None


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        with open(path) as f:
            return json.load(f), "measured"
    # fallback stated in /opt/skills/guides/B200_PROFILING.md
    return {"bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "hbm_gbs": 6650.0}, "fallback"


# ------------------------------------------------------------------------ clocks
class ClockSampler:
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int):
        self.idx = gpu_index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.idx), f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                 "-lms", "100"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": [], "samples": 0}
        if self.proc is None:
            return out
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        try:
            for line in open(self.path):
                f = [x.strip() for x in line.split(",")]
                if len(f) < 9:
                    continue
                try:
                    sm.append(float(f[1])); mx.append(float(f[2])); power.append(float(f[3]))
                except ValueError:
                    continue
                for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown",
                                      "sw_power_cap"), f[5:9]):
                    if val.lower().startswith("active"):
                        reasons.add(name)
            os.unlink(self.path)
        except Exception:
            pass
        if sm:
            out.update(sm_mhz=statistics.median(sm), sm_max_mhz=max(mx), reasons=sorted(reasons),
                       samples=len(sm), power_w_max=max(power))
        return out


# ------------------------------------------------------------------ CPU baseline
def cpu_rowblock_tflops(z, tau, seconds, rows_hint=0):
    """Times the oracle's fp32 row-block form on a bounded row sample; returns (tflops, rows, s)."""
    from oracle import oracle as O
    O.build()
    M, D = z.shape
    rows = rows_hint
    if not rows:
        probe = max(16, min(M, 64))
        for _ in range(3):  # calibrate in growing steps (the first call also spins up the threads)
            t0 = time.perf_counter()
            O.rowblock_f32(z, tau, 0, probe)
            dt = max(time.perf_counter() - t0, 1e-6)
            if dt > 0.5 or probe >= M:
                break
            probe = int(min(M, max(probe * 2, probe * 1.0 / dt)))
        rows = int(min(M, max(16, probe * seconds / dt)))
        rows = max(16, rows // 16 * 16)
    t0 = time.perf_counter()
    O.rowblock_f32(z, tau, 0, rows)
    dt = time.perf_counter() - t0
    return 8.0 * rows * M * D / dt / 1e12, rows, dt, O.num_threads()


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    M, D, tau, dt = workload(args)
    z = make_input(M, D)
    rows = 0
    per_step = max(1.0, min(10.0, 120.0 / max(1, args.steps + args.warmup)))
    for _ in range(args.warmup):
        _, rows, _, cores = cpu_rowblock_tflops(z, tau, per_step, rows)
    times = []
    cores = 1
    for _ in range(args.steps):
        _, rows, t, cores = cpu_rowblock_tflops(z, tau, per_step, rows)
        times.append(t)
    total = sum(times)
    val = 8.0 * rows * M * D * len(times) / total / 1e12
    sample = (f"rows [0,{rows}) of the M={M} problem per step, contract row-block form "
              f"(fwd + recompute + G.U + G^T.U_I = 8*rows*M*D FLOP), fp32 OpenMP")
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": {"workload": f"ntxent_fwd_bwd_M{M}_D{D}_tau{tau}_{dt}",
                                        "M": M, "D": D, "tau": tau},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference has no CPU NT-Xent and its CUDA sources do not compile; this is "
                "the oracle port on the host cores",
    }
    print(json.dumps(line))
    return 0


# ------------------------------------------------------------------------ native
def run_native(args):
    import torch
    import torch.distributed as dist

    import ntxent_b200
    from ntxent_b200 import capi

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the NT-Xent engine has no CPU path")
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    if args.gpus != world and rank == 0:
        print(f"[bench] --gpus {args.gpus} but WORLD_SIZE={world}; using {world}", file=sys.stderr)

    M, D, tau, dt = workload(args)
    tdtype = torch.bfloat16 if dt == "bf16" else torch.float32
    flags = capi.FLAG_FORCE_SIMT if args.simt else 0
    z_host_np = make_input(M, D)
    z_host_full = torch.from_numpy(z_host_np).to(tdtype)
    rows = M // world
    assert M % world == 0, "global batch must divide by the number of ranks"
    z_host = z_host_full[rank * rows:(rank + 1) * rows].contiguous().pin_memory()
    z_dev = z_host.to(dev)
    flush = None if args.no_flush else torch.empty(256 * 2**20, dtype=torch.uint8, device=dev)

    if world == 1:
        ws = capi.workspace(M, D, z_dev, flags)
        out = (torch.empty(1, dtype=torch.float32, device=dev),
               torch.empty(M, dtype=torch.float32, device=dev), torch.empty_like(z_dev))

        graphed = ntxent_b200.GraphedNTXent(M, D, tau, tdtype, dev) if args.graph else None

        def step(zin):
            if graphed is not None:
                return graphed(zin)[0]
            return capi.forward_backward(zin, tau, None, False, flags, ws, out)[0]
        launches_per_step = 6  # normalise, forward tiles, finalize, loss reduce, backward tiles, epilogue
    else:
        from ntxent_b200.distributed import RowPartitionedNTXent
        comm_used, comm_note = args.comm, None
        try:
            engine = RowPartitionedNTXent(M, D, tau, tdtype, dev, backward_mode=args.bwd_mode, flags=flags,
                                          use_cuda_graph=not args.no_graph, comm=args.comm)
        except Exception as exc:  # no peer mappings on this box: every rank fails the same rendezvous
            if args.comm == "nccl":
                raise
            comm_used, comm_note = "nccl", f"{args.comm} unavailable: {type(exc).__name__}: {str(exc)[:120]}"
            engine = RowPartitionedNTXent(M, D, tau, tdtype, dev, backward_mode=args.bwd_mode, flags=flags,
                                          use_cuda_graph=not args.no_graph, comm="nccl")

        def step(zin):
            return engine.forward_backward(zin)[0]
        launches_per_step = engine.launches_per_step

    def sync_all():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(fn, steps):
        """Per-step CUDA events on the launching stream; optional L2 flush between steps."""
        evs = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True))
               for _ in range(steps)]
        sync_all()
        for a, b in evs:
            if flush is not None:
                flush.fill_(1)
            a.record()
            fn()
            b.record()
        sync_all()
        return [a.elapsed_time(b) for a, b in evs]

    for _ in range(max(args.warmup, 3)):
        step(z_dev)
    sync_all()

    # --- device-resident timing, with per-kernel events and a clock record --------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    graph_mode = (world > 1 and not args.no_graph) or (world == 1 and args.graph)
    if not graph_mode:
        capi.profile_begin()
    t_steps = timed(lambda: step(z_dev), args.steps)
    if not graph_mode:
        prof = capi.profile_end()
    clocks = sampler.stop() if rank == 0 else {}
    total_ms = sum(t_steps)
    if graph_mode:
        # events cannot be recorded inside a replayed graph: take the per-kernel times from a
        # short eager pass of the same step (outside the timed region)
        if world > 1:
            engine.use_cuda_graph = False
        else:
            graphed, keep = None, graphed
        for _ in range(2):
            step(z_dev)
        sync_all()
        capi.profile_begin()
        for _ in range(5):
            step(z_dev)
        sync_all()
        prof = capi.profile_end()
        if world > 1:
            engine.use_cuda_graph = True
        else:
            graphed = keep
        sync_all()  # eager and replayed steps pick their buffer set independently: never back to back

    # --- end to end: pinned host z -> device -> fused op -> loss back on the host -------
    z_stage = torch.empty_like(z_dev)
    loss_host = torch.empty(1, dtype=torch.float32).pin_memory()

    def e2e_step():
        z_stage.copy_(z_host, non_blocking=True)
        loss = step(z_stage)
        loss_host.copy_(loss, non_blocking=True)
    for _ in range(3):
        e2e_step()
    t_e2e = timed(e2e_step, args.steps)
    e2e_ms = sum(t_e2e)

    if world > 1:
        t = torch.tensor([total_ms, e2e_ms, prof["fwd_ms"], prof["bwd_ms"]], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_ms, prof["fwd_ms"], prof["bwd_ms"] = [float(x) for x in t.tolist()]

    loss_val = float(loss_host.item())
    if world > 1:
        engine.release()  # a live CUDA graph with NCCL kernels would hang the communicator teardown
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return 0

    flop_alg = 8.0 * M * M * D
    ms_per_step = total_ms / args.steps
    value = flop_alg / (ms_per_step * 1e-3) / 1e12
    e2e_value = flop_alg / (e2e_ms / args.steps * 1e-3) / 1e12
    peaks, peak_kind = measured_peaks()
    peak = peaks["bf16_tflops"]  # burst figure: the kernel is timed alone by its own events
    if dt != "bf16" and not args.simt:
        peak = peak / 2  # tf32 dense peak is half the bf16 one
    plan = capi.plan_info(M, D, rows, capi.BF16 if dt == "bf16" else capi.F32, flags)
    # executed tensor FLOPs of the backward tile pass on this rank: S recompute + W.U
    bwd_factor = 4.0 if (world == 1 or args.bwd_mode == "symmetric") else 8.0
    prof_steps = 5 if graph_mode else args.steps
    bwd_flop_per_launch = bwd_factor * rows * M * D / max(1, prof["bwd_launches"] // prof_steps)
    bwd_ms = prof["bwd_ms"] / max(1, prof["bwd_launches"])
    fwd_ms = prof["fwd_ms"] / max(1, prof["fwd_launches"])
    achieved = bwd_flop_per_launch / (bwd_ms * 1e-3) / 1e12 if bwd_ms > 0 else 0.0
    roofline = {
        "bound": "tensor", "kernel": "backward tile pass (recompute S, W = P + P^T, W.U)",
        "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
        "peak_source": f"{peak_kind} MEASURED_PEAKS.json bf16_tflops (burst)"
                       + ("" if dt == "bf16" or args.simt else " / 2 for tf32"),
        # DRAM bytes of one launch from the committed ncu capture (same workload, 1 GPU); the
        # kernel is tensor-bound: its operands (8 MB of U) live in L2 for the whole launch
        "traffic": ncu_dram_traffic("v3_bwd") if (world == 1 and (M, D, dt) == (32768, 128, "bf16")
                                                   and not args.simt) else None,
        "flop_per_launch": bwd_flop_per_launch, "ms_per_launch": bwd_ms,
        "fwd_kernel": {"ms_per_launch": fwd_ms, "flop_per_launch": 2.0 * rows * M * D,
                       "achieved": (2.0 * rows * M * D / (fwd_ms * 1e-3) / 1e12) if fwd_ms > 0 else 0.0},
        "executed_flop_per_step": (2.0 + bwd_factor) * M * M * D,
        "executed_tflops": (2.0 + bwd_factor) * M * M * D / (ms_per_step * 1e-3) / 1e12,
        "kernel_family": "tcgen05" if plan["tcgen05"] else "simt-fp32",
    }
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if dt == "bf16" else "f32",
        "data": "synthetic",
        "config": {"workload": f"ntxent_fwd_bwd_M{M}_D{D}_tau{tau}_{dt}", "M": M, "D": D, "tau": tau,
                   "rows_per_gpu": rows, "flop_model": "8*M^2*D per step (BASELINE.md section 3)",
                   "l2_flush_between_steps": flush is not None, "plan": plan,
                   "backward_mode": "symmetric" if world == 1 else args.bwd_mode,
                   "cuda_graph": bool(graph_mode), "comm": (comm_used if comm_note is None else f"{comm_used} ({comm_note})") if world > 1 else None},
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": z_host.numel() * z_host.element_size() * world,
                "d2h_bytes_per_step": 4 * world, "ms_per_step": e2e_ms / args.steps},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": roofline, "clocks": clocks, "loss": loss_val,
    }
    if world == 1 and not args.no_cpu_baseline:
        v, r, secs, cores = cpu_rowblock_tflops(z_host_np, tau, args.cpu_seconds)
        line["cpu_baseline"] = {
            "value": v, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"rows [0,{r}) of the M={M} problem in {secs:.1f} s, contract row-block form "
                      f"(8*rows*M*D FLOP), fp32 OpenMP oracle (oracle/ntxent_oracle.c)"}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


def main():
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())
