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
    ap.add_argument("--comm", default="push", choices=["push", "nccl", "torch"],
                    help="multi-GPU exchange: 'push' = the C-ABI communicator, kernels store into the peers' "
                         "buffers (flag-ordered, no collective); 'nccl' = the same communicator issuing NCCL "
                         "collectives (all-gather U + all-gather l, or with --bwd-mode contract the "
                         "reduce-scatter overlapped with the row pass); 'torch' = torch.distributed "
                         "collectives around the C-ABI building blocks")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-flush", action="store_true")
    ap.add_argument("--no-graph", action="store_true",
                    help="issue every launch from the host instead of replaying the step as a CUDA graph "
                         "(the default at every GPU count, so that the scaling compares like with like)")
    ap.add_argument("--no-sustained", action="store_true")
    ap.add_argument("--sustained-seconds", type=float, default=2.5)
    ap.add_argument("--cpu-whole-seconds", type=float, default=40.0,
                    help="run the whole problem on the CPU (full dz parity) when that takes no longer")
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
    path = os.path.join(ROOT, "profiles", f"r2_ncu_{kernel_tag}_summary.csv")
    if not os.path.exists(path):
        path = os.path.join(ROOT, "profiles", f"r1_ncu_v3_{kernel_tag}_summary.csv")
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
def cpu_rowblock(z, tau, seconds, rows_hint=0, whole_if_under=0.0):
    """Times the oracle's fp32 row-block form (oracle/ntxent_oracle.c) on all host cores, on rows
    [0, rows) of the problem: the whole problem when that is estimated to take no longer than
    `whole_if_under` seconds, else a sample sized to ~`seconds`.
    -> dict(tflops, rows, seconds, cores, out=(lse, loss_partial, du_row, du_col))."""
    from oracle import oracle as O
    O.build()
    cores = O.use_all_cores()  # torchrun exports OMP_NUM_THREADS=1 to its ranks
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
        whole = dt * M / probe
        if whole <= max(whole_if_under, seconds):
            rows = M
        else:
            rows = int(min(M, max(16, probe * seconds / dt)))
            rows = max(16, rows // 16 * 16)
    t0 = time.perf_counter()
    out = O.rowblock_f32(z, tau, 0, rows)
    dt = time.perf_counter() - t0
    return dict(tflops=8.0 * rows * M * D / dt / 1e12, rows=rows, seconds=dt, cores=cores, out=out)


def sample_text(rows, M, secs=None):
    what = "the whole step (all rows)" if rows == M else f"rows [0,{rows}) of the M={M} problem per step"
    took = f" in {secs:.1f} s" if secs is not None else ""
    return (f"{what}{took}, contract row-block form (fwd + recompute + G.U + G^T.U_I = 8*rows*M*D FLOP), "
            f"fp32 OpenMP oracle (oracle/ntxent_oracle.c)")


def parity_vs_oracle(z_np, tau, cpu, loss_gpu, lse_gpu, dz_gpu):
    """What the CPU baseline computed is not thrown away: compare it with the GPU's result of the
    same input.  With the whole problem on the CPU, all of dz; with a row sample, l on the sample
    and dz on its first 256 rows (their column term formed in numpy from the GPU's l)."""
    import numpy as np
    M, D = z_np.shape
    rows = cpu["rows"]
    lse_c, loss_part, du_row, du_col = cpu["out"]
    zq = z_np.astype(np.float64)
    nrm = np.maximum(np.linalg.norm(zq, axis=1, keepdims=True), 1e-12)
    u = zq / nrm
    par = {"rows": rows, "lse_maxabs": float(np.abs(lse_gpu[:rows] - lse_c).max())}
    if rows == M:
        par["loss_rel"] = abs(loss_gpu - loss_part / M) / abs(loss_part / M)
        R = M
        du = (du_row.astype(np.float64) + du_col.astype(np.float64)) / (tau * M)
    else:
        R = min(rows, 256)
        S = (u[:R] @ u.T) / tau
        P = np.exp(S - lse_gpu[None, :].astype(np.float64))
        P[np.arange(R), np.arange(R)] = 0.0
        P[np.arange(R), (np.arange(R) + M // 2) % M] -= 1.0
        du = (du_row[:R].astype(np.float64) + P @ u) / (tau * M)
    ref = (du - (du * u[:R]).sum(1, keepdims=True) * u[:R]) / nrm[:R]
    par["dz_rows"] = R
    par["dz_maxrel"] = float(np.abs(dz_gpu[:R] - ref).max() / np.abs(ref).max())
    par["norm"] = "max|dz - ref| / max|ref|; reference = CPU oracle (fp32 row-block form)"
    return par


def base_config(M, D, tau, dt):
    return {"workload": f"ntxent_fwd_bwd_M{M}_D{D}_tau{tau}_{dt}", "M": M, "D": D, "tau": tau}


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    M, D, tau, dt = workload(args)
    z = make_input(M, D)
    if dt == "bf16":  # the values the GPU arm sees
        import torch
        z = torch.from_numpy(z).to(torch.bfloat16).float().numpy()
    n = max(1, args.steps + args.warmup)
    per_step = max(1.0, min(10.0, 150.0 / n))
    first = cpu_rowblock(z, tau, per_step, 0, whole_if_under=per_step)
    rows, cores = first["rows"], first["cores"]
    for _ in range(max(0, args.warmup - 1)):
        cpu_rowblock(z, tau, per_step, rows)
    times = [cpu_rowblock(z, tau, per_step, rows)["seconds"] for _ in range(args.steps)]
    total = sum(times)
    val = 8.0 * rows * M * D * len(times) / total / 1e12
    cfg = base_config(M, D, tau, dt)
    cfg["sample"] = sample_text(rows, M)
    line = {
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * total / len(times),
        "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": sample_text(rows, M)},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "the reference has no CPU NT-Xent and its CUDA sources do not compile; this is "
                "the oracle port on the host cores (a rate: FLOP of the rows it processed / time)",
    }
    OUT.emit(json.dumps(line))
    return 0


# ------------------------------------------------------------------------ native
def run_native(args):
    import numpy as np
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
    graph_mode = not args.no_graph

    engine = None
    if world == 1:
        ws = capi.workspace(M, D, z_dev, flags)
        out = (torch.empty(1, dtype=torch.float32, device=dev),
               torch.empty(M, dtype=torch.float32, device=dev), torch.empty_like(z_dev))
        graphed = ntxent_b200.GraphedNTXent(M, D, tau, tdtype, dev, flags=flags) if graph_mode else None
        use_graph = [graph_mode]

        def step(zin):
            """-> (loss[1], dz[rows, D]) on the device"""
            if use_graph[0]:
                return graphed(zin)
            loss, _, dz = capi.forward_backward(zin, tau, None, False, flags, ws, out)
            return loss, dz
        # normalise, forward tiles, finalize (+ loss, by its last block), backward tiles, epilogue; fp32
        # operands on the tf32 tiles add the transpose of U
        import math
        sym_fwd = (dt == "bf16" and M >= 4096 and math.log2(math.e) / tau <= 50.0 and not args.simt
                   and os.environ.get("NTXENT_NO_SYM_FORWARD") != "1")
        launches_per_step = 5 + (1 if dt != "bf16" and not args.simt else 0)
        comm_used = comm_note = None
    else:
        from ntxent_b200.distributed import RowPartitionedNTXent
        comm_used, comm_note = args.comm, None
        if comm_used == "push" and args.bwd_mode == "contract":
            comm_used = "nccl"  # the contract form is the NCCL exchange's
        try:
            engine = RowPartitionedNTXent(M, D, tau, tdtype, dev, backward_mode=args.bwd_mode, flags=flags,
                                          use_cuda_graph=graph_mode, comm=comm_used)
        except Exception as exc:  # no peer mappings on this box: every rank fails the same way
            if comm_used == "torch":
                raise
            comm_note = f"{comm_used} unavailable: {type(exc).__name__}: {str(exc)[:120]}"
            comm_used = "torch"
            engine = RowPartitionedNTXent(M, D, tau, tdtype, dev, backward_mode=args.bwd_mode, flags=flags,
                                          use_cuda_graph=graph_mode, comm="torch")

        def step(zin):
            return engine.forward_backward(zin)
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

    def set_graph(on):
        """Eager and replayed steps pick their buffer set independently: never back to back."""
        sync_all()
        if world > 1:
            engine.use_cuda_graph = on
        else:
            use_graph[0] = on
        sync_all()

    for _ in range(max(args.warmup, 3)):
        step(z_dev)
    sync_all()

    # --- device-resident timing, with a clock record ---------------------------------------
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    t_steps = timed(lambda: step(z_dev), args.steps)
    clocks = sampler.stop() if rank == 0 else {}
    total_ms = sum(t_steps)

    # --- per-kernel times: CUDA events around the two tile kernels need eager launches ------
    if graph_mode:
        set_graph(False)
        for _ in range(2):
            step(z_dev)
        sync_all()
    prof_steps = 5 if graph_mode else args.steps
    capi.profile_begin()
    timed(lambda: step(z_dev), prof_steps)
    prof = capi.profile_end()
    if graph_mode:
        set_graph(True)

    # --- sustained: the same step back to back for >= 2 s under the clock sampler -----------
    sustained = None
    if not args.no_sustained:
        n_sus = max(50, int(args.sustained_seconds * 1e3 / max(total_ms / args.steps, 1e-3)))
        if rank == 0:
            sampler = ClockSampler(local_rank)
            sampler.start()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        a.record()
        for _ in range(n_sus):
            step(z_dev)
        b.record()
        sync_all()
        sus_ms = a.elapsed_time(b)
        sus_clocks = sampler.stop() if rank == 0 else {}
        if world > 1:
            t = torch.tensor([sus_ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            sus_ms = float(t.item())
        sustained = {"ms_per_step": sus_ms / n_sus, "seconds": sus_ms / 1e3, "steps": n_sus,
                     "l2_flush_between_steps": False, "clocks": sus_clocks}

    # --- end to end: pinned host z -> device -> fused op -> loss AND dz back on the host ----
    # Two-deep pipeline: the H2D copy of step i+1 and the D2H copy of step i-1's dz run on their
    # own streams under the kernels of step i.  Timed as one region of K steps (first copy in to
    # last copy out), max over ranks.
    copy_in, copy_out = torch.cuda.Stream(device=dev), torch.cuda.Stream(device=dev)
    main = torch.cuda.current_stream()
    z_stage = [torch.empty_like(z_dev) for _ in range(2)]
    dz_stage = [torch.empty_like(z_dev) for _ in range(2)]
    dz_host = [torch.empty_like(z_host).pin_memory() for _ in range(2)]
    loss_host = torch.empty(2, dtype=torch.float32).pin_memory()

    def e2e_region(steps):
        ev_in = [torch.cuda.Event() for _ in range(steps)]
        ev_done = [torch.cuda.Event() for _ in range(steps)]
        ev_out = [torch.cuda.Event() for _ in range(steps)]
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sync_all()
        t0.record(main)
        copy_in.wait_event(t0)
        copy_out.wait_event(t0)
        for i in range(steps):
            b = i & 1
            with torch.cuda.stream(copy_in):
                if i >= 2:
                    copy_in.wait_event(ev_done[i - 2])  # z_stage[b] was read by step i-2
                z_stage[b].copy_(z_host, non_blocking=True)
                ev_in[i].record(copy_in)
            main.wait_event(ev_in[i])
            if i >= 2:
                main.wait_event(ev_out[i - 2])          # dz_stage[b] has left for the host
            loss, dz = step(z_stage[b])
            dz_stage[b].copy_(dz, non_blocking=True)    # out of the op's (static) output buffer
            loss_stage = loss.clone()
            ev_done[i].record(main)
            with torch.cuda.stream(copy_out):
                copy_out.wait_event(ev_done[i])
                dz_host[b].copy_(dz_stage[b], non_blocking=True)
                loss_host[b:b + 1].copy_(loss_stage, non_blocking=True)
                ev_out[i].record(copy_out)
        main.wait_event(ev_out[steps - 1])
        if steps > 1:
            main.wait_event(ev_out[steps - 2])
        t1.record(main)
        sync_all()
        return t0.elapsed_time(t1)

    e2e_region(3)
    e2e_ms = e2e_region(args.steps)

    if world > 1:
        t = torch.tensor([total_ms, e2e_ms, prof["fwd_ms"], prof["bwd_ms"]], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        total_ms, e2e_ms, prof["fwd_ms"], prof["bwd_ms"] = [float(x) for x in t.tolist()]

    loss_val = float(loss_host[(args.steps - 1) & 1].item())
    dz_last = dz_host[(args.steps - 1) & 1]

    # --- parity of what was just timed ------------------------------------------------------
    parity = None
    if world > 1:
        # every rank: its dz shard against a single-GPU run of the whole problem on its own device
        zf = z_host_full.to(dev)
        l1, _, dz1 = capi.forward_backward(zf, tau, None, False, flags)
        torch.cuda.synchronize()
        ref = dz1[rank * rows:(rank + 1) * rows].float()
        err = ((dz_last.to(dev).float() - ref).abs().max() / dz1.float().abs().max()).reshape(1)
        lerr = (torch.tensor([loss_val], device=dev) - l1).abs() / l1.abs()
        t = torch.cat([err, lerr.reshape(1)])
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        parity = {"dz_vs_1gpu_maxrel": float(t[0].item()), "loss_vs_1gpu_rel": float(t[1].item()),
                  "rows": M, "norm": "max over ranks of max|dz_shard - dz_1gpu| / max|dz_1gpu|",
                  "timed_out_on_rank": engine.timed_out_on()}
        del zf, dz1
    lse_last = None
    if world == 1:
        lse_last = out[1].float().cpu().numpy() if not graph_mode else graphed.out[1].float().cpu().numpy()
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
    peak_sus = peaks.get("bf16_tflops_sustained", peak)
    if dt != "bf16" and not args.simt:
        peak, peak_sus = peak / 2, peak_sus / 2  # tf32 dense peak is half the bf16 one
    plan = capi.plan_info(M, D, rows, capi.BF16 if dt == "bf16" else capi.F32, flags)
    # executed tensor FLOPs of the backward tile pass on this rank: S recompute + W.U
    bwd_factor = 4.0 if (world == 1 or args.bwd_mode == "symmetric") else 8.0
    fwd_factor = 2.0
    bwd_flop_per_launch = bwd_factor * rows * M * D / max(1, prof["bwd_launches"] // prof_steps)
    bwd_ms = prof["bwd_ms"] / max(1, prof["bwd_launches"])
    fwd_ms = prof["fwd_ms"] / max(1, prof["fwd_launches"])
    achieved = bwd_flop_per_launch / (bwd_ms * 1e-3) / 1e12 if bwd_ms > 0 else 0.0
    executed = (fwd_factor + bwd_factor) * M * M * D
    roofline = {
        "bound": "tensor", "kernel": "backward tile pass (recompute S, W = P + P^T, W.U)",
        "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
        "peak_source": f"{peak_kind} MEASURED_PEAKS.json bf16_tflops (burst)"
                       + ("" if dt == "bf16" or args.simt else " / 2 for tf32"),
        # DRAM bytes of one launch from the committed ncu capture (same workload, 1 GPU); the
        # kernel is tensor-bound: its operands (8 MB of U) live in L2 for the whole launch
        "traffic": ncu_dram_traffic("bwd") if (world == 1 and (M, D, dt) == (32768, 128, "bf16")
                                               and not args.simt) else None,
        "flop_per_launch": bwd_flop_per_launch, "ms_per_launch": bwd_ms,
        "fwd_kernel": {"ms_per_launch": fwd_ms, "flop_per_launch": fwd_factor * rows * M * D,
                       "achieved": (fwd_factor * rows * M * D / (fwd_ms * 1e-3) / 1e12) if fwd_ms > 0 else 0.0},
        "side_kernels_ms": ms_per_step - fwd_ms - bwd_ms,
        "executed_flop_per_step": executed,
        "executed_tflops": executed / (ms_per_step * 1e-3) / 1e12,
        "kernel_family": "tcgen05" if plan["tcgen05"] else "simt-fp32",
    }
    if sustained is not None:
        sustained["executed_tflops"] = executed / (sustained["ms_per_step"] * 1e-3) / 1e12
        sustained["frac_of_sustained_peak"] = sustained["executed_tflops"] / peak_sus
        sustained["sustained_peak"] = peak_sus
    cfg = base_config(M, D, tau, dt)
    cfg.update({"rows_per_gpu": rows, "flop_model": "8*M^2*D per step (BASELINE.md section 3)",
                "l2_flush_between_steps": flush is not None, "plan": plan,
                "backward_mode": "symmetric" if world == 1 else args.bwd_mode,
                "cuda_graph": bool(graph_mode), "pdl": os.environ.get("NTXENT_NO_PDL") != "1",
                "comm": (comm_used if comm_note is None else f"{comm_used} ({comm_note})") if world > 1 else None})
    zbytes = z_host.numel() * z_host.element_size() * world
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "bf16" if dt == "bf16" else "f32",
        "data": "synthetic", "config": cfg,
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": zbytes,
                "d2h_bytes_per_step": zbytes + 4 * world, "ms_per_step": e2e_ms / args.steps,
                "what": "pinned z -> H2D -> fused fwd+bwd -> loss and dz D2H to pinned memory, every step; "
                        "two-deep pipeline (copies of neighbouring steps overlap the kernels), one timed "
                        "region of K steps; inputs stream from the host, no L2 flush"},
        "gpu_launches": launches_per_step * args.steps,
        "roofline": roofline, "sustained": sustained, "clocks": clocks, "loss": loss_val,
    }
    if parity is not None:
        line["parity"] = parity
    if world == 1 and not args.no_cpu_baseline:
        zq = z_host_full.float().numpy()  # the values the device saw
        cpu = cpu_rowblock(zq, tau, args.cpu_seconds, whole_if_under=args.cpu_whole_seconds)
        line["cpu_baseline"] = {"value": cpu["tflops"], "unit": UNIT, "cores": cpu["cores"], "kind": "port",
                                "sample": sample_text(cpu["rows"], M, cpu["seconds"])}
        line["parity"] = parity_vs_oracle(zq, tau, cpu, loss_val, lse_last.astype(np.float64),
                                          dz_last.float().numpy().astype(np.float64))
    OUT.emit(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()
    return 0


class CleanStdout:
    """The driver reads ONE JSON line from stdout.  Libraries print there too (NCCL announces its
    version on the first communicator): send everything to stderr while the benchmark runs and
    write the line to the real stdout at the end."""

    def __init__(self):
        sys.stdout.flush()
        self.real = os.dup(1)
        os.dup2(2, 1)

    def emit(self, text):
        sys.stdout.flush()
        os.write(self.real, (text.rstrip("\n") + "\n").encode())


OUT = None


def main():
    global OUT
    OUT = CleanStdout()
    args = parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_native(args)


if __name__ == "__main__":
    sys.exit(main())
