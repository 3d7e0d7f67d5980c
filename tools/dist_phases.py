"""Per-phase GPU timing of the row-partitioned step (run under torchrun)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch, torch.distributed as dist
import ntxent_b200
from ntxent_b200 import capi
from ntxent_b200.distributed import RowPartitionedNTXent

rank = int(os.environ["RANK"]); world = int(os.environ["WORLD_SIZE"]); lr = int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
M, D, tau = 32768, 128, 0.1
rows = M // world
z = torch.from_numpy(np.random.default_rng(1234).standard_normal((M, D), dtype=np.float32)).to(torch.bfloat16)
zl = z[rank * rows:(rank + 1) * rows].to(dev)
eng = RowPartitionedNTXent(M, D, tau, torch.bfloat16, dev)
be = eng.backend

def ev():
    e = torch.cuda.Event(enable_timing=True); e.record(); return e

for _ in range(5):
    eng.forward_backward(zl)
torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
names = ["normalize", "allgather_U", "forward_rows", "allreduce+allgather_lse", "backward_rows", "finalize_dz"]
acc = np.zeros(len(names)); n = 20
t_cpu = 0.0
for _ in range(n):
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    e = [ev()]
    u, inv = be.normalize(zl); e.append(ev())
    dist.all_gather_into_tensor(eng.u_all, u); e.append(ev())
    lse, lp = be.forward_rows(eng.u_all, eng.row_begin, rows, tau, eng.ws); e.append(ev())
    loss = lp.clone(); dist.all_reduce(loss); dist.all_gather_into_tensor(eng.lse_all, lse); e.append(ev())
    du, _ = be.backward_rows(eng.u_all, eng.lse_all, eng.row_begin, rows, tau, 0, eng.ws); e.append(ev())
    dz = be.finalize_dz(du, zl, inv, None); e.append(ev())
    t_cpu += time.perf_counter() - t0
    torch.cuda.synchronize()
    acc += np.array([e[i].elapsed_time(e[i + 1]) for i in range(len(names))])
if rank == 0:
    print(f"[nccl] world={world} per-phase GPU ms:", {k: round(v / n, 4) for k, v in zip(names, acc)}, "sum", round(acc.sum() / n, 4),
          "cpu issue ms", round(1e3 * t_cpu / n, 4))

# the same step with the exchange fused into the kernels: peer stores ordered by barrier kernels ...
def run_symm(comm):
    eng = RowPartitionedNTXent(M, D, tau, torch.bfloat16, dev, comm=comm)
    sy = eng.symm
    for _ in range(5):
        eng.forward_backward(zl)
    flagged = comm == "symm"
    names = (["normalize_push+signal", "forward_rows_push(wait,signal)", "backward_rows_dz(wait)", "loss_sum"] if flagged
             else ["normalize_push", "barrier0", "forward_rows_push", "barrier1", "loss_sum", "backward_rows_dz"])
    acc = np.zeros(len(names)); t_cpu = 0.0
    for _ in range(n):
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        cur = sy.sets[sy.step & 1]; sy.step += 1
        t0 = time.perf_counter()
        e = [ev()]
        if flagged:
            inv = capi.normalize_push(zl, cur["u_peers"], eng.row_begin, 0, signal=cur["sig_u"]); e.append(ev())
            capi.forward_rows_push(cur["u_all"], eng.row_begin, rows, tau, cur["lse_peers"], cur["loss_peers"], rank, 0,
                                   eng.ws, u_ready=cur["wait_u"], lse_signal=cur["sig_lse"]); e.append(ev())
            dz = capi.backward_rows_dz(cur["u_all"], cur["lse_all"], eng.row_begin, rows, tau, zl, inv, None, 0, eng.ws,
                                       lse_ready=cur["wait_lse"]); e.append(ev())
            loss = cur["loss_parts"].sum(dim=0, keepdim=True); e.append(ev())
        else:
            inv = capi.normalize_push(zl, cur["u_peers"], eng.row_begin, 0); e.append(ev())
            sy.hdl.barrier(channel=0); e.append(ev())
            capi.forward_rows_push(cur["u_all"], eng.row_begin, rows, tau, cur["lse_peers"], cur["loss_peers"], rank, 0, eng.ws); e.append(ev())
            sy.hdl.barrier(channel=1); e.append(ev())
            loss = cur["loss_parts"].sum(dim=0, keepdim=True); e.append(ev())
            dz = capi.backward_rows_dz(cur["u_all"], cur["lse_all"], eng.row_begin, rows, tau, zl, inv, None, 0, eng.ws); e.append(ev())
        t_cpu += time.perf_counter() - t0
        torch.cuda.synchronize()
        acc += np.array([e[i].elapsed_time(e[i + 1]) for i in range(len(names))])
    if rank == 0:
        print(f"[{comm}] world={world} per-phase GPU ms:", {k: round(float(v) / n, 4) for k, v in zip(names, acc)}, "sum",
              round(acc.sum() / n, 4), "cpu issue ms", round(1e3 * t_cpu / n, 4))

run_symm("symm-barrier")
run_symm("symm")
dist.barrier(); torch.cuda.synchronize()
dist.destroy_process_group()
