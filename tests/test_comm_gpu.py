"""GPU tests of the C-ABI communicator (include/ntxent_b200.h: ntxent_b200_comm_* and
ntxent_b200_dist_forward_backward), called through ctypes.

  * one rank on one GPU (runs on the driver's single-GPU box): every exchange / backward form must
    reproduce the single-device entry point;
  * all GPUs of the box, single process (comm_init_all, one host thread per device);
  * all GPUs of the box, one process per rank (comm_unique_id + comm_init_rank, the id handed
    over through a file): every rank's shard against the single-GPU result and the fp64 oracle;
  * a stress run of the flag-ordered push path: 200 steps as CUDA-graph replays with a different
    rank held back by a spin kernel every step, every step's result bitwise equal to the first.
The multi-GPU ones skip on a box with one GPU."""
import os
import sys
import threading
import time

import numpy as np
import pytest
import torch
import torch.multiprocessing as mp

from conftest import ROOT, seeded_z

import ntxent_b200  # noqa: F401
from ntxent_b200 import capi

pytestmark = pytest.mark.gpu

FORMS = [("push", capi.EXCHANGE_PUSH, capi.BWD_SYMMETRIC), ("nccl-sym", capi.EXCHANGE_NCCL, capi.BWD_SYMMETRIC),
         ("nccl-contract", capi.EXCHANGE_NCCL, capi.BWD_CONTRACT)]


def _rel(a, b):
    return ((a.float() - b.float()).abs().max() / b.float().abs().max()).item()


@pytest.mark.parametrize("dtype,M,D,tau", [(torch.bfloat16, 2048, 128, 0.1), (torch.float32, 1024, 64, 0.2)])
def test_single_rank_communicator_matches_single_device(dtype, M, D, tau):
    z = torch.from_numpy(seeded_z(M, D, seed=11)).cuda().to(dtype)
    g = torch.tensor([0.5], device="cuda")
    # the whole-batch forward of this size runs on the same (full-matrix) tile kernel as a row block
    loss1, _, dz1 = capi.forward_backward(z, tau, g)
    (comm,) = capi.Communicator.init_all([torch.cuda.current_device()], M, D, dtype)
    try:
        assert (comm.rank, comm.nranks, comm.rows) == (0, 1, M)
        tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
        for name, ex, mode in FORMS:
            for _ in range(3):  # both buffer sets, and the first one again
                loss, dz = comm.forward_backward(z, tau, g, ex, mode)
            torch.cuda.synchronize()
            assert loss.item() == pytest.approx(loss1.item(), rel=1e-6), name
            assert _rel(dz, dz1) <= tol, name
        assert comm.status() == 0
        with pytest.raises(capi.NTXentError):  # the push exchange runs the symmetric form only
            comm.forward_backward(z, tau, g, capi.EXCHANGE_PUSH, capi.BWD_CONTRACT)
    finally:
        comm.destroy()


def test_communicator_argument_errors():
    import ctypes
    L = capi.lib()
    h = ctypes.c_void_p()
    devs = (ctypes.c_int * 2)(0, 0)
    assert L.ntxent_b200_comm_init_all(2, devs, 1024, 128, capi.BF16, 0, ctypes.byref(h)) == -1  # one rank per GPU
    one = (ctypes.c_int * 1)(0)
    assert L.ntxent_b200_comm_init_all(1, one, 1023, 128, capi.BF16, 0, ctypes.byref(h)) == -1
    assert L.ntxent_b200_comm_init_all(1, one, 1024, 100, capi.BF16, 0, ctypes.byref(h)) == -4
    assert L.ntxent_b200_comm_destroy(None) == -1


# ------------------------------------------------------------------ all GPUs, one process
def test_init_all_one_thread_per_device(oracle):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    M, D, tau, dtype = 4096, 128, 0.1, torch.bfloat16
    z = torch.from_numpy(seeded_z(M, D, seed=5)).to(dtype)
    ref_loss, _, ref_dz = oracle.forward_backward(z.double().numpy(), tau)
    comms = capi.Communicator.init_all(list(range(world)), M, D, dtype)
    rows = M // world
    results, errors = {}, []

    def work(r):
        try:
            torch.cuda.set_device(r)
            shard = z[r * rows:(r + 1) * rows].cuda(r)
            out = {}
            for name, ex, mode in FORMS:
                if ex == capi.EXCHANGE_NCCL and not comms[r].has_nccl:
                    continue
                for _ in range(3):
                    loss, dz = comms[r].forward_backward(shard, tau, None, ex, mode)
                torch.cuda.synchronize(r)
                out[name] = (loss.item(), dz.float().cpu().numpy())
            results[r] = out
        except Exception as exc:  # noqa: BLE001
            errors.append((r, repr(exc)))

    threads = [threading.Thread(target=work, args=(r,)) for r in range(world)]
    for t in threads:
        t.start()
    for t in threads:
        t.join(timeout=300)
    try:
        assert not errors, errors
        scale = np.abs(ref_dz).max()
        for r in range(world):
            assert "push" in results[r]
            for name, (loss, dz) in results[r].items():
                assert loss == pytest.approx(ref_loss, rel=2e-3), (r, name)
                assert np.abs(dz - ref_dz[r * rows:(r + 1) * rows]).max() <= 1e-2 * scale, (r, name)
            assert comms[r].status() == 0
    finally:
        for c in comms:
            c.destroy()


# ------------------------------------------------------------ all GPUs, one process per rank
def _rank_worker(rank, world, id_path, M, D, tau, dtype_name, out_dir, stress_steps):
    sys.path.insert(0, ROOT)
    torch.cuda.set_device(rank)
    import ntxent_b200  # noqa: F401,F811
    from ntxent_b200 import capi as C
    dtype = getattr(torch, dtype_name)
    if rank == 0:
        uid = C.comm_unique_id()
        with open(id_path + ".tmp", "wb") as f:
            f.write(uid)
        os.replace(id_path + ".tmp", id_path)
    else:
        t0 = time.time()
        while not os.path.exists(id_path):
            if time.time() - t0 > 120:
                raise RuntimeError("no unique id from rank 0")
            time.sleep(0.05)
        uid = open(id_path, "rb").read()
    comm = C.Communicator.init_rank(uid, world, rank, M, D, dtype, torch.device("cuda", rank))
    try:
        rows = M // world
        z = torch.from_numpy(seeded_z(M, D, seed=5)).to(dtype)
        shard = z[rank * rows:(rank + 1) * rows].cuda()
        out = {}
        for name, ex, mode in FORMS:
            for _ in range(3):
                loss, dz = comm.forward_backward(shard, tau, None, ex, mode)
            torch.cuda.synchronize()
            out["loss_" + name] = loss.float().cpu().numpy()
            out["dz_" + name] = dz.float().cpu().numpy()
        if stress_steps:
            # push path under skew: two consecutive steps (both buffer sets) captured as one CUDA
            # graph, replayed; before every replay a different rank is held back by a spin kernel
            loss_s = torch.empty(1, device="cuda")
            dz_s = torch.empty_like(shard)
            first = comm.forward_backward(shard, tau, None, C.EXCHANGE_PUSH, C.BWD_SYMMETRIC)
            first = (first[0].clone(), first[1].clone())
            comm.forward_backward(shard, tau, None, C.EXCHANGE_PUSH, C.BWD_SYMMETRIC)  # back on set 0
            torch.cuda.synchronize()
            side = torch.cuda.Stream()
            with torch.cuda.stream(side):
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, stream=side):
                    for _ in range(2):
                        comm.forward_backward(shard, tau, None, C.EXCHANGE_PUSH, C.BWD_SYMMETRIC, out=(loss_s, dz_s))
            torch.cuda.synchronize()
            same = True
            for it in range(stress_steps // 2):
                if it % world == rank:
                    torch.cuda._sleep(int(2e6))  # ~1 ms behind the others
                graph.replay()
                if it % 10 == 0 or it == stress_steps // 2 - 1:
                    torch.cuda.synchronize()
                    same = same and torch.equal(dz_s, first[1]) and torch.equal(loss_s, first[0])
            torch.cuda.synchronize()
            out["stress_same"] = np.array([int(same)])
            del graph
        out["status"] = np.array([comm.status()])
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), **out)
    finally:
        comm.destroy()


@pytest.mark.parametrize("dtype_name,M,D,tau", [("bfloat16", 4096, 128, 0.1), ("float32", 2048, 128, 0.5)])
def test_init_rank_processes_and_push_stress(tmp_path, oracle, dtype_name, M, D, tau):
    world = min(torch.cuda.device_count(), 8)
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    stress = 200 if dtype_name == "bfloat16" else 0
    mp.spawn(_rank_worker, args=(world, str(tmp_path / "uid"), M, D, tau, dtype_name, str(tmp_path), stress),
             nprocs=world, join=True)
    dtype = getattr(torch, dtype_name)
    z = torch.from_numpy(seeded_z(M, D, seed=5)).to(dtype)
    ref_loss, _, ref_dz = oracle.forward_backward(z.double().numpy(), tau)
    rows = M // world
    tol_l, tol_g = (2e-3, 1e-2) if dtype_name == "bfloat16" else (1e-5, 1e-4)
    scale = np.abs(ref_dz).max()
    for r in range(world):
        got = np.load(tmp_path / f"rank{r}.npz")
        assert int(got["status"][0]) == 0
        for name, _, _ in FORMS:
            assert float(got["loss_" + name][0]) == pytest.approx(ref_loss, rel=tol_l), (r, name)
            assert np.abs(got["dz_" + name] - ref_dz[r * rows:(r + 1) * rows]).max() <= tol_g * scale, (r, name)
        if stress:
            assert int(got["stress_same"][0]) == 1, f"rank {r}: a replayed step differed from the first"
