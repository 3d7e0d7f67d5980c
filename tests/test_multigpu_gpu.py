"""Real multi-GPU run of the row-partitioned path (NCCL), skipped on boxes with one GPU.

Each rank computes its shard through the CUDA backend; every rank's dz shard and the global loss
are compared with the single-GPU result and with the fp64 oracle (SURVEY.md section 8e,
"verification without a cluster", item 2)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, seeded_z

pytestmark = pytest.mark.gpu


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, M, D, tau, mode, dtype_name, out_dir):
    sys.path.insert(0, ROOT)
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank))
    try:
        import ntxent_b200  # noqa: F401
        from ntxent_b200.distributed import RowPartitionedNTXent, distributed_ntxent_loss
        dtype = getattr(torch, dtype_name)
        z = torch.from_numpy(seeded_z(M, D, seed=5)).to(dtype)
        rows = M // world
        shard = z[rank * rows:(rank + 1) * rows].cuda()
        dev = torch.device("cuda", rank)
        # the Python-level orchestration over torch.distributed collectives (the baseline) ...
        eng = RowPartitionedNTXent(M, D, tau, dtype, dev, backward_mode=mode, comm="torch")
        loss, dz = eng.forward_backward(shard)
        zr = shard.clone().requires_grad_(True)
        l2 = distributed_ntxent_loss(zr, tau, engine=eng)
        l2.backward()
        torch.cuda.synchronize()
        # ... whose CUDA-graph replay must reproduce the eager result bit for bit
        eng_g = RowPartitionedNTXent(M, D, tau, dtype, dev, backward_mode=mode, comm="torch", use_cuda_graph=True)
        for _ in range(2):
            loss_g, dz_g = eng_g.forward_backward(shard)
        torch.cuda.synchronize()
        same = torch.equal(loss_g, loss) and torch.equal(dz_g, dz)
        eng_g.release()
        assert same
        # the C-ABI communicator: kernels push into the peers' buffers (symmetric form), or NCCL
        # collectives with the reduce-scatter overlapped (contract form): same numbers as the baseline,
        # over several steps so that both rotating buffer sets are used
        tol = 1e-2 if dtype == torch.bfloat16 else 1e-5
        for comm in (["push", "nccl"] if mode == "symmetric" else ["nccl"]):
            eng_c = RowPartitionedNTXent(M, D, tau, dtype, dev, backward_mode=mode, comm=comm)
            for _ in range(3):
                loss_c, dz_c = eng_c.forward_backward(shard)
            torch.cuda.synchronize()
            loss_c, dz_c = loss_c.clone(), dz_c.clone()
            assert torch.allclose(loss_c, loss, rtol=1e-6), comm
            assert (dz_c.float() - dz.float()).abs().max() <= tol * dz.float().abs().max(), comm
            # through the autograd wrapper on the same engine
            zs = shard.clone().requires_grad_(True)
            ls = distributed_ntxent_loss(zs, tau, engine=eng_c)
            (ls * 2.0).backward()
            torch.cuda.synchronize()
            assert torch.allclose(ls.detach().reshape(1), loss.reshape(1), rtol=1e-6), comm
            assert torch.equal(zs.grad, (dz_c.float() * 2.0).to(dz_c.dtype)), comm
            # and its graph replay (one graph per buffer set), bit for bit over several steps,
            # switching between eager and replayed steps on one engine
            eng_c.use_cuda_graph = True
            same = True
            for _ in range(5):
                loss_g, dz_g = eng_c.forward_backward(shard)
                torch.cuda.synchronize()
                same = same and torch.equal(loss_g, loss_c) and torch.equal(dz_g, dz_c)
            eng_c.use_cuda_graph = False
            loss_e, dz_e = eng_c.forward_backward(shard)  # an odd number of replays, then eager again
            torch.cuda.synchronize()
            same = same and torch.equal(loss_e, loss_c) and torch.equal(dz_e, dz_c)
            assert eng_c.timed_out_on() is None
            eng_c.release()
            assert same, comm
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), loss=loss.float().cpu().numpy(),
                 dz=dz.float().cpu().numpy(), dz_ag=zr.grad.float().cpu().numpy())
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["symmetric", "contract"])
@pytest.mark.parametrize("dtype_name", ["bfloat16", "float32"])
def test_row_partition_nccl(tmp_path, oracle, mode, dtype_name):
    world = torch.cuda.device_count()
    if world < 2:
        pytest.skip("needs at least 2 GPUs")
    world = min(world, 8)
    M, D, tau = 2048, 128, 0.1
    mp.spawn(_worker, args=(world, _free_port(), M, D, tau, mode, dtype_name, str(tmp_path)), nprocs=world,
             join=True)
    dtype = getattr(torch, dtype_name)
    z = torch.from_numpy(seeded_z(M, D, seed=5)).to(dtype)
    ref_loss, _, ref_dz = oracle.forward_backward(z.double().numpy(), tau)
    rows = M // world
    tol_l, tol_g = (2e-3, 1e-2) if dtype_name == "bfloat16" else (1e-5, 1e-4)
    scale = np.abs(ref_dz).max()
    for r in range(world):
        got = np.load(tmp_path / f"rank{r}.npz")
        assert float(got["loss"][0]) == pytest.approx(ref_loss, rel=tol_l)
        sl = slice(r * rows, (r + 1) * rows)
        assert np.abs(got["dz"] - ref_dz[sl]).max() <= tol_g * scale
        assert np.abs(got["dz_ag"] - ref_dz[sl]).max() <= tol_g * scale
