"""World-size-2 gloo tests of the row-partitioned NT-Xent orchestration (distributed.py) on CPU.

What runs here is the HOST logic of the multi-GPU path: shard bookkeeping, the all-gather of U,
the all-gather of l / all-reduce of the loss, the reduce-scatter of the column term, the autograd
wrapper.  The compute steps are a numpy test double defined in this file (the product has no CPU
implementation; its only backend is the CUDA one).  Results are checked against the fp64 oracle.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, seeded_z


class NumpyBackend:
    """Test double with the same four methods as distributed.CudaBackend, in fp64 numpy."""
    launches = None

    def operand_dtype(self, dtype):
        return torch.float64

    def normalize(self, z_rows):
        z = z_rows.double().numpy()
        n = np.maximum(np.linalg.norm(z, axis=1, keepdims=True), 1e-12)
        return torch.from_numpy(z / n), torch.from_numpy((1.0 / n)[:, 0].astype(np.float32))

    @staticmethod
    def _s(u_all, row_begin, rows, tau):
        u = u_all.numpy()
        S = u[row_begin:row_begin + rows] @ u.T / tau
        mask = np.zeros_like(S, dtype=bool)
        mask[np.arange(rows), row_begin + np.arange(rows)] = True
        return u, S, mask

    def forward_rows(self, u_all, row_begin, rows, tau, ws=None):
        u, S, mask = self._s(u_all, row_begin, rows, tau)
        M = u.shape[0]
        Sm = np.where(mask, -np.inf, S)
        mx = Sm.max(axis=1, keepdims=True)
        lse = (mx + np.log(np.exp(Sm - mx).sum(axis=1, keepdims=True)))[:, 0]
        pos = (row_begin + np.arange(rows) + M // 2) % M
        part = float((lse - S[np.arange(rows), pos]).sum())
        return torch.from_numpy(lse.astype(np.float32)), torch.tensor([part], dtype=torch.float32)

    def backward_rows(self, u_all, lse_all, row_begin, rows, tau, mode, ws=None):
        u, S, mask = self._s(u_all, row_begin, rows, tau)
        M = u.shape[0]
        l = lse_all.double().numpy()
        idx = np.arange(rows)
        pos = (row_begin + idx + M // 2) % M
        P = np.where(mask, 0.0, np.exp(S - l[row_begin:row_begin + rows, None]))
        sc = 1.0 / (tau * M)
        if mode == 0:  # symmetric
            W = P + np.where(mask, 0.0, np.exp(S - l[None, :]))
            du = (W @ u - 2.0 * u[pos]) * sc
            return torch.from_numpy(du.astype(np.float32)), None
        G = P.copy()
        G[idx, pos] -= 1.0
        row_term = (G @ u) * sc
        col_term = (G.T @ u[row_begin:row_begin + rows]) * sc
        return torch.from_numpy(row_term.astype(np.float32)), torch.from_numpy(col_term.astype(np.float32))

    def finalize_dz(self, du_rows, z_rows, inv_norm, grad_out):
        z = z_rows.double().numpy()
        inv = inv_norm.double().numpy()[:, None]
        u = z * inv
        du = du_rows.double().numpy()
        g = 1.0 if grad_out is None else float(grad_out.reshape(-1)[0])
        dz = g * (du - (du * u).sum(1, keepdims=True) * u) * inv
        return torch.from_numpy(dz).to(z_rows.dtype)


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, M, D, tau, mode, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import ntxent_b200  # noqa: F401
        from ntxent_b200.distributed import RowPartitionedNTXent, distributed_ntxent_loss
        z = torch.from_numpy(seeded_z(M, D, seed=77, dtype=np.float64))
        rows = M // world
        shard = z[rank * rows:(rank + 1) * rows].clone()
        eng = RowPartitionedNTXent(M, D, tau, torch.float64, "cpu", backward_mode=mode, backend=NumpyBackend())
        assert (eng.rows, eng.row_begin) == (rows, rank * rows)
        loss, dz = eng.forward_backward(shard, torch.tensor([0.5]))
        # the autograd wrapper on the same engine
        zr = shard.clone().requires_grad_(True)
        l2 = distributed_ntxent_loss(zr, tau, engine=eng)
        (l2 * 0.5).backward()
        np.savez(os.path.join(out_dir, f"rank{rank}.npz"), loss=loss.numpy(), dz=dz.numpy(),
                 loss_ag=l2.detach().numpy(), dz_ag=zr.grad.numpy())
        with pytest.raises(ValueError):
            eng.forward(shard[:-1])
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("mode", ["symmetric", "contract"])
def test_row_partition_world2_gloo(tmp_path, oracle, mode):
    M, D, tau, world = 64, 16, 0.2, 2
    mp.spawn(_worker, args=(world, _free_port(), M, D, tau, mode, str(tmp_path)), nprocs=world, join=True)
    z = seeded_z(M, D, seed=77, dtype=np.float64)
    ref_loss, _, ref_dz = oracle.forward_backward(z, tau, grad_out=0.5)
    rows = M // world
    for r in range(world):
        got = np.load(tmp_path / f"rank{r}.npz")
        assert float(got["loss"][0]) == pytest.approx(ref_loss, rel=1e-6)      # same global loss everywhere
        assert float(got["loss_ag"]) == pytest.approx(ref_loss, rel=1e-6)
        sl = slice(r * rows, (r + 1) * rows)
        scale = np.abs(ref_dz).max()
        assert np.abs(got["dz"] - ref_dz[sl]).max() <= 1e-5 * scale
        assert np.abs(got["dz_ag"] - ref_dz[sl]).max() <= 1e-5 * scale


def test_engine_argument_checks():
    import ntxent_b200  # noqa: F401
    from ntxent_b200.distributed import RowPartitionedNTXent
    with pytest.raises(ValueError):
        RowPartitionedNTXent(64, 16, 0.2, torch.float32, "cpu", backward_mode="ring", backend=NumpyBackend())
    with pytest.raises(ValueError):
        RowPartitionedNTXent(63, 16, 0.2, torch.float32, "cpu", backend=NumpyBackend())
