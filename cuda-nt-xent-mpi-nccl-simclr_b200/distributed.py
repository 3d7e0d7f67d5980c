"""Row-partitioned (data-parallel) NT-Xent over torch.distributed -- SURVEY.md section 8e.

The reference names MPI/NCCL only as CMake link options (CMakeLists.txt:13-14,41-47,115-121) and
has no communicator, no collective and no multi-GPU test; this module is the new work that
BASELINE.json's north_star describes: one process per GPU, rank r owns rows
[r*M/R, (r+1)*M/R) of the stacked batch.

  forward   K0 normalise the shard -> all-gather U (operand dtype, 1 MiB per rank at 32768x128/8)
            -> forward tile pass over (own rows x all columns) -> l[own rows], partial loss
            -> all-gather l (M floats) + all-reduce of the partial loss (one float)
  backward  symmetric : dU_I = (G + G^T)_{I,:} U / tau from the gathered l; no second collective
            contract  : row term G_{I,:} U locally, column term G_{I,:}^T U_I for every global row
                        -> reduce-scatter (sum) -> add (the formulation the north_star words)
            then the normalisation Jacobian on the shard.

All compute goes through the C-ABI building blocks (capi.normalize / forward_rows /
backward_rows / finalize_dz).  The compute backend is an object so that the CPU tests can drive
the orchestration under gloo with a test double; the product default is the CUDA backend and
there is no CPU implementation behind it.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import capi


class CudaBackend:
    """The only product backend: hand-written sm_100a kernels behind the C-ABI."""

    launches = {"normalize": 1, "forward_rows": 3, "backward_rows_symmetric": 2,
                "backward_rows_contract": 4, "finalize_dz": 1}

    def __init__(self, flags: int = 0):
        self.flags = flags
        capi.lib()  # fail loudly now if the library is missing

    def normalize(self, z_rows):
        return capi.normalize(z_rows, self.flags)

    def forward_rows(self, u_all, row_begin, row_count, tau, ws=None):
        return capi.forward_rows(u_all, row_begin, row_count, tau, self.flags, ws)

    def backward_rows(self, u_all, lse_all, row_begin, row_count, tau, mode, ws=None):
        return capi.backward_rows(u_all, lse_all, row_begin, row_count, tau, mode, self.flags, ws)

    def finalize_dz(self, du_rows, z_rows, inv_norm, grad_out):
        return capi.finalize_dz(du_rows, z_rows, inv_norm, grad_out)

    def rows_workspace(self, M, D, rows, u_all):
        return capi.rows_workspace(M, D, rows, u_all, self.flags)


class _SymmetricBuffers:
    """Two rotating sets of (u_all, lse_all, loss_parts, flags) in symmetric memory: every rank maps
    every other rank's copy, so the kernels store their results into all of them directly
    (ntxent_b200_normalize_push / _forward_rows_push) and the NCCL all-gather / all-reduce
    disappear.  Producers and consumers are ordered by per-rank flags the kernels themselves
    publish and poll (include/ntxent_b200.h, ntxent_peer_signal / ntxent_peer_wait): no barrier
    kernel, and a consumer starts on the rows that have arrived.  Two sets, because a fast rank
    may already push step t+1 while a slow one still reads step t; by the time it pushes step
    t+2 into the first set again it has waited for every peer's step-t+1 push, which those peers
    issue after they are done with step t."""

    def __init__(self, M, D, op_dtype, world, rank, device, group):
        import torch.distributed._symmetric_memory as symm_mem
        item = torch.empty(0, dtype=op_dtype).element_size()
        al = lambda n: (n + 255) // 256 * 256  # noqa: E731
        u_bytes, lse_bytes, loss_bytes, sync_bytes = al(M * D * item), al(M * 4), 256, 256
        set_bytes = u_bytes + lse_bytes + loss_bytes + sync_bytes
        self.buf = symm_mem.empty(2 * set_bytes, dtype=torch.uint8, device=device)
        self.hdl = symm_mem.rendezvous(self.buf, group if group is not None else dist.group.WORLD)
        bases = [int(p) for p in self.hdl.buffer_ptrs]
        if len(bases) != world:
            raise capi.NTXentError("symmetric-memory rendezvous returned an unexpected number of peers")
        # flags and counters start at zero on every rank before anybody publishes anything
        self.buf.zero_()
        torch.cuda.synchronize(device)
        self.hdl.barrier(channel=0)
        torch.cuda.synchronize(device)
        local = bases[rank]
        rows = M // world
        self.sets = []
        for k in range(2):
            o = k * set_bytes
            o_lse, o_loss = o + u_bytes, o + u_bytes + lse_bytes
            o_sync = o_loss + loss_bytes  # flag_u[16] | flag_lse[16] | epoch_u . epoch_lse
            sig_u, sig_lse = capi.PeerSignal(), capi.PeerSignal()
            sig_u.flags = capi.PeerList.of([b + o_sync for b in bases])
            sig_u.epoch, sig_u.rank = local + o_sync + 128, rank
            sig_lse.flags = capi.PeerList.of([b + o_sync + 64 for b in bases])
            sig_lse.epoch, sig_lse.rank = local + o_sync + 136, rank
            # (word 36 of the sync block: set by a kernel whose wait for a peer timed out)
            wait_u = capi.PeerWait(local + o_sync, local + o_sync + 128, world, rows, local + o_sync + 144)
            wait_lse = capi.PeerWait(local + o_sync + 64, local + o_sync + 136, world, rows,
                                     local + o_sync + 144)
            self.sets.append(dict(
                u_all=self.buf[o:o + M * D * item].view(op_dtype).view(M, D),
                lse_all=self.buf[o_lse:o_lse + M * 4].view(torch.float32),
                loss_parts=self.buf[o_loss:o_loss + 4 * world].view(torch.float32),
                u_peers=capi.PeerList.of([b + o for b in bases]),
                lse_peers=capi.PeerList.of([b + o_lse for b in bases]),
                loss_peers=capi.PeerList.of([b + o_loss for b in bases]),
                sig_u=sig_u, sig_lse=sig_lse, wait_u=wait_u, wait_lse=wait_lse))
        self.step = 0
        self._sync_words = [self.buf[k * set_bytes + u_bytes + lse_bytes + loss_bytes:(k + 1) * set_bytes]
                            .view(torch.int32) for k in range(2)]

    def timed_out_on(self):
        """Rank a kernel of this process gave up waiting for (ntxent_b200_set_peer_wait_timeout_ms),
        or None.  Reads device memory: synchronises."""
        for w in self._sync_words:
            v = int(w[36].item())
            if v:
                return v - 1
        return None


def _reduce_scatter_sum(out: torch.Tensor, inp: torch.Tensor, group) -> None:
    """Sum-reduce-scatter; backends without the primitive (gloo) get all-reduce + slice."""
    try:
        dist.reduce_scatter_tensor(out, inp, op=dist.ReduceOp.SUM, group=group)
    except (RuntimeError, NotImplementedError):
        dist.all_reduce(inp, op=dist.ReduceOp.SUM, group=group)
        r = dist.get_rank(group)
        out.copy_(inp.view(dist.get_world_size(group), -1)[r].view_as(out))


class RowPartitionedNTXent:
    """Fused forward+backward of the global NT-Xent loss from each rank's row shard.

    `forward_backward(z_local)` returns (loss[1] -- the GLOBAL mean loss, identical on every
    rank --, dz_local -- d loss / d z_local).  Buffers for the collectives are allocated once.
    """

    def __init__(self, M: int, D: int, tau: float, dtype: torch.dtype, device, group=None,
                 backward_mode: str = "symmetric", flags: int = 0, backend=None,
                 use_mixed_precision: bool = False, use_cuda_graph: bool = False, comm: str = "nccl"):
        if backward_mode not in ("symmetric", "contract"):
            raise ValueError("backward_mode must be 'symmetric' or 'contract'")
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        if M % 2 or M % self.world:
            raise ValueError(f"global batch M={M} must divide evenly over {self.world} ranks")
        self.M, self.D, self.tau = int(M), int(D), float(tau)
        self.rows = self.M // self.world
        self.row_begin = self.rank * self.rows
        self.mode = backward_mode
        if use_mixed_precision:
            flags |= capi.FLAG_MIXED_PRECISION
        self.backend = backend if backend is not None else CudaBackend(flags)
        self.device = torch.device(device)
        op_dtype = torch.bfloat16 if (dtype == torch.bfloat16 or use_mixed_precision) else torch.float32
        if hasattr(self.backend, "operand_dtype"):  # test doubles may keep wider operands
            op_dtype = self.backend.operand_dtype(dtype)
        if comm not in ("nccl", "symm", "symm-barrier"):
            raise ValueError("comm must be 'nccl' (all-gather / all-reduce collectives), 'symm' (results "
                             "pushed into the peers' symmetric memory by the kernels, which also order "
                             "themselves through flags) or 'symm-barrier' (the same stores, ordered by "
                             "symmetric-memory barrier kernels)")
        self.comm = comm if self.world > 1 else "nccl"
        self.flags = flags
        self.symm = None
        if self.comm != "nccl":
            if backend is not None:
                raise ValueError("comm='symm' needs the CUDA backend")
            self.symm = _SymmetricBuffers(self.M, self.D, op_dtype, self.world, self.rank, self.device, group)
        self.u_all = torch.empty(self.M, self.D, dtype=op_dtype, device=self.device)
        self.lse_all = torch.empty(self.M, dtype=torch.float32, device=self.device)
        self.du_col_local = torch.empty(self.rows, self.D, dtype=torch.float32, device=self.device)
        self.ws = (self.backend.rows_workspace(self.M, self.D, self.rows, self.u_all)
                   if hasattr(self.backend, "rows_workspace") else None)
        sym = backward_mode == "symmetric"
        lc = getattr(self.backend, "launches", None)
        self.launches_per_step = (lc["normalize"] + lc["forward_rows"] + lc["finalize_dz"] +
                                  lc["backward_rows_symmetric" if sym else "backward_rows_contract"]
                                  ) if lc else 0
        # CUDA-graph replay of the whole step (kernels and collectives): at 4-8 ranks the step is
        # ~0.1-0.2 ms of GPU work behind ~0.3 ms of host-side launches, so the host has to go
        self.use_cuda_graph = bool(use_cuda_graph)
        self._graphs = {}
        self._gstep = 0
        if self.symm is not None and lc:
            signal = 1 if self.comm == "symm" else 0  # the one-block kernel that publishes U
            self.launches_per_step = (lc["normalize"] + signal + lc["forward_rows"] +
                                      (2 if sym else lc["backward_rows_contract"] + lc["finalize_dz"] + signal))

    # -- forward -------------------------------------------------------------------------
    def forward(self, z_local: torch.Tensor):
        """-> (loss[1], state); state feeds backward().  With more than one rank the returned loss
        is this rank's partial sum with its all-reduce still in flight: pass it through
        `_join_loss` (forward_backward and the autograd wrapper do) to get the global mean."""
        if tuple(z_local.shape) != (self.rows, self.D):
            raise ValueError(f"rank {self.rank} expects a [{self.rows}, {self.D}] shard, got {tuple(z_local.shape)}")
        if self.symm is not None:
            return self._forward_symm(z_local, defer_loss=False)
        u_local, inv_norm = self.backend.normalize(z_local)
        if self.world > 1:
            dist.all_gather_into_tensor(self.u_all, u_local.contiguous(), group=self.group)
        else:
            self.u_all.copy_(u_local)
        lse_local, loss_partial = self.backend.forward_rows(self.u_all, self.row_begin, self.rows,
                                                            self.tau, self.ws)
        loss = loss_partial.clone()
        if self.world > 1:
            if self.mode == "symmetric":
                dist.all_gather_into_tensor(self.lse_all, lse_local.contiguous(), group=self.group)
            else:  # contract: a rank only ever needs its own l
                self.lse_all[self.row_begin:self.row_begin + self.rows].copy_(lse_local)
            # nothing in the backward needs the loss value: its all-reduce is issued
            # asynchronously and joined by whoever reads the loss (`_LazyLoss`, or the end of
            # forward_backward), so the backward tiles never queue up behind it
            work = dist.all_reduce(loss, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self._pending_loss = work
        else:
            self.lse_all.copy_(lse_local)
        return loss, (z_local, inv_norm)

    def _forward_symm(self, z_local: torch.Tensor, defer_loss: bool):
        """Forward with the exchange inside the kernels: normalize_push publishes "my rows have
        landed", the forward tile kernel loads a column tile once its producers have published,
        the finalize/loss kernels publish l and the partial loss.  With defer_loss the loss parts
        are summed after the backward kernel (which waits for the same flags); otherwise a
        one-warp wait kernel orders the sum."""
        sy = self.symm
        k = sy.step & 1
        cur = sy.sets[k]
        sy.step += 1
        self.u_all, self.lse_all = cur["u_all"], cur["lse_all"]
        if self.comm == "symm-barrier":
            inv_norm = capi.normalize_push(z_local, cur["u_peers"], self.row_begin, self.flags)
            sy.hdl.barrier(channel=0)  # every rank's rows have landed in this rank's u_all
            capi.forward_rows_push(self.u_all, self.row_begin, self.rows, self.tau, cur["lse_peers"],
                                   cur["loss_peers"], self.rank, self.flags, self.ws)
            sy.hdl.barrier(channel=1)  # ... and every rank's l and partial loss
            return cur["loss_parts"].sum(dim=0, keepdim=True), (z_local, inv_norm, k)
        inv_norm = capi.normalize_push(z_local, cur["u_peers"], self.row_begin, self.flags,
                                       signal=cur["sig_u"])
        capi.forward_rows_push(self.u_all, self.row_begin, self.rows, self.tau, cur["lse_peers"],
                               cur["loss_peers"], self.rank, self.flags, self.ws,
                               u_ready=cur["wait_u"], lse_signal=cur["sig_lse"])
        loss = None
        if not defer_loss:
            capi.peer_wait(cur["wait_lse"], self.device)
            loss = cur["loss_parts"].sum(dim=0, keepdim=True)
        return loss, (z_local, inv_norm, k)

    def _join_loss(self, loss: torch.Tensor) -> torch.Tensor:
        work, self._pending_loss = getattr(self, "_pending_loss", None), None
        if work is not None:
            work.wait()  # makes the current stream wait for the collective; no host sync
        return loss / self.M

    # -- backward ------------------------------------------------------------------------
    def backward(self, state, grad_out: torch.Tensor | None = None) -> torch.Tensor:
        z_local, inv_norm = state[0], state[1]
        if self.symm is not None:
            cur = self.symm.sets[state[2]]
            self.u_all, self.lse_all = cur["u_all"], cur["lse_all"]
            if self.mode == "symmetric":  # tile kernel waits for the peers' l, epilogue emits dz
                return capi.backward_rows_dz(self.u_all, self.lse_all, self.row_begin, self.rows, self.tau,
                                             z_local, inv_norm, grad_out, self.flags, self.ws,
                                             lse_ready=cur["wait_lse"] if self.comm == "symm" else None)
        if self.mode == "symmetric":
            du_rows, _ = self.backend.backward_rows(self.u_all, self.lse_all, self.row_begin, self.rows,
                                                    self.tau, capi.BWD_SYMMETRIC, self.ws)
        else:
            du_rows, du_cols = self.backend.backward_rows(self.u_all, self.lse_all, self.row_begin,
                                                          self.rows, self.tau, capi.BWD_CONTRACT, self.ws)
            if self.world > 1:
                _reduce_scatter_sum(self.du_col_local, du_cols, self.group)
                du_rows = du_rows + self.du_col_local
            else:
                du_rows = du_rows + du_cols
        return self.backend.finalize_dz(du_rows, z_local, inv_norm, grad_out)

    def forward_backward(self, z_local: torch.Tensor, grad_out: torch.Tensor | None = None):
        if self.use_cuda_graph and z_local.is_cuda:
            return self._replay(z_local, grad_out)
        return self._step(z_local, grad_out)

    def _step(self, z_local, grad_out):
        if self.comm == "symm" and self.mode == "symmetric":
            _, state = self._forward_symm(z_local, defer_loss=True)
            dz = self.backward(state, grad_out)
            # the backward tile kernel has waited for every rank's l and partial loss
            loss = self.symm.sets[state[2]]["loss_parts"].sum(dim=0, keepdim=True)
            return loss / self.M, dz
        loss, state = self.forward(z_local)
        dz = self.backward(state, grad_out)
        return self._join_loss(loss), dz

    def release(self) -> None:
        """Drop the captured graphs and their static buffers.  Call this before
        torch.distributed.destroy_process_group(): tearing a NCCL communicator down while a
        live CUDA graph still holds its kernels hangs."""
        self._graphs = {}
        self._z_static = self._g_static = None
        if torch.cuda.is_available():
            torch.cuda.synchronize()

    def _replay(self, z_local, grad_out):
        """Capture forward+backward once per buffer set (static inputs), then copy-in + replay.
        The returned loss / dz are the graph's static outputs: consume them before the call
        after next.  Replayed and eager steps each alternate between the two buffer sets on
        their own count: when switching an engine between the two modes, put a cross-rank
        barrier in between (bench.py does) so that two consecutive steps never share a set."""
        k = (self._gstep & 1) if self.symm is not None else 0
        self._gstep += 1
        if not self._graphs:
            self._z_static = z_local.detach().clone()
            self._g_static = torch.ones(1, dtype=torch.float32, device=z_local.device)
            side = torch.cuda.Stream(device=z_local.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):  # warm-up on the side stream (communicators, lazy inits); both sets
                    self._step(self._z_static, self._g_static)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
        if k not in self._graphs:
            if self.symm is not None:
                self.symm.step = k  # the captured kernels are bound to buffer set k
            torch.cuda.synchronize()
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                loss, dz = self._step(self._z_static, self._g_static)
            self._graphs[k] = (graph, loss, dz)
        self._z_static.copy_(z_local, non_blocking=True)
        if grad_out is None:
            self._g_static.fill_(1.0)
        else:
            self._g_static.copy_(grad_out.reshape(-1)[:1], non_blocking=True)
        graph, loss, dz = self._graphs[k]
        graph.replay()
        return loss, dz


class _DistributedNTXentFunction(torch.autograd.Function):
    @staticmethod
    def forward(ctx, z_local, engine):
        loss, state = engine.forward(z_local.detach())
        ctx.engine, ctx.state = engine, state
        return engine._join_loss(loss)

    @staticmethod
    def backward(ctx, grad_loss):
        return ctx.engine.backward(ctx.state, grad_loss), None


def distributed_ntxent_loss(z_local: torch.Tensor, temperature: float = 0.07,
                            use_mixed_precision: bool = False, group=None,
                            backward_mode: str = "symmetric", engine: RowPartitionedNTXent | None = None):
    """Differentiable global NT-Xent loss (0-dim) from this rank's [M/R, D] shard.

    The value is the mean over all M global rows (the same on every rank); the gradient that
    flows to z_local is d(global loss)/d z_local.  Pass a prebuilt `engine` to reuse its buffers."""
    if engine is None:
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        engine = RowPartitionedNTXent(z_local.shape[0] * world, z_local.shape[1], temperature, z_local.dtype,
                                      z_local.device, group=group, backward_mode=backward_mode,
                                      use_mixed_precision=use_mixed_precision)
    return _DistributedNTXentFunction.apply(z_local, engine).squeeze(0)
