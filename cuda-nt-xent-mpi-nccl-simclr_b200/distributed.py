"""Row-partitioned (data-parallel) NT-Xent -- SURVEY.md section 8e.

The reference names MPI/NCCL only as CMake link options (CMakeLists.txt:13-14,41-47,115-121) and
has no communicator, no collective and no multi-GPU test; this module is the Python face of the new
work BASELINE.json's north_star describes: one process per GPU, rank r owns rows
[r*M/R, (r+1)*M/R) of the stacked batch.

Three ways to run the step (`RowPartitionedNTXent(comm=...)`):

  "push"   the C-ABI communicator (csrc/comm.cu, ntxent_b200_comm_*), exchange = PUSH: the kernels
           store U, l and the partial loss straight into every rank's peer-mapped buffers and order
           themselves through flags; the whole step is one C call and no collective is launched.
           The default on CUDA.
  "nccl"   the same communicator, exchange = NCCL: ncclAllGather of U, then ncclAllGather of l
           (symmetric backward) or ncclReduceScatter of the column term on a second stream while the
           row term is computed (contract backward) -- north_star's formulation.
  "torch"  the orchestration in Python over torch.distributed collectives and the C-ABI building
           blocks (capi.normalize / forward_rows / backward_rows / finalize_dz), with the compute
           behind a backend object: what the CPU tests drive under gloo with a test double.  The
           product backend is the CUDA one and there is no CPU implementation behind it.

The communicator's 128-byte id travels from rank 0 to the others through torch.distributed
(broadcast_object_list); nothing else of torch.distributed is used by "push" / "nccl".
"""
from __future__ import annotations

import torch
import torch.distributed as dist

from . import capi


class CudaBackend:
    """The only product backend of the "torch" mode: hand-written sm_100a kernels behind the C-ABI."""

    launches = {"normalize": 1, "forward_rows": 4, "backward_rows_symmetric": 2,
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
    rank --, dz_local -- d loss / d z_local).  Buffers are allocated once, at construction.
    """

    def __init__(self, M: int, D: int, tau: float, dtype: torch.dtype, device, group=None,
                 backward_mode: str = "symmetric", flags: int = 0, backend=None,
                 use_mixed_precision: bool = False, use_cuda_graph: bool = False, comm: str | None = None):
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
        self.device = torch.device(device)
        if comm == "symm":  # the name this mode had while it sat on torch's symmetric memory
            comm = "push"
        if comm is None:
            comm = "torch" if (backend is not None or self.device.type != "cuda") else "push"
            if comm == "push" and backward_mode != "symmetric":
                comm = "nccl"
        if comm not in ("push", "nccl", "torch"):
            raise ValueError("comm must be 'push' (kernels store into the peers' buffers, ordered by flags), "
                             "'nccl' (NCCL collectives issued by the C-ABI communicator) or 'torch' "
                             "(torch.distributed collectives around the C-ABI building blocks)")
        if comm == "push" and backward_mode != "symmetric":
            raise ValueError("comm='push' runs the symmetric backward; the contract form is comm='nccl' or 'torch'")
        if comm != "torch" and backend is not None:
            raise ValueError("a custom backend goes with comm='torch'")
        self.comm = comm
        if use_mixed_precision:
            flags |= capi.FLAG_MIXED_PRECISION
        self.flags = flags
        self.dtype = dtype
        self.use_cuda_graph = bool(use_cuda_graph)
        self._graphs = []
        self._gstep = 0
        self._last_was_replay = None
        self.ccomm = None
        if comm != "torch":
            self._init_communicator()
            self.launches_per_step = 6 if comm == "push" else (6 if backward_mode == "symmetric" else 10)
            return
        self.backend = backend if backend is not None else CudaBackend(flags)
        op_dtype = torch.bfloat16 if (dtype == torch.bfloat16 or use_mixed_precision) else torch.float32
        if hasattr(self.backend, "operand_dtype"):  # test doubles may keep wider operands
            op_dtype = self.backend.operand_dtype(dtype)
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

    # -- C-ABI communicator ---------------------------------------------------------------
    def _init_communicator(self):
        if self.world == 1:
            (self.ccomm,) = capi.Communicator.init_all([self.device.index or 0], self.M, self.D, self.dtype,
                                                       self.flags)
            return
        box = [capi.comm_unique_id() if self.rank == 0 else None]
        src = dist.get_global_rank(self.group, 0) if self.group is not None else 0
        dist.broadcast_object_list(box, src=src, group=self.group)
        self.ccomm = capi.Communicator.init_rank(box[0], self.world, self.rank, self.M, self.D, self.dtype,
                                                 self.device, self.flags)
        if self.comm == "nccl" and not self.ccomm.has_nccl:
            raise capi.NTXentError("comm='nccl' needs libnccl.so.2")

    def _ccall(self, z_local, grad_out, out=None):
        ex = capi.EXCHANGE_PUSH if self.comm == "push" else capi.EXCHANGE_NCCL
        mode = capi.BWD_SYMMETRIC if self.mode == "symmetric" else capi.BWD_CONTRACT
        return self.ccomm.forward_backward(z_local, self.tau, grad_out, ex, mode, out)

    def timed_out_on(self):
        """Rank a kernel of this process gave up waiting for (ntxent_b200_set_peer_wait_timeout_ms),
        or None.  Reads device memory: synchronises."""
        if self.ccomm is None:
            return None
        s = self.ccomm.status()
        return s - 1 if s > 0 else None

    # -- "torch" mode: forward ------------------------------------------------------------
    def forward(self, z_local: torch.Tensor):
        """-> (loss[1], state); state feeds backward().  With more than one rank the returned loss
        is this rank's partial sum with its all-reduce still in flight: pass it through
        `_join_loss` (forward_backward and the autograd wrapper do) to get the global mean.
        ("torch" mode only: the communicator modes run the step as one fused call.)"""
        if self.comm != "torch":
            raise capi.NTXentError("forward()/backward() are the 'torch' mode's two halves; the communicator "
                                   "modes have forward_backward() only")
        if tuple(z_local.shape) != (self.rows, self.D):
            raise ValueError(f"rank {self.rank} expects a [{self.rows}, {self.D}] shard, got {tuple(z_local.shape)}")
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
            # asynchronously and joined by whoever reads the loss, so the backward tiles never
            # queue up behind it
            work = dist.all_reduce(loss, op=dist.ReduceOp.SUM, group=self.group, async_op=True)
            self._pending_loss = work
        else:
            self.lse_all.copy_(lse_local)
        return loss, (z_local, inv_norm)

    def _join_loss(self, loss: torch.Tensor) -> torch.Tensor:
        work, self._pending_loss = getattr(self, "_pending_loss", None), None
        if work is not None:
            work.wait()  # makes the current stream wait for the collective; no host sync
        return loss / self.M

    # -- "torch" mode: backward -----------------------------------------------------------
    def backward(self, state, grad_out: torch.Tensor | None = None) -> torch.Tensor:
        z_local, inv_norm = state[0], state[1]
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

    # -- the fused step -------------------------------------------------------------------
    def forward_backward(self, z_local: torch.Tensor, grad_out: torch.Tensor | None = None):
        if tuple(z_local.shape) != (self.rows, self.D):
            raise ValueError(f"rank {self.rank} expects a [{self.rows}, {self.D}] shard, got {tuple(z_local.shape)}")
        replay = self.use_cuda_graph and z_local.is_cuda
        if self._last_was_replay is not None and replay != self._last_was_replay:
            self._fence()  # eager and replayed steps pick their buffer set independently
        self._last_was_replay = replay
        if replay:
            return self._replay(z_local, grad_out)
        return self._step(z_local, grad_out)

    def _step(self, z_local, grad_out, out=None):
        if self.comm != "torch":
            return self._ccall(z_local, grad_out, out)
        loss, state = self.forward(z_local)
        dz = self.backward(state, grad_out)
        return self._join_loss(loss), dz

    def _fence(self) -> None:
        """Every rank has finished everything it was given: after this, two consecutive steps may
        use the same buffer set (the push path otherwise relies on their alternation)."""
        if torch.cuda.is_available() and self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        if self.world > 1:
            dist.barrier(group=self.group)

    def release(self) -> None:
        """Drop the captured graphs and the communicator.  Call this before
        torch.distributed.destroy_process_group(): tearing a NCCL communicator down while a
        live CUDA graph still holds its kernels hangs."""
        self._graphs = []
        self._z_static = self._g_static = None
        if torch.cuda.is_available() and self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
        if self.ccomm is not None:
            self._fence()
            self.ccomm.destroy()
            self.ccomm = None

    def _replay(self, z_local, grad_out):
        """Capture the step once per buffer set (static inputs and outputs), then copy-in + replay,
        alternating between the two graphs.  The returned loss / dz are a graph's static outputs:
        consume them before the call after next."""
        if not self._graphs:
            self._z_static = z_local.detach().clone()
            self._g_static = torch.ones(1, dtype=torch.float32, device=z_local.device)
            side = torch.cuda.Stream(device=z_local.device)
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):
                for _ in range(2):  # warm-up on the side stream (communicators, lazy inits); both sets
                    self._step(self._z_static, self._g_static)
            torch.cuda.current_stream().wait_stream(side)
            self._fence()  # the captures below start on set 0 again: nobody may still be in a warm-up step
            for _ in range(2):
                out = (torch.empty(1, dtype=torch.float32, device=z_local.device), torch.empty_like(self._z_static))
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph):
                    loss, dz = self._step(self._z_static, self._g_static, out if self.comm != "torch" else None)
                self._graphs.append((graph, loss, dz))
            self._fence()
        k = self._gstep & 1
        self._gstep += 1
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
        if engine.comm != "torch":
            # the communicator runs forward and backward as one call: keep d loss / d z for backward
            loss, dz = engine.forward_backward(z_local.detach())
            ctx.dz = dz.clone() if engine.use_cuda_graph else dz
            return loss.clone() if engine.use_cuda_graph else loss
        loss, state = engine.forward(z_local.detach())
        ctx.engine, ctx.state = engine, state
        return engine._join_loss(loss)

    @staticmethod
    def backward(ctx, grad_loss):
        if getattr(ctx, "dz", None) is not None:
            return (ctx.dz.float() * grad_loss.reshape(-1)[:1].float()).to(ctx.dz.dtype), None
        return ctx.engine.backward(ctx.state, grad_loss), None


def distributed_ntxent_loss(z_local: torch.Tensor, temperature: float = 0.07,
                            use_mixed_precision: bool = False, group=None,
                            backward_mode: str = "symmetric", engine: RowPartitionedNTXent | None = None):
    """Differentiable global NT-Xent loss (0-dim) from this rank's [M/R, D] shard.

    The value is the mean over all M global rows (the same on every rank); the gradient that
    flows to z_local is d(global loss)/d z_local.  Pass a prebuilt `engine` to reuse its buffers
    (building one allocates the communicator: do it once, not per step)."""
    if engine is None:
        world = dist.get_world_size(group) if dist.is_initialized() else 1
        engine = RowPartitionedNTXent(z_local.shape[0] * world, z_local.shape[1], temperature, z_local.dtype,
                                      z_local.device, group=group, backward_mode=backward_mode,
                                      use_mixed_precision=use_mixed_precision)
    return _DistributedNTXentFunction.apply(z_local, engine).squeeze(0)
