"""ctypes face of the C-ABI in include/ntxent_b200.h.

torch is used only for device memory and streams: every call below hands raw device
pointers, sizes and the current CUDA stream to `libntxent_b200.so`.  There is no fallback
of any kind: a missing library or a non-CUDA tensor raises.
"""
from __future__ import annotations

import ctypes
import os

import torch

from . import _build

F32, BF16 = 0, 1
FLAG_MIXED_PRECISION = 1
FLAG_FORCE_SIMT = 2
FLAG_ONLINE_MAX = 4
FLAG_CROSS_VIEW = 8
BWD_SYMMETRIC, BWD_CONTRACT = 0, 1

EXPORTS = [
    "ntxent_b200_device_supported", "ntxent_b200_strerror", "ntxent_b200_version",
    "ntxent_b200_workspace_bytes", "ntxent_b200_forward", "ntxent_b200_backward",
    "ntxent_b200_forward_backward", "ntxent_b200_operand_size", "ntxent_b200_normalize",
    "ntxent_b200_rows_workspace_bytes", "ntxent_b200_forward_rows", "ntxent_b200_backward_rows",
    "ntxent_b200_finalize_dz", "ntxent_b200_profile_begin", "ntxent_b200_profile_end",
    "ntxent_b200_plan_info", "ntxent_b200_forward_static", "ntxent_b200_debug_set_trace",
    "ntxent_b200_normalize_push", "ntxent_b200_forward_rows_push", "ntxent_b200_peer_wait",
    "ntxent_b200_backward_rows_dz", "ntxent_b200_backward_tau",
    "ntxent_b200_set_peer_wait_timeout_ms", "ntxent_b200_backward_rows_part",
    "ntxent_b200_comm_init_all", "ntxent_b200_comm_unique_id", "ntxent_b200_comm_init_rank",
    "ntxent_b200_comm_destroy", "ntxent_b200_comm_info", "ntxent_b200_comm_status",
    "ntxent_b200_dist_forward_backward", "ntxent_b200_forward_backward_pos",
]
EXCHANGE_PUSH, EXCHANGE_NCCL = 0, 1

_lib = None
MAX_PEERS = 16


class PeerList(ctypes.Structure):
    """ntxent_peer_list: the same buffer's device pointer on every rank (symmetric memory)."""
    _fields_ = [("ptr", ctypes.c_void_p * MAX_PEERS), ("count", ctypes.c_int)]

    @classmethod
    def of(cls, ptrs):
        if not 1 <= len(ptrs) <= MAX_PEERS:
            raise NTXentError(f"between 1 and {MAX_PEERS} peers supported, got {len(ptrs)}")
        pl = cls()
        for i, p in enumerate(ptrs):
            pl.ptr[i] = int(p)
        pl.count = len(ptrs)
        return pl


class PeerSignal(ctypes.Structure):
    """ntxent_peer_signal: where a producing call publishes its call number on every rank."""
    _fields_ = [("flags", PeerList), ("epoch", ctypes.c_void_p), ("rank", ctypes.c_int),
                ("ticket", ctypes.c_void_p)]


class PeerWait(ctypes.Structure):
    """ntxent_peer_wait: the local flag array a consuming kernel polls, and the value to reach."""
    _fields_ = [("flags", ctypes.c_void_p), ("epoch", ctypes.c_void_p), ("ranks", ctypes.c_int),
                ("rows_per_rank", ctypes.c_int64), ("status", ctypes.c_void_p)]


def _ref(struct):
    return ctypes.byref(struct) if struct is not None else None


class NTXentError(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Load libntxent_b200.so (building it in-tree if the sources are newer)."""
    global _lib
    if _lib is not None:
        return _lib
    if os.environ.get("NTXENT_LIBRARY"):  # A/B runs against a variant build (_build.build_library(variant=))
        path = os.environ["NTXENT_LIBRARY"]
    elif os.environ.get("NTXENT_TRACE") == "1":  # tools/trace_bwd.py: the build with time stamps
        path = _build.LIB_PATH_TRACE
        if not os.path.exists(path):
            path = _build.build_library(trace=True)
    else:
        path = _build.LIB_PATH
        if not os.path.exists(path):
            path = _build.build_library()
    L = ctypes.CDLL(path)
    vp, i64, u32, f32, sz, ci = (ctypes.c_void_p, ctypes.c_int64, ctypes.c_uint, ctypes.c_float,
                                 ctypes.c_size_t, ctypes.c_int)
    L.ntxent_b200_device_supported.restype = ci
    L.ntxent_b200_strerror.restype = ctypes.c_char_p
    L.ntxent_b200_strerror.argtypes = [ci]
    L.ntxent_b200_version.restype = ctypes.c_char_p
    L.ntxent_b200_workspace_bytes.restype = sz
    L.ntxent_b200_workspace_bytes.argtypes = [i64, i64, ci, u32]
    L.ntxent_b200_forward.restype = ci
    L.ntxent_b200_forward.argtypes = [vp, i64, i64, f32, ci, u32, vp, sz, vp, vp, vp]
    L.ntxent_b200_backward.restype = ci
    L.ntxent_b200_backward.argtypes = [vp, vp, vp, i64, i64, f32, ci, u32, vp, sz, vp, vp, vp]
    L.ntxent_b200_backward_tau.restype = ci
    L.ntxent_b200_backward_tau.argtypes = [vp, vp, vp, i64, i64, f32, ci, u32, vp, sz, vp, vp, vp]
    L.ntxent_b200_forward_backward.restype = ci
    L.ntxent_b200_forward_backward.argtypes = [vp, vp, i64, i64, f32, ci, u32, vp, sz, vp, vp, vp,
                                               vp, vp]
    L.ntxent_b200_operand_size.restype = ci
    L.ntxent_b200_operand_size.argtypes = [ci, u32]
    L.ntxent_b200_normalize.restype = ci
    L.ntxent_b200_normalize.argtypes = [vp, i64, i64, ci, u32, vp, vp, vp]
    L.ntxent_b200_rows_workspace_bytes.restype = sz
    L.ntxent_b200_rows_workspace_bytes.argtypes = [i64, i64, i64, ci, u32]
    L.ntxent_b200_forward_rows.restype = ci
    L.ntxent_b200_forward_rows.argtypes = [vp, i64, i64, i64, i64, f32, ci, u32, vp, sz, vp, vp, vp]
    L.ntxent_b200_backward_rows.restype = ci
    L.ntxent_b200_backward_rows.argtypes = [vp, vp, i64, i64, i64, i64, f32, ci, ci, u32, vp, sz,
                                            vp, vp, vp]
    L.ntxent_b200_finalize_dz.restype = ci
    L.ntxent_b200_finalize_dz.argtypes = [vp, vp, vp, vp, i64, i64, ci, vp, vp]
    ip, fp = ctypes.POINTER(ctypes.c_int), ctypes.POINTER(ctypes.c_float)
    L.ntxent_b200_profile_begin.restype = ci
    L.ntxent_b200_profile_end.restype = ci
    L.ntxent_b200_profile_end.argtypes = [fp, ip, fp, ip]
    L.ntxent_b200_plan_info.restype = ci
    L.ntxent_b200_plan_info.argtypes = [i64, i64, i64, ci, u32, ip, ip, ip]
    pp = ctypes.POINTER(PeerList)
    L.ntxent_b200_normalize_push.restype = ci
    ps, pw = ctypes.POINTER(PeerSignal), ctypes.POINTER(PeerWait)
    L.ntxent_b200_normalize_push.argtypes = [vp, i64, i64, ci, u32, pp, i64, vp, ps, vp]
    L.ntxent_b200_forward_rows_push.restype = ci
    L.ntxent_b200_forward_rows_push.argtypes = [vp, i64, i64, i64, i64, f32, ci, u32, vp, sz, pp, pp, ci,
                                                pw, ps, vp]
    L.ntxent_b200_peer_wait.restype = ci
    L.ntxent_b200_peer_wait.argtypes = [pw, vp]
    L.ntxent_b200_set_peer_wait_timeout_ms.restype = ci
    L.ntxent_b200_set_peer_wait_timeout_ms.argtypes = [ctypes.c_uint64]
    L.ntxent_b200_backward_rows_dz.restype = ci
    L.ntxent_b200_backward_rows_dz.argtypes = [vp, vp, i64, i64, i64, i64, f32, ci, u32, vp, sz, vp, vp, vp,
                                               vp, pw, vp, ci, vp, vp]
    L.ntxent_b200_backward_rows_part.restype = ci
    L.ntxent_b200_backward_rows_part.argtypes = [vp, vp, i64, i64, i64, i64, f32, ci, ci, u32, vp, sz, vp, vp]
    vpp, i64p = ctypes.POINTER(ctypes.c_void_p), ctypes.POINTER(ctypes.c_int64)
    L.ntxent_b200_comm_init_all.restype = ci
    L.ntxent_b200_comm_init_all.argtypes = [ci, ip, i64, i64, ci, u32, vpp]
    L.ntxent_b200_comm_unique_id.restype = ci
    L.ntxent_b200_comm_unique_id.argtypes = [vp]
    L.ntxent_b200_comm_init_rank.restype = ci
    L.ntxent_b200_comm_init_rank.argtypes = [vp, ci, ci, i64, i64, ci, u32, vpp]
    L.ntxent_b200_comm_destroy.restype = ci
    L.ntxent_b200_comm_destroy.argtypes = [vp]
    L.ntxent_b200_comm_info.restype = ci
    L.ntxent_b200_comm_info.argtypes = [vp, ip, ip, i64p, ip]
    L.ntxent_b200_comm_status.restype = ci
    L.ntxent_b200_comm_status.argtypes = [vp]
    L.ntxent_b200_dist_forward_backward.restype = ci
    L.ntxent_b200_dist_forward_backward.argtypes = [vp, vp, vp, f32, ci, ci, vp, vp, vp]
    L.ntxent_b200_forward_backward_pos.restype = ci
    L.ntxent_b200_forward_backward_pos.argtypes = [vp, vp, vp, vp, vp, i64, i64, f32, ci, u32, vp, sz, vp, vp, vp,
                                                   vp]
    L.ntxent_b200_debug_set_trace.restype = ci
    L.ntxent_b200_debug_set_trace.argtypes = [vp]
    L.ntxent_b200_forward_static.restype = ci
    L.ntxent_b200_forward_static.argtypes = [vp, i64, i64, f32, ci, u32, vp, vp, vp]
    _lib = L
    return L


def profile_begin() -> None:
    check(lib().ntxent_b200_profile_begin())


def profile_end():
    """-> dict(fwd_ms, fwd_launches, bwd_ms, bwd_launches) summed since profile_begin()."""
    fm, bm, fn, bn = ctypes.c_float(), ctypes.c_float(), ctypes.c_int(), ctypes.c_int()
    check(lib().ntxent_b200_profile_end(ctypes.byref(fm), ctypes.byref(fn), ctypes.byref(bm),
                                        ctypes.byref(bn)))
    return dict(fwd_ms=fm.value, fwd_launches=fn.value, bwd_ms=bm.value, bwd_launches=bn.value)


def plan_info(M: int, D: int, row_count: int, dtype_code: int, flags: int = 0):
    tc, fs, bs = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
    check(lib().ntxent_b200_plan_info(M, D, row_count, dtype_code, flags, ctypes.byref(tc),
                                      ctypes.byref(fs), ctypes.byref(bs)))
    return dict(tcgen05=bool(tc.value), fwd_splits=fs.value, bwd_splits=bs.value)


def check(rc: int) -> None:
    if rc != 0:
        raise NTXentError(f"ntxent_b200 error {rc}: {lib().ntxent_b200_strerror(rc).decode()}")


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.float32:
        return F32
    if t.dtype == torch.bfloat16:
        return BF16
    raise NTXentError(f"unsupported dtype {t.dtype}: NT-Xent embeddings must be float32 or bfloat16")


def _require_cuda(*tensors) -> None:
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise NTXentError("ntxent_b200 runs on CUDA tensors only (there is no CPU path)")


def _ptr(t):
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _flags(use_mixed_precision: bool, extra: int) -> int:
    return (FLAG_MIXED_PRECISION if use_mixed_precision else 0) | int(extra)


def operand_dtype(z_dtype: torch.dtype, flags: int) -> torch.dtype:
    code = BF16 if z_dtype == torch.bfloat16 else F32
    return torch.bfloat16 if lib().ntxent_b200_operand_size(code, flags) == 2 else torch.float32


def workspace(M: int, D: int, z: torch.Tensor, flags: int) -> torch.Tensor:
    n = lib().ntxent_b200_workspace_bytes(M, D, _dtype_code(z), flags)
    if n == 0:
        raise NTXentError(f"unsupported NT-Xent shape M={M}, D={D} (need even M >= 2, D % 8 == 0, "
                          f"8 <= D <= 1024)")
    return torch.empty(n, dtype=torch.uint8, device=z.device)


def forward(z: torch.Tensor, tau: float, use_mixed_precision: bool = False, flags: int = 0,
            ws: torch.Tensor | None = None):
    """-> (loss[1] fp32, lse[M] fp32), both on z's device; asynchronous."""
    _require_cuda(z)
    z = z.contiguous()
    M, D = z.shape
    fl = _flags(use_mixed_precision, flags)
    with torch.cuda.device(z.device):
        ws = workspace(M, D, z, fl) if ws is None else ws
        loss = torch.empty(1, dtype=torch.float32, device=z.device)
        lse = torch.empty(M, dtype=torch.float32, device=z.device)
        check(lib().ntxent_b200_forward(_ptr(z), M, D, float(tau), _dtype_code(z), fl, _ptr(ws),
                                        ws.numel(), _ptr(loss), _ptr(lse), _stream()))
    return loss, lse


def backward(z: torch.Tensor, lse: torch.Tensor, grad_out: torch.Tensor | None, tau: float,
             use_mixed_precision: bool = False, flags: int = 0, ws: torch.Tensor | None = None,
             want_pos_prob: bool = True):
    """-> (dz[M,D] in z's dtype, pos_prob[M] fp32 or None)."""
    _require_cuda(z, lse, grad_out)
    z = z.contiguous()
    lse = lse.contiguous().float()
    M, D = z.shape
    if lse.numel() != M:
        raise NTXentError(f"backward expects the forward's row log-sum-exp of {M} entries, got a "
                          f"tensor of shape {tuple(lse.shape)} (an [M,M] softmax is never formed)")
    if grad_out is not None:
        grad_out = grad_out.reshape(-1)[:1].contiguous().float()
    fl = _flags(use_mixed_precision, flags)
    with torch.cuda.device(z.device):
        ws = workspace(M, D, z, fl) if ws is None else ws
        dz = torch.empty_like(z)
        pp = torch.empty(M, dtype=torch.float32, device=z.device) if want_pos_prob else None
        check(lib().ntxent_b200_backward(_ptr(z), _ptr(lse), _ptr(grad_out), M, D, float(tau),
                                         _dtype_code(z), fl, _ptr(ws), ws.numel(), _ptr(dz), _ptr(pp),
                                         _stream()))
    return dz, pp


def backward_tau(z: torch.Tensor, lse: torch.Tensor, grad_out: torch.Tensor | None, tau: float,
                 use_mixed_precision: bool = False, flags: int = 0, ws: torch.Tensor | None = None):
    """-> (dz[M,D] in z's dtype, tau_grad_rows[M] fp32); d loss / d tau = grad_out * tau_grad_rows.sum()."""
    _require_cuda(z, lse, grad_out)
    z = z.contiguous()
    lse = lse.contiguous().float()
    M, D = z.shape
    if lse.numel() != M:
        raise NTXentError(f"backward expects the forward's row log-sum-exp of {M} entries, got a "
                          f"tensor of shape {tuple(lse.shape)}")
    if grad_out is not None:
        grad_out = grad_out.reshape(-1)[:1].contiguous().float()
    fl = _flags(use_mixed_precision, flags)
    with torch.cuda.device(z.device):
        ws = workspace(M, D, z, fl) if ws is None else ws
        dz = torch.empty_like(z)
        rows = torch.empty(M, dtype=torch.float32, device=z.device)
        check(lib().ntxent_b200_backward_tau(_ptr(z), _ptr(lse), _ptr(grad_out), M, D, float(tau),
                                             _dtype_code(z), fl, _ptr(ws), ws.numel(), _ptr(dz), _ptr(rows),
                                             _stream()))
    return dz, rows


def forward_backward(z: torch.Tensor, tau: float, grad_out: torch.Tensor | None = None,
                     use_mixed_precision: bool = False, flags: int = 0,
                     ws: torch.Tensor | None = None, out=None):
    """Fused call -> (loss[1], lse[M], dz[M,D]).  `out` may carry preallocated (loss, lse, dz)."""
    _require_cuda(z, grad_out)
    z = z.contiguous()
    M, D = z.shape
    fl = _flags(use_mixed_precision, flags)
    with torch.cuda.device(z.device):
        ws = workspace(M, D, z, fl) if ws is None else ws
        if out is None:
            loss = torch.empty(1, dtype=torch.float32, device=z.device)
            lse = torch.empty(M, dtype=torch.float32, device=z.device)
            dz = torch.empty_like(z)
        else:
            loss, lse, dz = out
        check(lib().ntxent_b200_forward_backward(_ptr(z), _ptr(grad_out), M, D, float(tau),
                                                 _dtype_code(z), fl, _ptr(ws), ws.numel(), _ptr(loss),
                                                 _ptr(lse), _ptr(dz), None, _stream()))
    return loss, lse, dz


def inverse_positives(pos: torch.Tensor):
    """pos int[M] -> (pos int32, inv_ptr int32[M+1], inv_idx int32[M]): inv_idx[inv_ptr[a]:inv_ptr[a+1]]
    are the rows whose positive is row a (stable order, so the gradient is reproducible)."""
    M = pos.numel()
    pos64 = pos.reshape(-1).to(torch.int64)
    if M and (int(pos64.min()) < 0 or int(pos64.max()) >= M or bool((pos64 == torch.arange(M, device=pos.device)).any())):
        raise NTXentError("positives must be row indices in [0, M) and no row may be its own positive")
    inv_idx = torch.argsort(pos64, stable=True).to(torch.int32)
    inv_ptr = torch.zeros(M + 1, dtype=torch.int64, device=pos.device)
    inv_ptr[1:] = torch.cumsum(torch.bincount(pos64, minlength=M), 0)
    return pos64.to(torch.int32).contiguous(), inv_ptr.to(torch.int32).contiguous(), inv_idx.contiguous()


def forward_backward_pos(z: torch.Tensor, pos_lists, tau: float, grad_out: torch.Tensor | None = None,
                         use_mixed_precision: bool = False, flags: int = 0, ws: torch.Tensor | None = None,
                         want_dz: bool = True):
    """NT-Xent whose positives come from an index vector (ntxent_b200_forward_backward_pos).
    pos_lists = inverse_positives(pos).  -> (loss[1], lse[M], dz[M,D] or None)."""
    _require_cuda(z, grad_out, *pos_lists)
    z = z.contiguous()
    M, D = z.shape
    pos, inv_ptr, inv_idx = pos_lists
    if pos.numel() != M:
        raise NTXentError(f"need one positive per row: {pos.numel()} indices for {M} rows")
    if grad_out is not None:
        grad_out = grad_out.reshape(-1)[:1].contiguous().float()
    fl = _flags(use_mixed_precision, flags)
    with torch.cuda.device(z.device):
        ws = workspace(M, D, z, fl) if ws is None else ws
        loss = torch.empty(1, dtype=torch.float32, device=z.device)
        lse = torch.empty(M, dtype=torch.float32, device=z.device)
        dz = torch.empty_like(z) if want_dz else None
        check(lib().ntxent_b200_forward_backward_pos(_ptr(z), _ptr(pos), _ptr(inv_ptr), _ptr(inv_idx), _ptr(grad_out),
                                                     M, D, float(tau), _dtype_code(z), fl, _ptr(ws), ws.numel(),
                                                     _ptr(loss), _ptr(lse), _ptr(dz), _stream()))
    return loss, lse, dz


# ---- row-partitioned building blocks -----------------------------------------------
def normalize(z_rows: torch.Tensor, flags: int = 0):
    """-> (u[rows,D] in the operand dtype, inv_norm[rows] fp32)."""
    _require_cuda(z_rows)
    z_rows = z_rows.contiguous()
    rows, D = z_rows.shape
    with torch.cuda.device(z_rows.device):
        u = torch.empty(rows, D, dtype=operand_dtype(z_rows.dtype, flags), device=z_rows.device)
        inv = torch.empty(rows, dtype=torch.float32, device=z_rows.device)
        check(lib().ntxent_b200_normalize(_ptr(z_rows), rows, D, _dtype_code(z_rows), flags, _ptr(u),
                                          _ptr(inv), _stream()))
    return u, inv


def _op_code(u_all: torch.Tensor) -> int:
    return BF16 if u_all.dtype == torch.bfloat16 else F32


def rows_workspace(M: int, D: int, row_count: int, u_all: torch.Tensor, flags: int) -> torch.Tensor:
    n = lib().ntxent_b200_rows_workspace_bytes(M, D, row_count, _op_code(u_all), flags)
    if n == 0:
        raise NTXentError(f"unsupported NT-Xent row block M={M}, D={D}, rows={row_count}")
    return torch.empty(n, dtype=torch.uint8, device=u_all.device)


def forward_rows(u_all: torch.Tensor, row_begin: int, row_count: int, tau: float, flags: int = 0,
                 ws: torch.Tensor | None = None):
    """-> (lse_rows[row_count], loss_partial[1]) for rows [row_begin, row_begin+row_count)."""
    _require_cuda(u_all)
    u_all = u_all.contiguous()
    M, D = u_all.shape
    with torch.cuda.device(u_all.device):
        ws = rows_workspace(M, D, row_count, u_all, flags) if ws is None else ws
        lse = torch.empty(row_count, dtype=torch.float32, device=u_all.device)
        lp = torch.empty(1, dtype=torch.float32, device=u_all.device)
        check(lib().ntxent_b200_forward_rows(_ptr(u_all), M, D, row_begin, row_count, float(tau),
                                             _op_code(u_all), flags, _ptr(ws), ws.numel(), _ptr(lse),
                                             _ptr(lp), _stream()))
    return lse, lp


def backward_rows(u_all: torch.Tensor, lse_all: torch.Tensor, row_begin: int, row_count: int,
                  tau: float, mode: int = BWD_SYMMETRIC, flags: int = 0,
                  ws: torch.Tensor | None = None):
    """-> (du_rows[row_count,D] fp32, du_cols[M,D] fp32 or None)."""
    _require_cuda(u_all, lse_all)
    u_all = u_all.contiguous()
    M, D = u_all.shape
    with torch.cuda.device(u_all.device):
        ws = rows_workspace(M, D, row_count, u_all, flags) if ws is None else ws
        du_rows = torch.empty(row_count, D, dtype=torch.float32, device=u_all.device)
        du_cols = (torch.empty(M, D, dtype=torch.float32, device=u_all.device)
                   if mode == BWD_CONTRACT else None)
        check(lib().ntxent_b200_backward_rows(_ptr(u_all), _ptr(lse_all), M, D, row_begin, row_count,
                                              float(tau), mode, _op_code(u_all), flags, _ptr(ws),
                                              ws.numel(), _ptr(du_rows), _ptr(du_cols), _stream()))
    return du_rows, du_cols


def finalize_dz(du_rows: torch.Tensor, z_rows: torch.Tensor, inv_norm: torch.Tensor,
                grad_out: torch.Tensor | None = None) -> torch.Tensor:
    _require_cuda(du_rows, z_rows, inv_norm, grad_out)
    z_rows = z_rows.contiguous()
    rows, D = z_rows.shape
    if grad_out is not None:
        grad_out = grad_out.reshape(-1)[:1].contiguous().float()
    with torch.cuda.device(z_rows.device):
        dz = torch.empty_like(z_rows)
        check(lib().ntxent_b200_finalize_dz(_ptr(du_rows.contiguous()), _ptr(z_rows), _ptr(inv_norm),
                                            _ptr(grad_out), rows, D, _dtype_code(z_rows), _ptr(dz),
                                            _stream()))
    return dz


STATIC_SCRATCH_BYTES = 8 << 20


def forward_static(z: torch.Tensor, tau: float, use_mixed_precision: bool = False, flags: int = 0):
    """Forward through ntxent_b200_forward_static: no workspace is passed, the library's built-in
    scratch array (STATIC_SCRATCH_BYTES per GPU) is used instead.  For callers that cannot
    allocate; larger problems raise.  -> (loss[1], lse[M])."""
    _require_cuda(z)
    z = z.contiguous()
    M, D = z.shape
    fl = _flags(use_mixed_precision, flags)
    with torch.cuda.device(z.device):
        loss = torch.empty(1, dtype=torch.float32, device=z.device)
        lse = torch.empty(M, dtype=torch.float32, device=z.device)
        check(lib().ntxent_b200_forward_static(_ptr(z), M, D, float(tau), _dtype_code(z), fl,
                                               _ptr(loss), _ptr(lse), _stream()))
    return loss, lse


def normalize_push(z_rows: torch.Tensor, u_all_peers: "PeerList", row_begin: int, flags: int = 0,
                   inv_norm: torch.Tensor | None = None, signal: "PeerSignal | None" = None) -> torch.Tensor:
    """Normalise this rank's rows and store them at `row_begin` of every rank's u_all; -> inv_norm.
    With `signal` the kernel also publishes "rows landed" to every rank's flag array."""
    _require_cuda(z_rows)
    z_rows = z_rows.contiguous()
    rows, D = z_rows.shape
    with torch.cuda.device(z_rows.device):
        if inv_norm is None:
            inv_norm = torch.empty(rows, dtype=torch.float32, device=z_rows.device)
        check(lib().ntxent_b200_normalize_push(_ptr(z_rows), rows, D, _dtype_code(z_rows), flags,
                                               ctypes.byref(u_all_peers), row_begin, _ptr(inv_norm),
                                               _ref(signal), _stream()))
    return inv_norm


def forward_rows_push(u_all: torch.Tensor, row_begin: int, row_count: int, tau: float,
                      lse_all_peers: "PeerList", loss_parts_peers: "PeerList", rank: int, flags: int = 0,
                      ws: torch.Tensor | None = None, u_ready: "PeerWait | None" = None,
                      lse_signal: "PeerSignal | None" = None) -> None:
    """forward_rows whose l_i and partial loss are stored into every rank's lse_all / loss_parts.
    `u_ready`: the tile kernel waits, tile by tile, for the peers' normalize_push; `lse_signal`:
    publishes this call's results."""
    _require_cuda(u_all)
    M, D = u_all.shape
    with torch.cuda.device(u_all.device):
        ws = rows_workspace(M, D, row_count, u_all, flags) if ws is None else ws
        check(lib().ntxent_b200_forward_rows_push(_ptr(u_all), M, D, row_begin, row_count, float(tau),
                                                  _op_code(u_all), flags, _ptr(ws), ws.numel(),
                                                  ctypes.byref(lse_all_peers), ctypes.byref(loss_parts_peers),
                                                  rank, _ref(u_ready), _ref(lse_signal), _stream()))


def set_peer_wait_timeout_ms(ms: int) -> None:
    """How long a kernel waits for a peer's flag before it marks the step void (0: for ever)."""
    check(lib().ntxent_b200_set_peer_wait_timeout_ms(int(ms)))


def peer_wait(ready: "PeerWait", device) -> None:
    """Make the current stream wait until every slot of `ready` has been published."""
    with torch.cuda.device(device):
        check(lib().ntxent_b200_peer_wait(ctypes.byref(ready), _stream()))


def backward_rows_dz(u_all: torch.Tensor, lse_all: torch.Tensor, row_begin: int, row_count: int, tau: float,
                     z_rows: torch.Tensor, inv_norm: torch.Tensor, grad_out: torch.Tensor | None = None,
                     flags: int = 0, ws: torch.Tensor | None = None,
                     lse_ready: "PeerWait | None" = None) -> torch.Tensor:
    """Symmetric-form backward of rows [row_begin, row_begin+row_count) straight to dz (z dtype)."""
    _require_cuda(u_all, lse_all, z_rows, inv_norm, grad_out)
    M, D = u_all.shape
    z_rows = z_rows.contiguous()
    if grad_out is not None:
        grad_out = grad_out.reshape(-1)[:1].contiguous().float()
    with torch.cuda.device(u_all.device):
        ws = rows_workspace(M, D, row_count, u_all, flags) if ws is None else ws
        dz = torch.empty_like(z_rows)
        check(lib().ntxent_b200_backward_rows_dz(_ptr(u_all), _ptr(lse_all), M, D, row_begin, row_count,
                                                 float(tau), _dtype_code(z_rows), flags, _ptr(ws), ws.numel(),
                                                 _ptr(z_rows), _ptr(inv_norm), _ptr(grad_out), _ptr(dz),
                                                 _ref(lse_ready), None, 0, None, _stream()))
    return dz


# ---- communicator (include/ntxent_b200.h, "Communicator") ----------------------------------
def comm_unique_id() -> bytes:
    """128 opaque bytes made by rank 0; hand them to every rank (e.g. a torch.distributed broadcast)."""
    buf = ctypes.create_string_buffer(128)
    check(lib().ntxent_b200_comm_unique_id(buf))
    return buf.raw


class Communicator:
    """One rank's handle on ntxent_b200_comm_*: owns the peer-mapped buffers, the row workspace and
    (when NCCL could be loaded) a NCCL communicator.  `forward_backward(z_rows)` is the whole
    row-partitioned step as one C-ABI call."""

    def __init__(self, handle, device, dtype):
        self.handle, self.device, self.dtype = handle, torch.device(device), dtype
        rank, nranks, has = ctypes.c_int(), ctypes.c_int(), ctypes.c_int()
        rows = ctypes.c_int64()
        check(lib().ntxent_b200_comm_info(handle, ctypes.byref(rank), ctypes.byref(nranks), ctypes.byref(rows),
                                          ctypes.byref(has)))
        self.rank, self.nranks, self.rows, self.has_nccl = rank.value, nranks.value, rows.value, bool(has.value)

    @classmethod
    def init_rank(cls, unique_id: bytes, nranks: int, rank: int, M: int, D: int, dtype: torch.dtype, device,
                  flags: int = 0):
        h = ctypes.c_void_p()
        code = BF16 if dtype == torch.bfloat16 else F32
        with torch.cuda.device(device):
            check(lib().ntxent_b200_comm_init_rank(ctypes.c_char_p(unique_id), nranks, rank, M, D, code, flags,
                                                   ctypes.byref(h)))
        return cls(h, device, dtype)

    @classmethod
    def init_all(cls, devices, M: int, D: int, dtype: torch.dtype, flags: int = 0):
        """Single process: one communicator per device index in `devices`."""
        n = len(devices)
        devs = (ctypes.c_int * n)(*[int(d) for d in devices])
        hs = (ctypes.c_void_p * n)()
        code = BF16 if dtype == torch.bfloat16 else F32
        check(lib().ntxent_b200_comm_init_all(n, devs, M, D, code, flags, hs))
        return [cls(ctypes.c_void_p(hs[i]), torch.device("cuda", int(devices[i])), dtype) for i in range(n)]

    def forward_backward(self, z_rows: torch.Tensor, tau: float, grad_out: torch.Tensor | None = None,
                         exchange: int = EXCHANGE_PUSH, bwd_mode: int = BWD_SYMMETRIC, out=None):
        """-> (loss[1] global mean, dz_rows).  `out` may carry preallocated (loss, dz)."""
        _require_cuda(z_rows, grad_out)
        z_rows = z_rows.contiguous()
        if grad_out is not None:
            grad_out = grad_out.reshape(-1)[:1].contiguous().float()
        with torch.cuda.device(self.device):
            if out is None:
                loss = torch.empty(1, dtype=torch.float32, device=self.device)
                dz = torch.empty_like(z_rows)
            else:
                loss, dz = out
            check(lib().ntxent_b200_dist_forward_backward(self.handle, _ptr(z_rows), _ptr(grad_out), float(tau),
                                                          exchange, bwd_mode, _ptr(loss), _ptr(dz), _stream()))
        return loss, dz

    def status(self) -> int:
        """0, or 1 + the rank a kernel of this rank gave up waiting for.  Synchronises."""
        return int(lib().ntxent_b200_comm_status(self.handle))

    def destroy(self) -> None:
        if self.handle is not None:
            check(lib().ntxent_b200_comm_destroy(self.handle))
            self.handle = None
