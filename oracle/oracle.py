"""Python face of the CPU oracle.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl
reference legs may import this module (see oracle/ntxent_oracle.c for the full
statement, including "parity unpinned" and what the restatement follows in
/root/reference).  Two restatements live here:

* ``libntxent_oracle.so`` (C, fp64 for parity + fp32/OpenMP row-block form for
  the timed CPU baseline), loaded through ctypes;
* ``numpy_forward_backward`` -- an independent dense numpy fp64 version for
  small cases, used to cross-check the C code.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libntxent_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the C oracle in place (gcc via oracle/Makefile)."""
    src = os.path.join(_HERE, "ntxent_oracle.c")
    if (force or not os.path.exists(_LIB_PATH)
            or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src)):
        subprocess.run(["make", "-C", _HERE, "-B", "libntxent_oracle.so"], check=True,
                       stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = ctypes.CDLL(_LIB_PATH)
        dp = ctypes.POINTER(ctypes.c_double)
        fp = ctypes.POINTER(ctypes.c_float)
        i64 = ctypes.c_int64
        L.ntxent_oracle_num_threads.restype = ctypes.c_int
        L.ntxent_oracle_set_num_threads.argtypes = [ctypes.c_int]
        L.ntxent_oracle_set_num_threads.restype = None
        L.ntxent_oracle_forward_f64.argtypes = [dp, i64, i64, ctypes.c_double, dp, dp, dp]
        L.ntxent_oracle_forward_f64.restype = ctypes.c_int
        L.ntxent_oracle_backward_f64.argtypes = [dp, dp, ctypes.c_double, i64, i64,
                                                 ctypes.c_double, dp, dp]
        L.ntxent_oracle_backward_f64.restype = ctypes.c_int
        L.ntxent_oracle_dtau_f64.argtypes = [dp, dp, ctypes.c_double, i64, i64, ctypes.c_double, dp]
        L.ntxent_oracle_dtau_f64.restype = ctypes.c_int
        L.ntxent_oracle_rowblock_f32.argtypes = [fp, i64, i64, ctypes.c_float, i64, i64,
                                                 fp, fp, fp, fp]
        L.ntxent_oracle_rowblock_f32.restype = ctypes.c_int
        ip64 = ctypes.POINTER(ctypes.c_int64)
        L.ntxent_oracle_forward_pos_f64.argtypes = [dp, ip64, i64, i64, i64, ctypes.c_double, dp, dp]
        L.ntxent_oracle_forward_pos_f64.restype = ctypes.c_int
        L.ntxent_oracle_backward_pos_f64.argtypes = [dp, ip64, i64, dp, ctypes.c_double, i64, i64, ctypes.c_double, dp]
        L.ntxent_oracle_backward_pos_f64.restype = ctypes.c_int
        L.ntxent_literal_reference_forward_f64.argtypes = [dp, i64, i64, ctypes.c_double, dp]
        L.ntxent_literal_reference_forward_f64.restype = ctypes.c_int
        _lib = L
    return _lib


def _dptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_double))


def _fptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


def num_threads() -> int:
    return int(lib().ntxent_oracle_num_threads())


def use_all_cores() -> int:
    """Run the OpenMP loops on every host core, whatever OMP_NUM_THREADS the launcher exported
    (torchrun sets it to 1 for its ranks).  Returns the thread count now in use."""
    lib().ntxent_oracle_set_num_threads(os.cpu_count() or 1)
    return num_threads()


def forward(z, tau: float):
    """fp64 forward.  Returns (loss, lse[M], pos_logit[M])."""
    z = np.ascontiguousarray(z, dtype=np.float64)
    M, D = z.shape
    loss = np.zeros(1)
    lse = np.zeros(M)
    sp = np.zeros(M)
    rc = lib().ntxent_oracle_forward_f64(_dptr(z), M, D, float(tau), _dptr(loss), _dptr(lse), _dptr(sp))
    if rc != 0:
        raise ValueError(f"ntxent_oracle_forward_f64 rc={rc}")
    return float(loss[0]), lse, sp


def backward(z, lse, tau: float, grad_out: float = 1.0):
    """fp64 backward.  Returns (dz[M,D], pos_prob[M])."""
    z = np.ascontiguousarray(z, dtype=np.float64)
    lse = np.ascontiguousarray(lse, dtype=np.float64)
    M, D = z.shape
    dz = np.zeros((M, D))
    pp = np.zeros(M)
    rc = lib().ntxent_oracle_backward_f64(_dptr(z), _dptr(lse), float(grad_out), M, D, float(tau),
                                          _dptr(dz), _dptr(pp))
    if rc != 0:
        raise ValueError(f"ntxent_oracle_backward_f64 rc={rc}")
    return dz, pp


def dtau(z, lse, tau: float, grad_out: float = 1.0) -> float:
    """fp64 d loss / d tau (the contract of ntxent_b200_backward_tau)."""
    z = np.ascontiguousarray(z, dtype=np.float64)
    lse = np.ascontiguousarray(lse, dtype=np.float64)
    M, D = z.shape
    out = np.zeros(1)
    rc = lib().ntxent_oracle_dtau_f64(_dptr(z), _dptr(lse), float(grad_out), M, D, float(tau), _dptr(out))
    if rc != 0:
        raise ValueError(f"ntxent_oracle_dtau_f64 rc={rc}")
    return float(out[0])


def forward_backward(z, tau: float, grad_out: float = 1.0):
    loss, lse, _ = forward(z, tau)
    dz, _ = backward(z, lse, tau, grad_out)
    return loss, lse, dz


def forward_backward_pos(z, pos, tau: float, grad_out: float = 1.0, cross_n: int = 0):
    """fp64 loss / l / dz with the positive of row i given by pos[i].
    cross_n > 0: denominators over the other view only (rows < cross_n against rows >= cross_n).
    Returns (loss, lse[M], dz[M,D])."""
    z = np.ascontiguousarray(z, dtype=np.float64)
    pos = np.ascontiguousarray(pos, dtype=np.int64)
    M, D = z.shape
    loss, lse, dz = np.zeros(1), np.zeros(M), np.zeros((M, D))
    ip = pos.ctypes.data_as(ctypes.POINTER(ctypes.c_int64))
    rc = lib().ntxent_oracle_forward_pos_f64(_dptr(z), ip, int(cross_n), M, D, float(tau), _dptr(loss), _dptr(lse))
    if rc == 0:
        rc = lib().ntxent_oracle_backward_pos_f64(_dptr(z), ip, int(cross_n), _dptr(lse), float(grad_out), M, D, float(tau),
                                                  _dptr(dz))
    if rc != 0:
        raise ValueError(f"ntxent_oracle_*_pos_f64 rc={rc}")
    return float(loss[0]), lse, dz


def forward_backward_cross(z, tau: float, grad_out: float = 1.0):
    """fp64 symmetric two-tower InfoNCE over the stacked [q; k] rows (the cross-view form of NT-Xent)."""
    M = np.asarray(z).shape[0]
    return forward_backward_pos(z, (np.arange(M) + M // 2) % M, tau, grad_out, cross_n=M // 2)


def rowblock_f32(z, tau: float, row_begin: int, row_count: int):
    """fp32/OpenMP row-block contract form (the timed CPU baseline).

    Returns (lse[row_count], loss_partial, du_row[row_count,D], du_col[M,D]);
    du_* are unscaled (multiply by 1/(tau*M) to get dU contributions).
    """
    z = np.ascontiguousarray(z, dtype=np.float32)
    M, D = z.shape
    lse = np.zeros(row_count, dtype=np.float32)
    lp = np.zeros(1, dtype=np.float32)
    du_row = np.zeros((row_count, D), dtype=np.float32)
    du_col = np.zeros((M, D), dtype=np.float32)
    rc = lib().ntxent_oracle_rowblock_f32(_fptr(z), M, D, float(tau), int(row_begin), int(row_count),
                                          _fptr(lse), _fptr(lp), _fptr(du_row), _fptr(du_col))
    if rc != 0:
        raise ValueError(f"ntxent_oracle_rowblock_f32 rc={rc}")
    return lse, float(lp[0]), du_row, du_col


def literal_reference_forward(z, T: float) -> float:
    """What src/ntxent_kernel.cu literally evaluates (NOT NT-Xent); documentation only."""
    z = np.ascontiguousarray(z, dtype=np.float64)
    B, D = z.shape
    out = np.zeros(1)
    rc = lib().ntxent_literal_reference_forward_f64(_dptr(z), B, D, float(T), _dptr(out))
    if rc != 0:
        raise ValueError(f"ntxent_literal_reference_forward_f64 rc={rc}")
    return float(out[0])


# --------------------------------------------------------------------------
# Independent numpy restatement (dense, fp64, small cases only).
# --------------------------------------------------------------------------
def numpy_forward_backward(z, tau: float, grad_out: float = 1.0, eps: float = 1e-12, pos=None,
                           cross_view: bool = False):
    """Dense numpy fp64 NT-Xent.  Returns dict(loss, lse, dz, dU, pos_prob, dtau, ...); dtau is
    d loss / d tau = -(1/tau) sum_ij G_ij S_ij.  `pos`: the positive of
    every row as an index vector (default (i + N) mod M).  `cross_view`: the denominators run over
    the OTHER view only (rows 0..N-1 against N..2N-1 and back): the symmetric two-tower InfoNCE of
    CLIP, (L(q->k) + L(k->q)) / 2 with q = rows 0..N-1 and k = rows N..2N-1."""
    z = np.asarray(z, dtype=np.float64)
    M, D = z.shape
    assert M % 2 == 0 or pos is not None
    N = M // 2
    nrm = np.maximum(np.linalg.norm(z, axis=1, keepdims=True), eps)
    u = z / nrm
    S = (u @ u.T) / tau
    idx = np.arange(M)
    pos = (idx + N) % M if pos is None else np.asarray(pos, dtype=np.int64)
    Sm = S.copy()
    Sm[idx, idx] = -np.inf
    if cross_view:
        same = (idx[:, None] >= N) == (idx[None, :] >= N)
        Sm[same] = -np.inf
    mx = Sm.max(axis=1, keepdims=True)
    lse = (mx + np.log(np.exp(Sm - mx).sum(axis=1, keepdims=True)))[:, 0]
    loss = float(np.mean(lse - S[idx, pos]))
    P = np.exp(Sm - lse[:, None])
    Y = np.zeros_like(P)
    Y[idx, pos] = 1.0
    G = (P - Y) / M
    dU = (G + G.T) @ u / tau
    dz = grad_out * (dU - (dU * u).sum(axis=1, keepdims=True) * u) / nrm
    S0 = np.where(np.isfinite(Sm), S, 0.0)  # G's diagonal is zero
    dtau = float(-grad_out * (G * S0).sum() / tau)
    return dict(loss=loss, lse=lse, dz=dz, dU=dU, pos_prob=P[idx, pos], G=G, u=u, dtau=dtau)
