"""GPU test of ntxent_b200_forward_static, the forward entry point that allocates nothing.

Some hosts of this library cannot give it a workspace: they run inside someone else's memory
pool or had all device memory handed out at start-up, and are not allowed to call cudaMalloc.  For them the
library carries its forward scratch inside itself (8 MiB per GPU).  This file checks, through
the C-ABI, that the entry point returns what ntxent_b200_forward returns -- on two calls in a
row with different inputs and shapes, because the built-in array is shared and keeps the
previous call's contents -- and that a problem beyond the limit is refused with a clear error
instead of overrunning the array.

The file name sorts last on purpose: a run that stops at the first failure has then already
run every other test.  It passes on its own in a fresh process as well as after the others.
"""
import ctypes

import numpy as np
import pytest
import torch

from conftest import seeded_z

from ntxent_b200 import capi

pytestmark = pytest.mark.gpu


def _forward_static_raw(z, tau, flags=0):
    """Straight ctypes call: device pointers, sizes and the stream, nothing else."""
    L = capi.lib()
    M, D = z.shape
    loss = torch.empty(1, dtype=torch.float32, device=z.device)
    lse = torch.empty(M, dtype=torch.float32, device=z.device)
    code = capi.BF16 if z.dtype == torch.bfloat16 else capi.F32
    rc = L.ntxent_b200_forward_static(ctypes.c_void_p(z.data_ptr()), M, D, float(tau), code, flags,
                                      ctypes.c_void_p(loss.data_ptr()), ctypes.c_void_p(lse.data_ptr()),
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
    return rc, loss, lse


def test_forward_static_matches_forward_on_consecutive_calls():
    # tolerance of the fp32 parity tests (loss 1e-5 relative); the two entry points launch the
    # same kernels, so in practice the results are identical
    cases = [
        (torch.from_numpy(seeded_z(1024, 128, seed=41)).cuda(), 0.1),
        (torch.from_numpy(seeded_z(384, 64, seed=42)).cuda().to(torch.bfloat16), 0.07),
        (torch.from_numpy(seeded_z(1024, 128, seed=43)).cuda(), 0.5),
    ]
    for z, tau in cases:  # back to back: each call finds the previous call's leftovers
        rc, loss_s, lse_s = _forward_static_raw(z, tau)
        assert rc == 0, capi.lib().ntxent_b200_strerror(rc)
        loss_w, lse_w = capi.forward(z, tau)
        torch.cuda.synchronize()
        assert loss_s.item() == pytest.approx(loss_w.item(), rel=1e-5)
        np.testing.assert_allclose(lse_s.cpu().numpy(), lse_w.cpu().numpy(), rtol=1e-5, atol=1e-5)
    # the same input again, twice in a row, gives the same bits
    z, tau = cases[0]
    a = _forward_static_raw(z, tau)
    b = _forward_static_raw(z, tau)
    torch.cuda.synchronize()
    assert a[0] == 0 and b[0] == 0
    assert torch.equal(a[1], b[1]) and torch.equal(a[2], b[2])


def test_forward_static_refuses_what_does_not_fit():
    L = capi.lib()
    z = torch.zeros(32768, 128, dtype=torch.bfloat16, device="cuda")
    rc, _, _ = _forward_static_raw(z, 0.1)
    assert rc == -5
    msg = L.ntxent_b200_strerror(rc).decode()
    assert "NTXENT_STATIC_SCRATCH_BYTES" in msg and "ntxent_b200_forward" in msg
    with pytest.raises(capi.NTXentError, match="scratch"):
        capi.forward_static(z, 0.1)
    # bad arguments are still argument errors
    assert _forward_static_raw(torch.zeros(7, 128, device="cuda"), 0.1)[0] == -1
