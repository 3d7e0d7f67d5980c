"""CPU-side checks of the C-ABI library: it loads, exports every symbol include/*.h declares,
sizes workspaces and rejects bad arguments -- no compute (there is no GPU here), and no
fallback: a CPU tensor must raise."""
import ctypes
import glob
import os
import re

import pytest
import torch

from conftest import ROOT

import ntxent_b200
from ntxent_b200 import capi


def _declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = open(h).read()
        text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
        names |= set(re.findall(r"\b(ntxent_b200_[a-z_0-9]+)\s*\(", text))
    return names


def test_library_loads_and_exports_every_declared_symbol():
    L = capi.lib()
    declared = _declared_symbols()
    assert len(declared) >= 20
    for name in declared:
        assert hasattr(L, name), f"{name} declared in include/ but not exported"
    assert set(capi.EXPORTS) == declared
    assert b"sm_100a" in L.ntxent_b200_version()


def test_strerror_and_codes():
    L = capi.lib()
    assert L.ntxent_b200_strerror(0) == b"ok"
    for code in (-1, -2, -3, -4):
        assert len(L.ntxent_b200_strerror(code)) > 5
    assert b"unknown" in L.ntxent_b200_strerror(-99)


def test_workspace_sizing():
    L = capi.lib()
    # O(M*D) memory, never O(M^2): the reference's three [M,M] fp32 temporaries would be 12.9 GB here
    n = L.ntxent_b200_workspace_bytes(32768, 128, capi.BF16, 0)
    assert 0 < n < 512 * 2**20
    assert L.ntxent_b200_workspace_bytes(131072, 256, capi.BF16, 0) < 4 * 2**30
    assert L.ntxent_b200_workspace_bytes(31, 128, capi.F32, 0) == 0      # odd M
    assert L.ntxent_b200_workspace_bytes(32, 12, capi.F32, 0) == 0       # D % 8 != 0
    assert L.ntxent_b200_workspace_bytes(32, 128, 7, 0) == 0             # unknown dtype
    assert L.ntxent_b200_rows_workspace_bytes(4096, 128, 1024, capi.BF16, 0) > 0
    assert L.ntxent_b200_rows_workspace_bytes(4096, 128, 8192, capi.BF16, 0) == 0
    assert L.ntxent_b200_operand_size(capi.F32, 0) == 4
    assert L.ntxent_b200_operand_size(capi.F32, capi.FLAG_MIXED_PRECISION) == 2
    assert L.ntxent_b200_operand_size(capi.BF16, 0) == 2


def test_argument_errors_return_codes_without_touching_the_device():
    L = capi.lib()
    null = ctypes.c_void_p(None)
    one = ctypes.c_void_p(256)  # never dereferenced: argument checks come first
    odd = ctypes.c_void_p(260)  # misaligned
    assert L.ntxent_b200_forward(odd, 64, 128, 0.5, 0, 0, one, 1 << 30, one, one, null) == -1
    assert L.ntxent_b200_forward(one, 64, 128, 0.5, 0, 0, odd, 1 << 30, one, one, null) == -1
    assert L.ntxent_b200_forward(null, 64, 128, 0.5, 0, 0, one, 1 << 30, one, one, null) == -1
    assert L.ntxent_b200_forward(one, 63, 128, 0.5, 0, 0, one, 1 << 30, one, one, null) == -1
    assert L.ntxent_b200_forward(one, 64, 128, -1.0, 0, 0, one, 1 << 30, one, one, null) == -1
    assert L.ntxent_b200_forward(one, 64, 100, 0.5, 0, 0, one, 1 << 30, one, one, null) == -4
    assert L.ntxent_b200_forward(one, 64, 128, 0.5, 0, 0, one, 16, one, one, null) == -3
    assert L.ntxent_b200_backward(one, null, null, 64, 128, 0.5, 0, 0, one, 1 << 30, one, null, null) == -1
    assert L.ntxent_b200_backward_rows(one, one, 64, 128, 0, 32, 0.5, 5, 0, 0, one, 1 << 30, one, null,
                                       null) == -1  # unknown mode
    assert L.ntxent_b200_backward_rows(one, one, 64, 128, 0, 32, 0.5, capi.BWD_CONTRACT, 0, 0, one,
                                       1 << 30, one, null, null) == -1  # contract needs du_cols
    assert L.ntxent_b200_forward_rows(one, 64, 128, 48, 32, 0.5, 0, 0, one, 1 << 30, one, one, null) == -1


def test_compute_calls_report_an_unsupported_device():
    """Valid arguments on a machine without an sm_100 GPU: NTXENT_ERR_UNSUPPORTED_DEVICE, not a raw
    "no kernel image" from the first launch (and nothing is launched)."""
    if torch.cuda.is_available() and torch.cuda.get_device_capability()[0] == 10:
        pytest.skip("sm_100 device present")
    L = capi.lib()
    one, null = ctypes.c_void_p(256), ctypes.c_void_p(None)
    assert L.ntxent_b200_forward(one, 64, 128, 0.5, 0, 0, one, 1 << 30, one, one, null) == -2
    assert L.ntxent_b200_forward_static(one, 64, 128, 0.5, 0, 0, one, one, null) == -2
    assert L.ntxent_b200_normalize(one, 64, 128, 0, 0, one, one, null) == -2
    assert b"sm_100" in L.ntxent_b200_strerror(-2)


def test_no_cpu_fallback():
    z = torch.randn(8, 16)
    with pytest.raises(capi.NTXentError, match="CUDA"):
        ntxent_b200.forward(z, 0.5)
    with pytest.raises(capi.NTXentError, match="CUDA"):
        ntxent_b200.backward(z, torch.zeros(8), torch.ones(1), 0.5)


def test_device_supported_is_false_without_a_gpu():
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    assert capi.lib().ntxent_b200_device_supported() == 0
    assert ntxent_b200.check_tensor_core_support() is False


def _cta_of_tile(t, total, ctas):
    # tc_kernels.cu, bwd_cta_of_tile: CTA c owns the linear tiles [c*T/G, (c+1)*T/G)
    return ((t + 1) * ctas - 1) // total


@pytest.mark.parametrize("M,rows", [(32768, 32768), (32768, 16384), (32768, 4096), (8192, 8192), (2048, 2048),
                                    (4096, 512), (1030, 1030), (256, 256), (131072, 16384)])
def test_backward_tile_dealing_invariants(M, rows):
    """The tcgen05 backward deals the (row tile, column tile) pairs to persistent CTAs in equal
    contiguous shares and the epilogue re-derives, per row tile, how many partials to add.  This
    restates the dealing (csrc/tc_kernels.cu, BwdUnit) and ties it to the library's own plan:
    every tile is owned exactly once, shares differ by at most one tile, the slots of a row tile
    are 0..count-1 with count = the epilogue's formula, and the largest count is the plan's split
    number (which sizes the workspace)."""
    plan = capi.plan_info(M, 128, rows, capi.BF16)
    assert plan["tcgen05"]
    row_tiles, col_tiles = -(-rows // 128), -(-M // 128)
    total = row_tiles * col_tiles
    ctas = min(148, total)  # device_sm_count() is 148 when there is no device
    begins = [c * total // ctas for c in range(ctas + 1)]
    assert begins[0] == 0 and begins[-1] == total
    shares = [b - a for a, b in zip(begins, begins[1:])]
    assert max(shares) - min(shares) <= 1 and min(shares) >= 1
    for c in range(ctas):  # the closed form the kernel and the epilogue use agrees with the ranges
        assert _cta_of_tile(begins[c], total, ctas) == c
        assert _cta_of_tile(begins[c + 1] - 1, total, ctas) == c
    worst = 0
    for rt in range(row_tiles):
        first, last = rt * col_tiles, (rt + 1) * col_tiles - 1
        owners = list(range(_cta_of_tile(first, total, ctas), _cta_of_tile(last, total, ctas) + 1))
        slots = [c - owners[0] for c in owners]
        assert slots == list(range(len(owners)))
        # epilogue_kernels.cu: ((first + col_tiles) * G - 1) / T - ((first + 1) * G - 1) / T + 1
        assert ((first + col_tiles) * ctas - 1) // total - ((first + 1) * ctas - 1) // total + 1 == len(owners)
        worst = max(worst, len(owners))
    assert worst == plan["bwd_splits"]
