"""Out-of-bounds writes, checked by hand: compute-sanitizer is closed on this GPU pool, so every
buffer an entry point writes (workspace, loss, lse, dz, the row-block outputs) is carved out of one
arena with a guard band of a known byte pattern on either side; after the call every band must be
intact, and the outputs must equal those of a call on ordinary allocations (so a write that
landed short of its buffer would show too).  Sizes are chosen ragged with respect to the 128-row
tiles and the column splits: the tail tiles are where an unmasked store would go wrong."""
import pytest
import torch

from conftest import seeded_z

import ntxent_b200  # noqa: F401
from ntxent_b200 import capi

pytestmark = pytest.mark.gpu

BAND = 4096  # bytes either side of every buffer
PATTERN = 0xA5


class Arena:
    """Buffers carved from one uint8 allocation, 256-byte aligned, a guard band between neighbours."""

    def __init__(self, device):
        self.device = device
        self.specs = []

    def want(self, name, shape, dtype):
        self.specs.append((name, tuple(shape), dtype))

    def build(self):
        off = BAND
        self.slots = {}
        for name, shape, dtype in self.specs:
            n = 1
            for s in shape:
                n *= s
            nbytes = n * torch.empty((), dtype=dtype).element_size()
            self.slots[name] = (off, nbytes, shape, dtype)
            off += (nbytes + 255) // 256 * 256 + BAND
        self.raw = torch.full((off,), PATTERN, dtype=torch.uint8, device=self.device)
        return self

    def get(self, name):
        off, nbytes, shape, dtype = self.slots[name]
        return self.raw[off:off + nbytes].view(dtype).view(shape)

    def check(self):
        keep = torch.ones_like(self.raw, dtype=torch.bool)
        for off, nbytes, _, _ in self.slots.values():
            keep[off:off + nbytes] = False
        bad = ((self.raw != PATTERN) & keep).nonzero().flatten()
        if bad.numel():
            first = int(bad[0])
            owner = min(self.slots.items(), key=lambda kv: min(abs(first - kv[1][0]), abs(first - kv[1][0] - kv[1][1])))
            raise AssertionError(f"{bad.numel()} guard bytes overwritten, first at arena offset {first}, "
                                 f"next to '{owner[0]}' [{owner[1][0]}, {owner[1][0] + owner[1][1]})")


CASES = [
    # M, D, dtype, tau, flags
    (512, 128, torch.float32, 0.5, 0),
    (770, 72, torch.float32, 0.5, 0),                       # SIMT family, ragged rows and D
    (1280, 64, torch.float32, 0.5, 0),                      # tf32 tiles
    (1026, 128, torch.bfloat16, 0.1, 0),                    # one row pair past a tile boundary
    (4096, 128, torch.bfloat16, 0.1, 0),                    # upper-triangle forward
    (4354, 128, torch.bfloat16, 0.1, 0),                    # upper-triangle forward, ragged last tile
    (2048, 256, torch.bfloat16, 0.1, 0),
    (1024, 128, torch.bfloat16, 0.1, capi.FLAG_CROSS_VIEW),
    (2050, 128, torch.float32, 0.2, capi.FLAG_FORCE_SIMT),
]


@pytest.mark.parametrize("M,D,dtype,tau,flags", CASES)
def test_whole_batch_entry_points_stay_inside_their_buffers(M, D, dtype, tau, flags):
    dev = torch.device("cuda")
    z = torch.from_numpy(seeded_z(M, D, seed=21)).to(dev).to(dtype)
    g = torch.tensor([0.75], device=dev)
    fl = capi._flags(False, flags)
    nws = capi.lib().ntxent_b200_workspace_bytes(M, D, capi._dtype_code(z), fl)
    assert nws > 0
    a = Arena(dev)
    a.want("ws", (nws,), torch.uint8)
    a.want("loss", (1,), torch.float32)
    a.want("lse", (M,), torch.float32)
    a.want("dz", (M, D), dtype)
    a.want("loss2", (1,), torch.float32)
    a.want("lse2", (M,), torch.float32)
    a.want("dz2", (M, D), dtype)
    a.build()
    # fused call
    capi.forward_backward(z, tau, g, False, flags, ws=a.get("ws"), out=(a.get("loss"), a.get("lse"), a.get("dz")))
    torch.cuda.synchronize()
    a.check()
    ref_loss, ref_lse, ref_dz = capi.forward_backward(z, tau, g, False, flags)
    assert torch.equal(a.get("loss"), ref_loss) and torch.equal(a.get("lse"), ref_lse)
    assert torch.equal(a.get("dz"), ref_dz)
    # forward, then backward, as two calls on the same arena workspace
    L = capi.lib()
    capi.check(L.ntxent_b200_forward(capi._ptr(z), M, D, float(tau), capi._dtype_code(z), fl, capi._ptr(a.get("ws")),
                                     nws, capi._ptr(a.get("loss2")), capi._ptr(a.get("lse2")), capi._stream()))
    capi.check(L.ntxent_b200_backward(capi._ptr(z), capi._ptr(a.get("lse2")), capi._ptr(g), M, D, float(tau),
                                      capi._dtype_code(z), fl, capi._ptr(a.get("ws")), nws, capi._ptr(a.get("dz2")),
                                      None, capi._stream()))
    torch.cuda.synchronize()
    a.check()
    assert torch.allclose(a.get("lse2"), ref_lse, rtol=0, atol=2e-6)
    scale = ref_dz.float().abs().max()
    assert (a.get("dz2").float() - ref_dz.float()).abs().max() <= 2e-3 * scale


@pytest.mark.parametrize("M,D,dtype,tau", [(1024, 128, torch.bfloat16, 0.1), (1152, 64, torch.float32, 0.5),
                                           (770, 72, torch.float32, 0.5)])
def test_index_vector_positives_stay_inside_their_buffers(M, D, dtype, tau):
    dev = torch.device("cuda")
    z = torch.from_numpy(seeded_z(M, D, seed=22)).to(dev).to(dtype)
    lists = capi.inverse_positives((torch.arange(M, device=dev) * 7 + 3) % M)
    fl = capi._flags(False, 0)
    nws = capi.lib().ntxent_b200_workspace_bytes(M, D, capi._dtype_code(z), fl)
    a = Arena(dev)
    a.want("ws", (nws,), torch.uint8)
    a.build()
    loss, lse, dz = capi.forward_backward_pos(z, lists, tau, ws=a.get("ws"))
    torch.cuda.synchronize()
    a.check()
    loss2, lse2, dz2 = capi.forward_backward_pos(z, lists, tau)
    assert torch.equal(loss, loss2) and torch.equal(dz, dz2)


@pytest.mark.parametrize("M,D,dtype,tau,row_begin,rows", [(2048, 128, torch.bfloat16, 0.1, 512, 512),
                                                           (1536, 128, torch.bfloat16, 0.1, 1152, 384),
                                                           (1024, 64, torch.float32, 0.5, 256, 256),
                                                           (1030, 72, torch.float32, 0.5, 515, 515)])
def test_row_block_entry_points_stay_inside_their_buffers(M, D, dtype, tau, row_begin, rows):
    dev = torch.device("cuda")
    z = torch.from_numpy(seeded_z(M, D, seed=23)).to(dev).to(dtype)
    u, inv = capi.normalize(z)
    L = capi.lib()
    op = capi._op_code(u)
    nws = L.ntxent_b200_rows_workspace_bytes(M, D, rows, op, 0)
    assert nws > 0
    a = Arena(dev)
    a.want("u", (M, D), u.dtype)
    a.want("inv", (M,), torch.float32)
    a.want("ws", (nws,), torch.uint8)
    a.want("lse_rows", (rows,), torch.float32)
    a.want("lp", (1,), torch.float32)
    a.want("du_rows", (rows, D), torch.float32)
    a.want("du_cols", (M, D), torch.float32)
    a.want("dz_rows", (rows, D), dtype)
    a.build()
    capi.check(L.ntxent_b200_normalize(capi._ptr(z), M, D, capi._dtype_code(z), 0, capi._ptr(a.get("u")),
                                       capi._ptr(a.get("inv")), capi._stream()))
    torch.cuda.synchronize()
    a.check()
    assert torch.equal(a.get("u"), u) and torch.equal(a.get("inv"), inv)
    capi.check(L.ntxent_b200_forward_rows(capi._ptr(u), M, D, row_begin, rows, float(tau), op, 0, capi._ptr(a.get("ws")),
                                          nws, capi._ptr(a.get("lse_rows")), capi._ptr(a.get("lp")), capi._stream()))
    torch.cuda.synchronize()
    a.check()
    lse_all, _ = capi.forward_rows(u, 0, M, tau)
    assert torch.allclose(a.get("lse_rows"), lse_all[row_begin:row_begin + rows], rtol=0, atol=2e-5)
    for mode, cols in ((capi.BWD_SYMMETRIC, None), (capi.BWD_CONTRACT, a.get("du_cols"))):
        capi.check(L.ntxent_b200_backward_rows(capi._ptr(u), capi._ptr(lse_all), M, D, row_begin, rows, float(tau), mode,
                                               op, 0, capi._ptr(a.get("ws")), nws, capi._ptr(a.get("du_rows")),
                                               capi._ptr(cols), capi._stream()))
        torch.cuda.synchronize()
        a.check()
    zr = z[row_begin:row_begin + rows].contiguous()
    invr = inv[row_begin:row_begin + rows].contiguous()
    capi.check(L.ntxent_b200_finalize_dz(capi._ptr(a.get("du_rows")), capi._ptr(zr), capi._ptr(invr), None, rows, D,
                                         capi._dtype_code(z), capi._ptr(a.get("dz_rows")), capi._stream()))
    torch.cuda.synchronize()
    a.check()
    assert torch.isfinite(a.get("dz_rows").float()).all()
