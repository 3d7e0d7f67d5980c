"""Shows that tests/test_parity_gpu.py::test_config_c4_full_size notices ONE lost 128 x 128 tile.

Builds a variant of the library whose backward tile kernel drops the weights of tile (row tile
100, column tile 37) of the 2N = 32768 problem (-DNTXENT_DROP_TILE, csrc/tc_kernels.cu), then
compares dz of the product build and of the faulty build with the rounding-faithful restatement
(tests/emulation.py) and with each other.  The test's bound is 2e-4; the product build must stay
below it and the faulty build above it.  Run on a B200:  python tools/drop_tile_proof.py
"""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CHILD = r"""
import sys, numpy as np, torch
sys.path.insert(0, %(root)r); sys.path.insert(0, %(root)r + "/tests")
import ntxent_b200
from ntxent_b200 import capi
from conftest import seeded_z
from emulation import emulate_bf16_step
M, D, tau = 32768, 128, 0.1
# bf16 values stored as fp32 + use_mixed_precision: the bf16 kernels, with dz returned unquantised
z = torch.from_numpy(seeded_z(M, D)).cuda().to(torch.bfloat16).float()
loss, lse, dz = capi.forward_backward(z, tau, None, True)
el, _, edz = emulate_bf16_step(z, tau)
d = (dz.float() - edz).abs()
err = (d.max() / edz.abs().max()).item()
tile = int(d.reshape(M // 128, -1).max(1).values.argmax().item())
print(f"dz vs restatement: {err:.3e} (worst row tile {tile}); bound 2e-4 -> {'PASS' if err <= 2e-4 else 'FAIL'}")
"""


def main():
    import ntxent_b200  # noqa: F401
    from ntxent_b200 import _build
    bad = _build.build_library(variant="droptile", defines=["-DNTXENT_DROP_TILE=1"])
    for name, lib in (("product build", _build.LIB_PATH), ("build that drops tile (100, 37)", bad)):
        env = dict(os.environ, NTXENT_LIBRARY=lib)
        out = subprocess.run([sys.executable, "-c", CHILD % {"root": ROOT}], env=env, capture_output=True,
                             text=True)
        print(f"{name}: {out.stdout.strip() or out.stderr.strip()[-400:]}")


if __name__ == "__main__":
    main()
