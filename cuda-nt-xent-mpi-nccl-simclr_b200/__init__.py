"""ntxent_b200 -- B200-native NT-Xent loss engine behind the reference's operator surface.

The directory name carries the upstream project's name and cannot be imported with a plain
`import` statement; `ntxent_b200.py` at the repository root loads it under the module name
`ntxent_b200`.

Layers:
  csrc/*.cu            hand-written sm_100a kernels + the extern "C" boundary (include/ntxent_b200.h)
  csrc/binding.cpp     C++ torch extension `ntxent_cuda` with the reference's forward/backward
  capi.py              ctypes access to the same C-ABI (tests, bench)
  ops.py               the operator surface in Python: forward / backward / autograd / nn.Module
  distributed.py       row-partitioned data-parallel loss over torch.distributed (NCCL)
"""
from . import capi  # noqa: F401
from .ops import (GraphedNTXent, NTXentLoss, backward, check_tensor_core_support,  # noqa: F401
                  forward, forward_with_state, ntxent_loss, two_tower_loss)

__all__ = ["capi", "forward", "backward", "forward_with_state", "check_tensor_core_support",
           "ntxent_loss", "two_tower_loss", "NTXentLoss", "GraphedNTXent"]
