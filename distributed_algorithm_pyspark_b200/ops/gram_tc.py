"""Python side of the tcgen05 Gram kernel (csrc/gram_tc.cu)."""
from __future__ import annotations

import torch

from .. import _native


def supported(A: torch.Tensor) -> bool:
    lib = _native.load()
    if not hasattr(lib, "dapb_gram_tc"):
        return False
    return A.is_cuda and A.dtype == torch.bfloat16 and A.shape[1] == 512 and A.shape[0] >= 4096


class GramTC:
    def __init__(self, A: torch.Tensor, dist):
        raise NotImplementedError
