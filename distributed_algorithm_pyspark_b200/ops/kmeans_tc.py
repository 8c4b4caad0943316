"""Python side of the tcgen05 k-means assignment kernel (csrc/kmeans_tc.cu)."""
from __future__ import annotations

import torch

from .. import _native


def supported(X: torch.Tensor, k: int) -> bool:
    """The tensor-core kernel covers bf16 points with d a multiple of 64 (<= 256) and k a multiple of 256."""
    lib = _native.load()
    if not hasattr(lib, "dapb_kmeans_assign_tc"):
        return False
    return (X.is_cuda and X.dtype == torch.bfloat16 and X.shape[1] % 64 == 0 and X.shape[1] <= 256
            and k % 256 == 0 and k <= 4096 and X.shape[0] >= 128)


class KMeansTC:
    def __init__(self, X: torch.Tensor, k: int):
        raise NotImplementedError
