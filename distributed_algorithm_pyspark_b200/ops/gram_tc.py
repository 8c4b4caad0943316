"""Python side of the tcgen05 Gram kernel (csrc/gram_tc.cu)."""
from __future__ import annotations

import torch

from .. import _native


def supported(A: torch.Tensor, world: int = 1) -> bool:
    lib = _native.load()
    if not hasattr(lib, "dapb_gram_tc"):
        return False
    if A.dim() == 2 and (A.shape[1] * A.shape[1]) % (4 * max(world, 1)):
        return False                    # the in-kernel two-shot all-reduce cuts G into `world` slices of float4s
    return (A.is_cuda and A.dtype == torch.bfloat16 and A.dim() == 2 and A.is_contiguous() and A.shape[1] in (128, 256, 512)
            and A.shape[0] >= 64 and A.data_ptr() % 16 == 0)


class GramTC:
    """``G = A^T A`` summed over every rank's rows, result replicated (fp32 [P, P])."""

    def __init__(self, A: torch.Tensor, dist=None, heap=None, splits: int = 0):
        from ..parallel.dist import current
        self.dist = dist or current()
        self.lib = _native.load()
        self.A = A
        self.n, self.P = A.shape
        self.splits = splits
        self.G = torch.zeros(self.P, self.P, dtype=torch.float32, device=A.device)
        self.tickets = torch.zeros(1024, dtype=torch.int32, device=A.device)      # DAPB_GRAM_TC_SCRATCH_WORDS
        self.region = None
        from ..parallel.symm import default_heap
        self.heap = heap or default_heap()
        if self.dist.world > 1:
            nbytes = int(self.lib.dapb_gram_tc_payload_bytes(self.dist.world, self.P))
            self.region = self.heap.region(f"gram_tc_{self.P}", nbytes, (self.P // 128) * _native.MAX_RANKS)

    def __call__(self, allreduce: bool = True) -> torch.Tensor:
        reg = self.region
        _native.check(self.lib.dapb_gram_tc(self.heap.comm_ref(), self.A.data_ptr(), self.n, self.P, self.G.data_ptr(),
                                            self.tickets.data_ptr(), self.splits, reg.payload_off if reg else 0,
                                            reg.flag_base if reg else 0, reg.next_epoch() if (reg and allreduce) else 1,
                                            int(allreduce), _native.stream_ptr()), "gram_tc")
        return self.G
