"""Python side of the tcgen05 k-means assignment kernel (csrc/kmeans_tc.cu)."""
from __future__ import annotations

import torch

from .. import _native


def supported(X: torch.Tensor, k: int) -> bool:
    """bf16 or fp32 points with d in {64, 128}, 1 <= k <= 16384 (padded to a multiple of 128 inside), at least one full tile of
    rows.  fp32 points additionally need room for a resident bf16 shadow copy (half the size of X) and enough
    work (k * d >= 8192) for the tensor-core kernel to beat the exact CUDA-core scan."""
    lib = _native.load()
    if not hasattr(lib, "dapb_kmeans_assign_tc"):
        return False
    ok = (X.is_cuda and X.shape[1] in (64, 128) and 1 <= k <= 16384 and X.shape[0] >= 128 and X.data_ptr() % 16 == 0)
    if not ok:
        return False
    if X.dtype == torch.bfloat16:
        return True
    if X.dtype == torch.float32 and k * X.shape[1] >= 8192:
        free, _ = torch.cuda.mem_get_info(X.device)
        return free > X.numel() * 2 + (2 << 30)
    return False


class KMeansTC:
    """Holds the prepared centroid operand; ``assign_accumulate`` = prepare + fused assignment/accumulation."""

    def __init__(self, X: torch.Tensor, k: int, grid: int = 0):
        self.lib = _native.load()
        self.k = int(k)
        self.n, self.d = X.shape
        self.grid = grid
        # fp32 points: a resident bf16 shadow copy feeds the MMAs (distances to bf16 resolution, like the bf16 path),
        # the per-cluster sums are accumulated from the fp32 originals inside the same kernel
        self.X32 = None
        if X.dtype == torch.float32:
            self.X32 = X
            Xb = torch.empty(X.shape, dtype=torch.bfloat16, device=X.device)
            step = max(1, (1 << 28) // max(self.d, 1))
            for i in range(0, self.n, step):
                Xb[i:i + step].copy_(X[i:i + step])
            X = Xb
        self.X = X
        k_pad = (self.k + 127) // 128 * 128                     # the MMA's N granularity; padding centroids never win
        self.b_operand = torch.empty(k_pad, self.d, dtype=torch.bfloat16, device=X.device)
        self.cnorm = torch.empty(9 * k_pad + 8, dtype=torch.float32, device=X.device)      # norms + aux + K=16 norm slabs
        # max ||x||^2 over the resident shard (X never changes): bounds every distance, so the kernel can compare
        # fixed-exponent bit patterns instead of floats.  One chunked pass at construction.
        xmax = 0.0
        step = max(1, (1 << 28) // max(self.d, 1))
        for i in range(0, self.n, step):
            xc = X[i:i + step].float()
            xmax = max(xmax, float((xc * xc).sum(1).max().item()))
        self.xnorm_max = xmax

    def prepare(self, centers: torch.Tensor):
        """fp32 centres -> bf16(-2c) operand + biased norms (only needed when the centres did not come out of
        ``dapb_kmeans_finalize_tc``, which emits both itself)."""
        _native.check(self.lib.dapb_kmeans_tc_prepare(centers.data_ptr(), self.k, self.d, self.b_operand.data_ptr(),
                                                      self.cnorm.data_ptr(), self.xnorm_max, _native.stream_ptr()),
                      "kmeans_tc_prepare")

    def assign_accumulate(self, centers: torch.Tensor, sums: torch.Tensor, counts: torch.Tensor, assign=None,
                          prepared: bool = False):
        st = _native.stream_ptr()
        if not prepared:
            self.prepare(centers)
        if self.X32 is not None:
            _native.check(self.lib.dapb_kmeans_assign_tc_f32(self.X.data_ptr(), self.X32.data_ptr(), self.n, self.d,
                                                             self.b_operand.data_ptr(), self.cnorm.data_ptr(), self.k,
                                                             _native.ptr(assign), sums.data_ptr(), counts.data_ptr(),
                                                             self.grid, st), "kmeans_assign_tc_f32")
            return
        _native.check(self.lib.dapb_kmeans_assign_tc(self.X.data_ptr(), self.n, self.d, self.b_operand.data_ptr(),
                                                     self.cnorm.data_ptr(), self.k, _native.ptr(assign), sums.data_ptr(),
                                                     counts.data_ptr(), self.grid, st), "kmeans_assign_tc")
