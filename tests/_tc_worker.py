"""Runs one tensor-core kernel check in its own process (a faulting kernel must not poison the pytest process).
usage: _tc_worker.py kmeans n d k grid | kmeans_ties n d k | gram n p grid | lr n D grid"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from distributed_algorithm_pyspark_b200 import _native  # noqa: E402


def kmeans(n, d, k, grid):
    from distributed_algorithm_pyspark_b200.ops.kmeans_tc import KMeansTC
    dev = "cuda"
    torch.manual_seed(n + d + k)
    centres = torch.randn(k, d, device=dev) * 3
    lab = torch.randint(0, k, (n,), device=dev)
    X = (centres[lab] + torch.randn(n, d, device=dev)).to(torch.bfloat16).contiguous()
    C = (centres + 0.3 * torch.randn(k, d, device=dev)).contiguous()          # current centres (fp32)
    op = KMeansTC(X, k, grid=grid)
    sums = torch.zeros(k, d, device=dev)
    counts = torch.zeros(k, dtype=torch.int32, device=dev)
    assign = torch.full((n,), -1, dtype=torch.int32, device=dev)
    op.assign_accumulate(C, sums, counts, assign)
    torch.cuda.synchronize()
    cb = C.to(torch.bfloat16).float()
    cn = (cb * cb).sum(1)
    dist = cn[None, :] - 2.0 * (X.float() @ cb.T)
    ref = dist.argmin(1)
    got = assign.long()
    bad_range = int(((got < 0) | (got >= k)).sum().item())
    got_c = got.clamp(0, k - 1)
    gap = dist.gather(1, got_c[:, None]).squeeze(1) - dist.gather(1, ref[:, None]).squeeze(1)
    tol = dist.abs().max(1).values * 2.0 ** -11 + 1e-3
    wrong = int((gap > tol).sum().item())
    mism = int((got != ref).sum().item())
    ref_sums = torch.zeros_like(sums).index_add_(0, got_c, X.float())
    sum_err = float((sums - ref_sums).abs().max().item() / (ref_sums.abs().max().item() + 1e-9))
    cnt_ok = bool(torch.equal(counts.long(), torch.bincount(got_c, minlength=k)))
    return {"n": n, "bad_range": bad_range, "wrong": wrong, "mismatch_vs_fp32": mism, "sum_rel_err": sum_err,
            "counts_ok": cnt_ok, "purity": float((got == lab).float().mean().item())}


def kmeans_ties(n, d, k):
    """Exact distance ties between centroids that sit in DIFFERENT 32-column groups of a chunk (and in different chunks):
    duplicates of centroid j placed at j + 37 and j + 200 must never take a row -- lowest index wins (k-means.py:25)."""
    from distributed_algorithm_pyspark_b200.ops.kmeans_tc import KMeansTC
    dev = "cuda"
    torch.manual_seed(7)
    centres = (torch.randn(k, d, device=dev) * 3).to(torch.bfloat16).float()
    dup_src = torch.arange(0, 20, device=dev)
    centres[dup_src + 37] = centres[dup_src]
    centres[dup_src + 200] = centres[dup_src]
    lab = torch.randint(0, 20, (n,), device=dev)
    X = (centres[lab] + 0.1 * torch.randn(n, d, device=dev)).to(torch.bfloat16).contiguous()
    op = KMeansTC(X, k)
    sums = torch.zeros(k, d, device=dev)
    counts = torch.zeros(k, dtype=torch.int32, device=dev)
    assign = torch.full((n,), -1, dtype=torch.int32, device=dev)
    op.assign_accumulate(centres.contiguous(), sums, counts, assign)
    torch.cuda.synchronize()
    got = assign.long()
    return {"to_duplicates": int(((got >= 37) & (got < 57)).sum().item() + ((got >= 200) & (got < 220)).sum().item()),
            "to_originals": int((got < 20).sum().item()), "n": n}


def gram(n, P, splits):
    from distributed_algorithm_pyspark_b200.ops.gram_tc import GramTC
    from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed
    dist = init_distributed()
    dev = "cuda"
    torch.manual_seed(n + P)
    A = torch.randn(n, P, device=dev).to(torch.bfloat16).contiguous()
    op = GramTC(A, dist, splits=splits)
    G1 = op().clone()
    G2 = op().clone()                       # second call: tickets / accumulator state must have been reset
    torch.cuda.synchronize()
    ref = A.double().T @ A.double()
    scale = float(ref.abs().max().item())
    return {"err1": float((G1.double() - ref).abs().max().item()) / scale,
            "err2": float((G2.double() - ref).abs().max().item()) / scale,
            "sym": float((G1 - G1.T).abs().max().item()) / scale}


def lr(n, frac_pct, grid):
    """tcgen05 logistic step vs PyTorch fp32 (gradient) and vs the CUDA-core kernel (6 SSGD steps)."""
    from distributed_algorithm_pyspark_b200.config import LogisticConfig
    from distributed_algorithm_pyspark_b200.models.logistic import ShardedLogisticEngine
    from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed
    from distributed_algorithm_pyspark_b200.utils.philox import philox_row_mask_torch
    dist = init_distributed()
    dev, D = "cuda", 1024
    torch.manual_seed(n)
    X = (torch.randn(n, D, device=dev) * 0.5).to(torch.bfloat16).contiguous()
    X[:, -1] = 1
    y = (torch.rand(n, device=dev) < 0.4).float()
    frac = frac_pct / 100.0
    cfg = LogisticConfig.for_algo("ssgd" if frac < 1 else "gd", mini_batch_fraction=frac, sample_seed=42, seed=1, eta=0.05 if frac < 1 else 1e-6)
    eng = ShardedLogisticEngine(X, y, cfg, row_base=500, dist=dist, backend="fused", use_tc=True, grid=grid)
    assert eng.use_tc
    w0 = (torch.randn(D, device=dev) * 0.05).contiguous()
    eng.w.copy_(w0)
    out = {}
    for it in (0, 2):
        eng._launch(_native.FIN_GRAD_ONLY, eng.w, it)
        torch.cuda.synchronize()
        keep = philox_row_mask_torch(n, frac, 42, it, 500, device=dev) if frac < 1 else torch.ones(n, dtype=torch.bool, device=dev)
        Xf = X.float()
        r = (1.0 / (torch.exp(-(Xf @ w0)) + 1.0) - y) * keep.float()
        g_ref = r @ Xf
        out[f"cnt_ok_{it}"] = int(eng.last_count.item()) == int(keep.sum().item())
        out[f"grad_err_{it}"] = float((eng.g - g_ref).abs().max().item() / (g_ref.abs().max().item() + 1e-12))
    ref = ShardedLogisticEngine(X, y, cfg, row_base=500, dist=dist, backend="fused", use_tc=False)
    ref.w.copy_(w0)
    eng.w.copy_(w0)
    eng.t = ref.t = 0
    for _ in range(6):
        eng.step()
        ref.step()
    torch.cuda.synchronize()
    out["step_err"] = float((eng.w - ref.w).abs().max().item() / (ref.w.abs().max().item() + 1e-12))
    return out


def main():
    kind = sys.argv[1]
    args = list(map(int, sys.argv[2:]))
    out = {"kmeans": kmeans, "kmeans_ties": kmeans_ties, "gram": gram, "lr": lr}[kind](*args)
    print("RESULT " + json.dumps(out))


if __name__ == "__main__":
    main()
