"""Per-role wait-time breakdown of the k-means tensor-core kernel (CTA 0), plus event timing."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributed_algorithm_pyspark_b200 import _native
lib = _native.load()
n, d, k = int(sys.argv[1]) if len(sys.argv) > 1 else 16_000_000, 128, 1024
dev = "cuda"
X = torch.randn(n, d, device=dev).to(torch.bfloat16)
C = torch.randn(k, d, device=dev)
B = torch.empty(k, d, dtype=torch.bfloat16, device=dev); cn = torch.empty(k + 8, device=dev)
sums = torch.zeros(k, d, device=dev); counts = torch.zeros(k, dtype=torch.int32, device=dev)
dbg = torch.zeros(64, dtype=torch.int64, device=dev)
st = _native.stream_ptr()
_native.check(lib.dapb_kmeans_tc_prepare(C.data_ptr(), k, d, B.data_ptr(), cn.data_ptr(),
                                         float((X[:1 << 20].float() ** 2).sum(1).max().item()) * 1.5, st))
for rep in range(3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    _native.check(lib.dapb_kmeans_assign_tc_dbg(X.data_ptr(), n, d, B.data_ptr(), cn.data_ptr(), k, 0, sums.data_ptr(),
                                                counts.data_ptr(), 0, dbg.data_ptr(), st))
    b.record(); torch.cuda.synchronize()
ms = a.elapsed_time(b)
tiles_per_cta = n / 128 / 148
print(f"n={n}: {ms:.3f} ms, {n/ms/1e6:.2f} G rows/s, {2*n*k*d/ms/1e9:.1f} TFLOP/s")
v = dbg.cpu().view(8, 8)
names = {0: ("X producer", ["wait a_empty"]), 1: ("MMA issuer", ["wait b_full", "wait a_full", "wait t_empty"]),
         3: ("C producer", ["wait b_empty"]), 4: ("arg-min warp", ["wait t_full", "wait tmem ld"]), 5: ("scatter warp", ["wait as_full"])}
for r, (nm, labels) in names.items():
    tot = int(v[r, 4])
    print(f"{nm:14s} total {tot/1e6:8.2f} Mcyc ({tot/tiles_per_cta:8.0f} cyc/tile): " +
          ", ".join(f"{l} {100*int(v[r,i])/max(tot,1):5.1f}%" for i, l in enumerate(labels)) +
          (f", mma issue {100*int(v[r,3])/max(tot,1):5.1f}%, commits {100*int(v[r,5])/max(tot,1):5.1f}%" if r == 1 else ""))
