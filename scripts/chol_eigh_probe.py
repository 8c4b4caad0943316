"""Times dapb_cholesky / dapb_eigh alone (CUDA events, 20 repetitions) for a 512 x 512 fp32 Gram matrix."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_algorithm_pyspark_b200 import _native
lib = _native.load()
dev = "cuda"
n = 512
A = torch.randn(4096, n, device=dev)
G = (A.T @ A).contiguous()
info = torch.zeros(1, dtype=torch.int32, device=dev)
Ls = [G.clone() for _ in range(22)]
def chol(L):
    _native.check(lib.dapb_cholesky(L.data_ptr(), n, 0, info.data_ptr(), _native.stream_ptr()))
chol(Ls[0]); chol(Ls[1]); torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for i in range(20): chol(Ls[2 + i])
b.record(); torch.cuda.synchronize()
out = {"chol_ms": a.elapsed_time(b) / 20, "grid_env": os.environ.get("DAPB_CHOL_GRID")}
dbg = torch.zeros(8, dtype=torch.int64, device=dev)
Ld = G.clone()
_native.check(lib.dapb_cholesky_dbg(Ld.data_ptr(), n, 0, info.data_ptr(), dbg.data_ptr(), _native.stream_ptr()))
torch.cuda.synchronize()
out["chol_cycles_factor_solve_sync_update_sync"] = dbg[:5].tolist()
ref = torch.linalg.cholesky(G.double())
out["chol_err"] = float((Ls[5].double() - ref).abs().max() / ref.abs().max())
work = torch.empty(2 * n * n + 8, device=dev); ev = torch.empty(n, device=dev); evec = torch.empty(n, n, device=dev)
sw = torch.zeros(1, dtype=torch.int32, device=dev)
G64 = G.double().contiguous()
work = torch.empty(2 * n * n + 8, device=dev, dtype=torch.float64); ev = torch.empty(n, device=dev, dtype=torch.float64); evec = torch.empty(n, n, device=dev, dtype=torch.float64)
def eigh():
    _native.check(lib.dapb_eigh(G64.data_ptr(), n, 1, work.data_ptr(), ev.data_ptr(), evec.data_ptr(), 0.0, 0, sw.data_ptr(), _native.stream_ptr()))
eigh(); torch.cuda.synchronize()
a.record(); eigh(); b.record(); torch.cuda.synchronize()
wr = torch.linalg.eigvalsh(G.double())
out.update({"eigh_ms": a.elapsed_time(b), "eigh_b_env": os.environ.get("DAPB_EIGH_B"), "eigh_sweeps": int(sw.item()), "eigh_err": float((ev.double() - wr).abs().max() / wr.abs().max())})
a.record(); torch.linalg.eigh(G); b.record(); torch.cuda.synchronize()
a.record(); torch.linalg.eigh(G); b.record(); torch.cuda.synchronize()
out["torch_eigh_fp32_ms"] = a.elapsed_time(b)
a.record(); torch.linalg.cholesky(G); b.record(); torch.cuda.synchronize()
a.record(); torch.linalg.cholesky(G); b.record(); torch.cuda.synchronize()
out["torch_cholesky_fp32_ms"] = a.elapsed_time(b)
print(json.dumps(out))
