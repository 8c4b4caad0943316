"""Per-role cycle counters of gram_tc_kernel (DAPB_GRAM_DBG=1): total MMA-loop cycles and cycles the MMA thread waited for data."""
import os, sys, json
os.environ.setdefault("DAPB_GRAM_DBG", "1")
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from distributed_algorithm_pyspark_b200.ops.gram_tc import GramTC
from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed
dist = init_distributed()
n, P = 1 << 24, 512
A = torch.empty(n, P, dtype=torch.bfloat16, device="cuda")
for r0 in range(0, n, 1 << 22):
    A[r0:r0 + (1 << 22)].normal_()
op = GramTC(A, dist)
op(); op(); torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record(); op(); b.record(); torch.cuda.synchronize()
d = op.tickets[512:512 + 4 * 148].view(-1, 4).cpu()
out = {"ms": a.elapsed_time(b), "env": {k: v for k, v in os.environ.items() if k.startswith("DAPB_GRAM")}}
for role in range(5):
    m = d[:, 0] == role
    rows = d[m]
    rows = rows[rows[:, 1] > 0]
    if rows.numel():
        out[f"role{role}"] = {"ctas": int(rows.shape[0]), "loop_kcycles_mean": float(rows[:, 1].float().mean()),
                              "loop_kcycles_max": int(rows[:, 1].max()), "wait_full_kcycles_mean": float(rows[:, 2].float().mean())}
print(json.dumps(out))
