"""Wall time of the ten reference-compatible entry scripts with their default constants, per backend.
usage: python scripts/time_scripts.py oracle|torch|fused [out.json]   (oracle/torch run on the CPU, fused needs a GPU)"""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SCRIPTS = ["optimization/ssgd.py", "optimization/ma.py", "optimization/bmuf.py", "optimization/easgd.py",
           "machine_learning/logistic_regression.py", "machine_learning/k-means.py", "graph_computation/pagerank.py",
           "graph_computation/transitive_closure.py", "matrix_computation/matrix_decomposition.py",
           "randomized_algorithm/monte_carlo.py"]
backend = sys.argv[1] if len(sys.argv) > 1 else "oracle"
out = {}
for s in SCRIPTS:
    cmd = [sys.executable, os.path.join(ROOT, s), "--backend", backend, "--seed", "0", "--quiet", "--no-save"]
    best = None
    for rep in range(int(os.environ.get("REPS", 1))):
        t0 = time.perf_counter()
        r = subprocess.run(cmd, capture_output=True, text=True, timeout=1800)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    last = [l for l in r.stdout.splitlines() if l.strip()][-1][:100] if r.stdout.strip() else r.stderr[-200:]
    out[s] = {"wall_s": round(best, 3), "rc": r.returncode, "last_line": last}
    print(f"{s:48s} {backend:7s} {best:8.2f} s  rc={r.returncode}  {last}", flush=True)
if len(sys.argv) > 2:
    json.dump({"backend": backend, "scripts": out}, open(sys.argv[2], "w"), indent=1)
