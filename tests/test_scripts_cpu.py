"""The ten entry points keep the reference's stdout format."""
import os
import re
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(script, *extra, tmp):
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="")
    out = subprocess.run([sys.executable, os.path.join(ROOT, script), "--seed", "0", "--out-dir", str(tmp), *extra],
                         capture_output=True, text=True, timeout=600, env=env)
    assert out.returncode == 0, out.stderr[-2000:]
    return out.stdout


@pytest.mark.parametrize("script,png", [
    ("optimization/ssgd.py", "ssgd_acc_plot"), ("optimization/ma.py", "ma_acc_plot"),
    ("optimization/bmuf.py", "bmuf_acc_plot"), ("optimization/easgd.py", "easgd_acc_plot"),
    ("machine_learning/logistic_regression.py", "logistic_regression_acc_plot")])
@pytest.mark.parametrize("backend", ["torch", "oracle"])
def test_logistic_scripts_stdout(tmp_path, script, png, backend):
    out = _run(script, "--backend", backend, "--n-iterations", "3", tmp=tmp_path)
    lines = out.strip().splitlines()
    assert lines[0].startswith("Initial w: [")
    assert "On iteration 1" in lines and "On iteration 3" in lines
    assert any(re.fullmatch(r"iterations: 2, accuracy: 0\.\d{6}", l) for l in lines)
    assert any(l.startswith("Final w: [") for l in lines)
    assert re.fullmatch(r"Final acc: \d\.\d{6}", lines[-1])
    assert (tmp_path / (png + ".npy")).exists()


def test_other_scripts_stdout(tmp_path):
    assert "Final centers: [array([10.,  2.]), array([1., 2.])]" in _run("machine_learning/k-means.py", "--backend", "torch", tmp=tmp_path)
    out = _run("graph_computation/pagerank.py", "--backend", "torch", tmp=tmp_path)
    assert out.strip().splitlines() == ["1 has rank: 0.38891305880091237.", "2 has rank: 0.214416470596171.",
                                        "3 has rank: 0.3966704706029163."]
    assert _run("graph_computation/transitive_closure.py", "--backend", "torch", tmp=tmp_path).strip() == "The original graph has 9 paths"
    out = _run("matrix_computation/matrix_decomposition.py", "--backend", "torch", tmp=tmp_path).strip().splitlines()
    assert len(out) == 5 and out[0] == "iterations: 0, rmse: 0.224488" and out[4].startswith("iterations: 4, rmse: 0.0314")
    assert re.fullmatch(r"Pi is roughly 3\.1\d{5}", _run("randomized_algorithm/monte_carlo.py", "--backend", "torch", tmp=tmp_path).strip())
    assert (tmp_path / "matrix_decomposition_result.npz").exists() and (tmp_path / "kmeans_clusters_display.npy").exists()


def test_bench_reference_arm_contract():
    """`bench.py --impl reference` must print one JSON line with impl = reference and exit 0 (the reference cannot be
    installed offline: ten PySpark scripts, no pyspark / JVM in the image)."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "3",
                          "--warmup", "3"], capture_output=True, text=True, timeout=120)
    assert out.returncode == 0, out.stderr[-500:]
    line = json.loads(out.stdout.strip().splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and "unavailable" in line
