import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_gpu = torch.cuda.is_available()
        n_gpu = torch.cuda.device_count() if has_gpu else 0
    except Exception:
        has_gpu, n_gpu = False, 0
    skip_gpu = pytest.mark.skip(reason="no CUDA device")
    skip_multi = pytest.mark.skip(reason="needs >= 2 CUDA devices")
    for item in items:
        if "gpu" in item.keywords and not has_gpu:
            item.add_marker(skip_gpu)
        if "multigpu" in item.keywords and n_gpu < 2:
            item.add_marker(skip_multi)


@pytest.fixture(scope="session")
def dist():
    from distributed_algorithm_pyspark_b200.parallel.dist import init_distributed
    return init_distributed()


@pytest.fixture(scope="session")
def breast_cancer():
    from distributed_algorithm_pyspark_b200.data.datasets import load_breast_cancer_split
    return load_breast_cancer_split()
