import numpy as np
import torch

from distributed_algorithm_pyspark_b200.utils.philox import (bernoulli_threshold, philox4x32_10_np,
                                                              philox_row_mask_np, philox_row_mask_torch)


def test_philox_known_answer():
    # Random123 known-answer test: counter = key = 0
    out = philox4x32_10_np(np.array([0]), 0, 0, 0, 0, 0)
    assert [int(v[0]) for v in out] == [0x6627E8D5, 0xE169C58D, 0xBC57AC4C, 0x9B00DBD8]
    out = philox4x32_10_np(np.array([0xFFFFFFFF]), 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF, 0xFFFFFFFF)
    assert [int(v[0]) for v in out] == [0x408F276D, 0x41C83B0E, 0xA20BC7C6, 0x6D5451FD]


def test_numpy_and_torch_masks_agree():
    for seed, it, off in [(42, 0, 0), (42, 7, 12345), ((1 << 40) + 3, 99, (1 << 33) + 5)]:
        a = philox_row_mask_np(4096, 0.1, seed, it, off)
        b = philox_row_mask_torch(4096, 0.1, seed, it, off).numpy()
        assert (a == b).all()
        assert 300 < a.sum() < 520


def test_threshold_edges():
    assert bernoulli_threshold(0.0) == 0 and bernoulli_threshold(1.0) == 1 << 24
    assert philox_row_mask_np(100, 1.0, 1, 0).all() and not philox_row_mask_np(100, 0.0, 1, 0).any()
