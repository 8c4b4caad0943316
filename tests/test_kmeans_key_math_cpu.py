"""The integer-key arg-min of the tcgen05 k-means kernel (csrc/kmeans_tc.cu), restated in NumPy float32 / uint32.

The kernel never compares floats: the accumulator is pre-loaded with ||c||^2 + C1 where C1 = 1.5 * 2^E puts every biased
distance into one binade, and the epilogue takes the minimum of key = bits * 32 + column (mod 2^32) per 32-column group,
tracking which group won.  These tests check, on the host, the three facts that make this correct:
the Cauchy-Schwarz bias really confines all values to one exponent, key order == (value, column) order, and the padding
centroids' norm is above every real distance."""
import numpy as np
import pytest


def bf16_round(a):
    """Round-to-nearest-even float32 -> bf16 -> float32."""
    u = np.asarray(a, dtype=np.float32).view(np.uint32).astype(np.uint64)
    u = (u + 0x7FFF + ((u >> 16) & 1)) & 0xFFFF0000
    return u.astype(np.uint32).view(np.float32)


def bias_from_bounds(cmax, xmax):
    """kmeans_tc_bias_kernel: M, e, C1 and the padding norm, in float32 arithmetic."""
    cmax, xmax = np.float32(cmax), np.float32(xmax)
    M = np.float32(1.01) * (cmax + np.float32(2.0) * np.sqrt(xmax * cmax, dtype=np.float32)) + np.float32(1e-30)
    _, e = np.frexp(M)                                  # M = f * 2^e, f in [0.5, 1)
    C1 = np.ldexp(np.float32(1.5), int(e) + 1)
    pad = C1 + np.ldexp(np.float32(0.9999), int(e))
    return M, int(e), np.float32(C1), np.float32(pad)


def kernel_argmin(biased):
    """Epilogue restated: per 32-column group min of (bits * 32 + column_in_group) mod 2^32; a later group only wins
    with a strictly smaller key."""
    n, k = biased.shape
    bits = biased.astype(np.float32).view(np.uint32).astype(np.uint64)
    best = np.full(n, 0xFFFFFFFF, dtype=np.uint64)
    grp = np.zeros(n, dtype=np.int64)
    for g in range(k // 32):
        key = (bits[:, g * 32:(g + 1) * 32] * 32 + np.arange(32, dtype=np.uint64)) & 0xFFFFFFFF
        hm = key.min(axis=1)
        win = hm < best
        best[win], grp[win] = hm[win], g
    return grp * 32 + (best & 31).astype(np.int64)


@pytest.mark.parametrize("seed,d,k,scale", [(0, 128, 256, 1.0), (1, 64, 128, 30.0), (2, 128, 1024, 1e-3), (3, 128, 384, 1e4)])
def test_biased_distances_share_one_exponent_and_keys_order_correctly(seed, d, k, scale):
    rs = np.random.RandomState(seed)
    n = 300
    C = bf16_round(rs.randn(k, d) * scale)
    X = bf16_round(C[rs.randint(0, k, n)] + rs.randn(n, d) * scale * 0.5)
    cn = (C.astype(np.float32) ** 2).sum(1, dtype=np.float32)
    xmax = float((X.astype(np.float64) ** 2).sum(1).max())
    M, e, C1, pad = bias_from_bounds(cn.max(), xmax)
    dist = cn[None, :].astype(np.float64) - 2.0 * X.astype(np.float64) @ C.astype(np.float64).T
    assert np.abs(dist).max() <= float(M) / 1.009                 # the bound holds with its slack
    biased = (np.float32(C1) + cn)[None, :] + (-2.0 * X.astype(np.float64) @ C.astype(np.float64).T).astype(np.float32)
    biased = biased.astype(np.float32)
    expo = (biased.view(np.uint32) >> 23) & 0xFF
    assert expo.min() == expo.max() == 127 + e + 1                # one binade: [2^(e+1), 2^(e+2))
    got = kernel_argmin(biased)
    # reference: smallest biased value, ties -> lowest index (np.argmin); biased values are what the kernel sees
    assert np.array_equal(got, biased.argmin(axis=1))
    # and it agrees with the exact distances up to the quantisation step 2^(e+1-23)
    step = 2.0 ** (e + 1 - 23)
    gap = dist[np.arange(n), got] - dist.min(axis=1)
    assert gap.max() <= 2 * step
    # padding centroids (norm `pad`, zero operand row) never win and stay inside the binade
    assert pad > biased.max() and ((np.array([pad], np.float32).view(np.uint32) >> 23) & 0xFF)[0] == 127 + e + 1


def test_padding_and_degenerate_inputs():
    for cmax, xmax in [(0.0, 0.0), (1e-20, 1e-20), (1.0, 0.0), (0.0, 5.0), (3e8, 7e9)]:
        M, e, C1, pad = bias_from_bounds(cmax, xmax)
        assert np.isfinite(C1) and C1 > 0 and M < 2.0 ** e and pad < 2.0 ** (e + 2) and pad > C1
