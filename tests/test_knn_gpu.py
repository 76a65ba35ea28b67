"""GPU parity of the brute-force Hamming kNN-2 (C ABI) against the CPU oracle: bit-exact
distances and indices, ties, degenerate sizes."""
import numpy as np
import pytest

from orb_slam2_annotate_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("m,n", [(1, 1), (7, 1), (1000, 20000), (5000, 777), (300000, 600), (80000, 900)])
def test_knn2_matches_oracle(orbfe, oracle, m, n):
    t = synth.descriptors(n, 1)
    q = synth.planted_descriptors(t, m, 2)
    mt = orbfe.ORBmatcher(0.7)
    idx, d1, d2 = mt.knn2(q, t)
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.7)
    assert np.array_equal(d1, o1) and np.array_equal(d2, o2) and np.array_equal(idx, oi)
    if m >= 1000 and n >= 600:
        assert (idx >= 0).sum() > 0


def test_knn2_ties_and_extremes(orbfe, oracle):
    t = np.zeros((64, 32), np.uint8)
    t[10:] = 0xFF
    q = np.zeros((4, 32), np.uint8)
    q[1] = 0xFF          # distance 256 to rows 0..9, 0 to rows 10..
    q[2, 0] = 1          # distance 1 to rows 0..9: lowest index must win, d2 == d1
    mt = orbfe.ORBmatcher(0.9)
    idx, d1, d2 = mt.knn2(q, t, th=50, ratio=0.9)
    oi, o1, o2 = oracle.knn2(q, t, 50, 0.9)
    assert np.array_equal(idx, oi) and np.array_equal(d1, o1) and np.array_equal(d2, o2)
    # empty train set
    idx, d1, d2 = mt.knn2(q, np.zeros((0, 32), np.uint8))
    assert (idx == -1).all() and (d1 == 256).all() and (d2 == 256).all()


def test_descriptor_distance_host(orbfe, oracle):
    rng = np.random.default_rng(0)
    for _ in range(200):
        a = rng.integers(0, 256, 32, dtype=np.uint8); b = rng.integers(0, 256, 32, dtype=np.uint8)
        assert orbfe.ORBmatcher.DescriptorDistance(a, b) == oracle.descriptor_distance(a, b)


def test_knn2_full_size_planted_property(orbfe):
    """The bench's kNN shape (131072 x 131072, one GPU's share of configs[4]) through a size-independent property:
    every query is a train row with k <= 12 flipped bits, so its best match is that row at distance exactly k, the
    runner-up is far (random 256-bit rows sit near 128), and the ratio test accepts it."""
    n = 131072
    rng = np.random.default_rng(11)
    t = rng.integers(0, 256, (n, 32), dtype=np.uint8)
    src = rng.permutation(n)
    q = t[src].copy()
    k = rng.integers(0, 13, n)
    for i in np.nonzero(k)[0]:
        bits = rng.choice(256, k[i], replace=False)
        np.bitwise_xor.at(q[i], bits >> 3, (1 << (bits & 7)).astype(np.uint8))
    idx, d1, d2 = orbfe.ORBmatcher(0.7).knn2(q, t, th=50, ratio=0.7)
    assert np.array_equal(d1, k.astype(d1.dtype))
    assert np.array_equal(idx, src.astype(idx.dtype))
    assert d2.min() > 60
