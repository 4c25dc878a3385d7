"""CPU-only: the order-preserving integer keys that carry fp32 / fp64 min and max through the switch in the posted reduce
(`PostedKey32::code`, `PostedKey64::code` in src/coll/coll_posted.cuh), restated with numpy on the same bit patterns.

k = b ^ ((b >> 31) & 0x7fffffff)   (arithmetic shift; 63 / 0x7fff...f for fp64)

Checked: the map is its own inverse; signed-integer order of the keys equals IEEE order of the values for everything that is not
a NaN (infinities, denormals, the largest and smallest normals included); -0 orders directly below +0; NaNs order beyond the
infinity of their sign, so a min never returns a positive NaN unless every operand is one; and an integer min / max over keys,
decoded, equals numpy's min / max on NaN-free data — which is what the device path computes with multimem.ld_reduce.min.s32/.s64."""
import numpy as np
import pytest


def key(bits: np.ndarray) -> np.ndarray:
    """bits: int32 or int64 view of the floats."""
    width = bits.dtype.itemsize * 8
    mask = np.array((1 << (width - 1)) - 1, dtype=bits.dtype)
    return bits ^ ((bits >> (width - 1)) & mask)


def specials(ft):
    fi = np.finfo(ft)
    vals = [0.0, -0.0, fi.tiny, -fi.tiny, fi.tiny / 8, -fi.tiny / 8, fi.smallest_subnormal, -fi.smallest_subnormal, 1.0, -1.0, 1.0 + fi.eps, -1.0 - fi.eps, fi.max, fi.min,
            np.inf, -np.inf, 3.5, -3.5, 1e-30, -1e-30]
    return np.array(vals, dtype=ft)


@pytest.mark.parametrize("ft,it", [(np.float32, np.int32), (np.float64, np.int64)])
def test_keys_are_an_order_preserving_involution(ft, it):
    rng = np.random.default_rng(7)
    rand_bits = rng.integers(np.iinfo(it).min, np.iinfo(it).max, size=20000, dtype=it)
    x = np.concatenate([specials(ft), rand_bits.view(ft)])
    x = x[~np.isnan(x)]
    b = x.view(it)
    k = key(b)
    assert np.array_equal(key(k), b)  # its own inverse
    order = np.argsort(k, kind="stable")
    xs = x[order]
    assert np.all(xs[:-1] <= xs[1:])  # integer order of the keys == IEEE order of the values
    # strictness: different values never share a key, and -0 sits directly below +0
    assert len(np.unique(k)) == len(np.unique(b))
    kz = key(np.array([-0.0, 0.0], dtype=ft).view(it))
    assert kz[0] == -1 and kz[1] == 0


@pytest.mark.parametrize("ft,it", [(np.float32, np.int32), (np.float64, np.int64)])
def test_nans_order_beyond_the_infinities(ft, it):
    nan_pos = np.array([np.nan], dtype=ft)
    nan_neg = -nan_pos
    inf = np.array([np.inf, -np.inf], dtype=ft)
    k_inf = key(inf.view(it))
    assert key(nan_pos.view(it))[0] > k_inf[0]
    assert key(nan_neg.view(it))[0] < k_inf[1]


@pytest.mark.parametrize("ft,it", [(np.float32, np.int32), (np.float64, np.int64)])
@pytest.mark.parametrize("P", [2, 4, 8])
def test_integer_min_max_over_keys_is_the_float_min_max(ft, it, P):
    rng = np.random.default_rng(100 + P)
    n = 4096
    ranks = [(rng.standard_normal(n) * 10.0 ** rng.integers(-30, 30, size=n)).astype(ft) for _ in range(P)]
    for r in range(P):  # sprinkle specials over every rank at different places
        s = specials(ft)
        ranks[r][r::P][: len(s)] = np.roll(s, r)[: len(ranks[r][r::P])]
    keys = np.stack([key(x.view(it)) for x in ranks])
    stacked = np.stack(ranks)
    for red_k, red_f in ((np.min, np.min), (np.max, np.max)):
        got = key(red_k(keys, axis=0)).view(ft)
        want = red_f(stacked, axis=0)
        # equal as values everywhere; equal as bits except where the answer is a zero (-0 < +0 for keys, numpy picks the first)
        assert np.array_equal(got, want)
        nz = want != 0
        assert np.array_equal(got.view(it)[nz], want.view(it)[nz])
