"""Host-side checks of arithmetic identities the CUDA shading code relies on (no GPU)."""
import numpy as np


def fdiv_shortcut(a, b):
    """pt_math.cuh:fdiv in numpy fp32: a zero numerator is replaced by 2^-40 and the quotient multiplied by a."""
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    zero = a == 0
    with np.errstate(all="ignore"):
        q = np.where(zero, np.float32(2.0 ** -40), a) / b
        z = a * q
    return np.where(zero, z, q).astype(np.float32)


def test_division_shortcut_algorithm_is_ieee_division():
    """The identity behind pt_math.cuh:fdiv, in IEEE fp32 on the host: for a == +-0 the product a * (2^-40 / b) is the
    signed zero 0 / b gives for every finite non-zero b (2^-40 / b cannot overflow: b >= 2^-149) and for b = inf, and
    NaN for b = 0 and b = NaN.  tests/test_pt_gpu.py::test_division_shortcut_is_ieee_division runs the device code."""
    rng = np.random.RandomState(1)
    special = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 1e-39, 3e38, -3e38,
                        1.17549435e-38, 2.0 ** -40, 2.0 ** -100, 2.0 ** 100, 0.1, -7.5], np.float32)
    rnd = lambda n: rng.randint(0, 2 ** 32, n, dtype=np.uint64).astype(np.uint32).view(np.float32)
    a = np.concatenate([np.repeat(special, special.size), rnd(1 << 20), np.zeros(1 << 16, np.float32),
                        -np.zeros(1 << 16, np.float32)])
    b = np.concatenate([np.tile(special, special.size), rnd(1 << 20), rnd(1 << 17)])
    with np.errstate(all="ignore"):
        want = (a / b).astype(np.float32)
    got = fdiv_shortcut(a, b)
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    assert np.array_equal(got[~nan].view(np.uint32), want[~nan].view(np.uint32))


def test_quantised_plane_conversion_is_exact_and_biased_outwards():
    """traverse.cuh:quant_byte_to_float: byte b dropped into mantissa bits [15:8] of 1.0f is exactly 1 + b * 2^-15, and
    with the HIGH-plane constant (1.0f + 2 ulp) exactly 2^-7 quantisation step more."""
    b = np.arange(256, dtype=np.uint32)
    lo = (np.uint32(0x3F800000) | (b << 8)).view(np.float32).astype(np.float64)
    hi = (np.uint32(0x3F800002) | (b << 8)).view(np.float32).astype(np.float64)
    assert np.array_equal(lo, 1.0 + b * 2.0 ** -15)
    assert np.array_equal((hi - lo) * 2.0 ** 15, np.full(256, 2.0 ** -7))


def test_division_through_a_double_reciprocal_is_ieee_division():
    """pt_math.cuh:fdiv (PT_FDIV_MODE 2) as an algorithm, on the host: (float)((double)a * r) with r within 1.5 ulp
    (double) of 1 / b is the correctly rounded fp32 quotient for every finite non-zero b -- a quotient of two 24-bit
    numbers cannot lie within 2^-49 (relatively) of a float midpoint, the double result is within 2^-51 of it.  The
    reciprocal is perturbed randomly inside its error bound, as the device's Newton iteration may land anywhere in
    it.  The device code itself: tests/test_pt_gpu.py::test_division_shortcut_is_ieee_division."""
    rng = np.random.RandomState(3)
    n = 1 << 22
    rnd = lambda k: rng.randint(0, 2 ** 32, k, dtype=np.uint64).astype(np.uint32).view(np.float32)
    unit = lambda k: (rng.rand(k).astype(np.float32) * np.float32(4.0) - np.float32(2.0))
    special = np.array([0.0, -0.0, 1.0, -1.0, np.inf, -np.inf, np.nan, 1e-45, -1e-45, 1e-39, 3e38, -3e38,
                        1.17549435e-38, 2.0 ** -40, 2.0 ** -100, 2.0 ** 100, 0.1, -7.5, 3.0, 1.0 / 3.0], np.float32)
    a = np.concatenate([np.repeat(special, special.size), rnd(n), unit(n), unit(n), np.zeros(1 << 12, np.float32)])
    b = np.concatenate([np.tile(special, special.size), rnd(n), unit(n), rnd(n), unit(1 << 12)])
    ok = np.isfinite(b) & (b != 0)          # the device takes the plain division for b = 0, inf, NaN
    a, b = a[ok], b[ok]
    with np.errstate(all="ignore"):
        want = (a / b).astype(np.float32)
        r = 1.0 / b.astype(np.float64)
        r = r * (1.0 + (rng.rand(r.size) - 0.5) * 3.0 * 2.0 ** -53)
        got = (a.astype(np.float64) * r).astype(np.float32)
    nan = np.isnan(want)
    assert np.array_equal(np.isnan(got), nan)
    assert np.array_equal(got[~nan].view(np.uint32), want[~nan].view(np.uint32))
