"""Exactness claims the CUDA kernels rely on (csrc/intersect.cuh, csrc/vecmath.h), proved on the CPU."""
import numpy as np


def _neighbours(x32):
    x = np.float32(x32)
    return [np.nextafter(x, np.float32(-np.inf), dtype=np.float32), x, np.nextafter(x, np.float32(np.inf), dtype=np.float32)]


def test_double_literal_compares_equal_float_compares():
    """(double)x < 0.001 <=> x < 0.001f, (double)x <= -0.01 <=> x < -0.01f, (double)x > 0.00001 <=> x > 0.00001f
    for EVERY float x: it is enough to check the floats around the literal."""
    for lit, dbl_op, flt_op in ((0.001, lambda x: float(x) < 0.001, lambda x: x < np.float32(0.001)),
                                (-0.01, lambda x: float(x) <= -0.01, lambda x: x < np.float32(-0.01)),
                                (0.00001, lambda x: float(x) > 0.00001, lambda x: x > np.float32(0.00001))):
        centre = np.float32(lit)
        cands = []
        for c in _neighbours(centre):
            cands += _neighbours(c)
        for x in cands:
            assert dbl_op(x) == bool(flt_op(x)), (lit, x)
        # monotone predicates: agreement around the threshold implies agreement everywhere
        assert float(cands[0]) < lit < float(cands[-1])


def test_fp64_reciprocal_rounded_equals_fp32_reciprocal():
    """(float)(1.0/(double)n) == correctly rounded 1.0f/n: structs.h:82-84's FP64 divide can be a
    single-precision __frcp_rn on the device."""
    rng = np.random.default_rng(5)
    bits = rng.integers(0x00800000, 0x7F000000, size=4_000_000, dtype=np.uint32)
    n = bits.view(np.float32)
    a = (1.0 / n.astype(np.float64)).astype(np.float32)
    b = np.float32(1.0) / n
    assert (a.view(np.uint32) == b.view(np.uint32)).all()
    # exhaustive over one binade (the mantissa pattern is all that matters)
    m = (np.arange(1 << 23, dtype=np.uint32) | np.uint32(0x3F800000)).view(np.float32)
    a = (1.0 / m.astype(np.float64)).astype(np.float32)
    b = np.float32(1.0) / m
    assert (a.view(np.uint32) == b.view(np.uint32)).all()


def test_max_double_zero_equals_fmaxf():
    x = np.array([-1.5, -0.0, 0.0, 1e-30, 0.25, 3.0], dtype=np.float32)
    assert (np.maximum(0.0, x.astype(np.float64)).astype(np.float32) == np.maximum(np.float32(0), x)).all()


def test_pack_roundtrip_all_bytes():
    """v/255*255 truncates back to v for all 256 values when done as two FP32 multiplies (sky texels,
    reference kernel.cu:163, 256-257)."""
    v = np.arange(256, dtype=np.float32)
    back = (v * np.float32(1.0 / 255.0)) * np.float32(255.0)
    assert (np.clip(back, 0, 255).astype(np.int32) == np.arange(256)).all()


def test_constant_division_by_fma_correction_is_ieee_division():
    """rt_div_const (csrc/trace_common.cuh): q0 = a*rc, q = q0 + (a - q0*b)*rc equals RN(a/b) for the two
    divisors of the sky lookup (reference kernel.cu:156-157, PI = 3.141592).  Exact rational arithmetic
    stands in for the two FMAs."""
    from fractions import Fraction as F
    PI = 3.141592
    rng = np.random.default_rng(11)
    for b in (PI, float(np.float32(2.0)) * PI):
        rc = 1.0 / b
        vals = list(rng.uniform(0.0, 6.4, 4000)) + [float(np.float64(np.arcsin(np.float32(v))) + PI / 2.0)
                                                     for v in rng.uniform(-1, 1, 2000).astype(np.float32)]
        for a in vals:
            q0 = a * rc
            r = float(F(a) - F(q0) * F(b))            # fma(-q0, b, a): one rounding
            q = float(F(q0) + F(r) * F(rc))           # fma(r, rc, q0): one rounding
            assert q == a / b, (a, b)


def test_fxaa_division_by_twelve_by_fma_correction():
    """fxaa_core.cuh: (double)sum / 12.0 (reference kernel.cu:372) as q0 + (sum - 12 q0) rc with rc = RN(1/12),
    for float sums in the range of eight luminances plus doubled axis terms (0 .. 12) and beyond."""
    from fractions import Fraction as F
    rng = np.random.default_rng(3)
    rc = 1.0 / 12.0
    vals = np.concatenate([rng.uniform(0, 12.5, 20000), rng.uniform(0, 0.05, 2000), [0.0, 12.0, 1.0, 3.0, 1e-30]]).astype(np.float32)
    for a32 in vals:
        a = float(a32)
        q0 = a * rc
        r = float(F(a) - F(q0) * 12)
        q = float(F(q0) + F(r) * F(rc))
        assert q == a / 12.0, a


def _rn32(x):
    """exact rational -> nearest float32 (ties to even), as a float"""
    from fractions import Fraction as F
    import math
    if x == 0:
        return 0.0
    sgn = -1 if x < 0 else 1
    x = abs(x)
    e = math.floor(math.log2(float(x)))
    while F(2) ** e > x:
        e -= 1
    while F(2) ** (e + 1) <= x:
        e += 1
    scale = F(2) ** (23 - e)
    n = x * scale
    fl = n.numerator // n.denominator
    rem = n - fl
    if rem > F(1, 2) or (rem == F(1, 2) and fl % 2 == 1):
        fl += 1
    return sgn * float(F(fl) / scale)


def test_fxaa_division_by_twelve_in_fp32():
    """fxaa_core.cuh computes the reference's (float)((double)sum / 12.0) (kernel.cu:372) entirely in FP32:
    q = fma(fma(-q0, 12, sum), rc, q0), q0 = RN(sum rc), rc = RN32(1/12).  Both claims exactly: the FP32 sequence is the
    correctly rounded quotient, and rounding via double first gives the same float."""
    from fractions import Fraction as F
    rng = np.random.default_rng(7)
    rc = float(np.float32(1.0 / 12.0))
    assert rc == 0.083333335816860198974609375
    # the sums FXAA forms are sums of nine luminances k/255-like values; plus random floats over the range and beyond
    vals = np.concatenate([rng.uniform(0, 12.5, 12000), rng.uniform(0, 0.05, 1500), rng.integers(0, 3060, 1500) / 255.0,
                           [0.0, 12.0, 1.0, 3.0, 0.25, 11.999999, 2.8e-4]]).astype(np.float32)
    for a32 in vals:
        a = float(a32)
        q0 = _rn32(F(a) * F(rc))
        r = _rn32(F(a) - F(q0) * 12)
        assert F(r) == F(a) - F(q0) * 12, a                      # the remainder is exact
        q = _rn32(F(q0) + F(r) * F(rc))
        want = float(np.float32(np.float64(a32) / 12.0))           # the reference: divide in double, round to float
        assert q == want, a
        assert q == _rn32(F(a) / 12), a                            # = the correctly rounded FP32 quotient


def test_fxaa_pack_by_round_to_zero_add():
    """fxaa_core.cuh fx_pack: the blend (1 - b) c + b n of two bytes never leaves [0, 255.0001], so the reference's
    clamp + truncate (kernel.cu:26-32) is the low mantissa byte of v + 2^23 rounded towards zero.  Checked on the FP32
    values the blend actually produces for every byte pair and a sweep of weights."""
    c, n = np.meshgrid(np.arange(256, dtype=np.float32), np.arange(256, dtype=np.float32))
    for b in np.concatenate([np.linspace(0, 1, 41), [1e-7, 0.5, 0.99999994]]).astype(np.float32):
        omb = np.float32(1.0) - b
        # fma(omb, c, b * n): emulate in float64 (products of a float32 and a byte are exact there, the sum is not
        # rounded twice in any case that matters for a floor below 256: compare against the exact floor instead)
        v = (omb.astype(np.float64) * c + (b * n).astype(np.float32).astype(np.float64)).astype(np.float32)
        assert v.min() >= 0.0 and v.max() < 256.0
        want = np.floor(np.minimum(np.maximum(v, 0.0), 255.0)).astype(np.uint32)     # clamp + truncate
        # v + 2^23 towards zero: exact sum has at most 24 + 8 bits, truncation = floor for v >= 0
        got = (np.floor(v.astype(np.float64)) + 8388608.0).astype(np.float32).view(np.uint32) & 255
        assert (got == want).all(), b


def test_fmodf_one_equals_x_minus_trunc():
    x = np.concatenate([np.random.default_rng(5).uniform(-3, 3, 20000), [0.0, 1.0, 1.9999999, -0.5]]).astype(np.float32)
    assert (np.fmod(x, np.float32(1.0)) == x - np.trunc(x)).all()
