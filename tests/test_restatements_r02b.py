"""CPU: two pieces of arithmetic that device code of round 2 relies on, restated in numpy.

1. The factored flux operands of maxwell_stage_class_mma (yfd_maxwell_mma.cuh): with z = ny Hx - nx Hy, q1 = cE Ez - vH z,
   q2 = cH z - vE Ez, the three per-field face operands of maxwell_stage_class (methods/dg_maxwell_2D.hpp:356-410 as the
   streamed kernel states them) are  ny q1,  -nx q1,  q2 -- so one block product serves both H fields.
2. The wire format of the cross-rank sums (yfd_cg_shared.cuh: xrank_sum): a double travels as two 8-byte packets
   { 32 data bits | 32-bit epoch tag }; decoding returns the same bits, tags of epochs e and e + 2 (same buffer) always differ
   and are never the zero of a fresh region."""
import numpy as np


def test_factored_flux_operands_equal_the_three_per_field_combinations():
    rng = np.random.default_rng(0)
    for _ in range(200):
        nx, ny = rng.standard_normal(2)
        nrm = np.hypot(nx, ny); nx, ny = nx / nrm, ny / nrm
        cE, cH, vH, vE = rng.uniform(0.0, 2.0, 4)
        Hx, Hy, Ez = rng.standard_normal((3, 15))
        # maxwell_stage_class: w = ahx Hx + ahy Hy + aez Ez per output field
        w_hx = (ny * cE) * Ez + (-vH * ny * ny) * Hx + (vH * nx * ny) * Hy
        w_hy = (-nx * cE) * Ez + (vH * nx * ny) * Hx + (-vH * nx * nx) * Hy
        w_ez = (cH * ny) * Hx + (-cH * nx) * Hy + (-vE) * Ez
        z = ny * Hx - nx * Hy
        q1 = cE * Ez - vH * z
        q2 = cH * z - vE * Ez
        scale = max(np.abs(w_hx).max(), np.abs(w_hy).max(), np.abs(w_ez).max(), 1.0)
        assert np.abs(ny * q1 - w_hx).max() <= 8e-16 * scale
        assert np.abs(-nx * q1 - w_hy).max() <= 8e-16 * scale
        assert np.abs(q2 - w_ez).max() <= 8e-16 * scale
        # a linear map applied to q1 and scaled afterwards = the map applied to the scaled operand (what the kernel exploits)
        B = rng.standard_normal((15, 15))
        assert np.abs(ny * (B @ q1) - B @ w_hx).max() <= 1e-13 * scale * np.abs(B).sum(axis=1).max()


def _tag(e):
    return ((int(e) & 0x7fffffff) | 0x80000000) << 32


def _encode(v, e):
    bits = int(np.float64(v).view(np.uint64))
    return (bits & 0xffffffff) | _tag(e), (bits >> 32) | _tag(e)


def _decode(lo, hi, e):
    mask = 0xffffffff00000000
    if (lo & mask) != _tag(e) or (hi & mask) != _tag(e):
        return None
    return np.uint64(((lo & 0xffffffff) | (hi << 32)) & 0xffffffffffffffff).view(np.float64)


def test_tagged_packets_carry_a_double_and_its_epoch():
    rng = np.random.default_rng(1)
    vals = list(rng.standard_normal(50) * 10.0 ** rng.integers(-300, 300, 50)) + [0.0, -0.0, np.inf, -np.inf, 5e-324, np.nan]
    for e in (1, 2, 3, 0x7fffffff, 0x80000000, 0x80000001, 0xffffffff, 2 ** 40 + 5):
        assert _tag(e) != 0 and _tag(e) != _tag(e + 2)
        for v in vals:
            lo, hi = _encode(v, e)
            got = _decode(lo, hi, e)
            assert got is not None and np.float64(v).view(np.uint64) == np.float64(got).view(np.uint64)
            # a stale packet (previous epoch of the same buffer) or a half-arrived pair is not accepted
            assert _decode(lo, hi, e + 2) is None
            lo_old, hi_old = _encode(v * 2 if np.isfinite(v) else 1.0, e - 2 if e > 2 else e + 4)
            assert _decode(lo, hi_old, e) is None and _decode(lo_old, hi, e) is None
    # a fresh (zeroed) region never looks like a packet
    assert _decode(0, 0, 1) is None
