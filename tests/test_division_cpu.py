"""CPU check of the arithmetic fact behind `div_wide` (csrc/grid.cuh): dividing two binary32 numbers in binary64
and rounding the quotient to binary32 gives the correctly rounded binary32 quotient -- double rounding is
innocuous when the wide format has at least 2*24+2 bits -- including subnormal results.  The GPU kernels use it
as their exact fallback for dividends outside the verified window (tests/test_gpu_parity.py checks those bitwise
against the oracle); here numpy's float32 division is the judge."""
import numpy as np


def test_division_through_double_is_correctly_rounded():
    rng = np.random.default_rng(20261107)
    n = 2_000_000
    anywhere = rng.integers(0, 0x7F800000, n, dtype=np.uint32).view(np.float32)      # every finite magnitude
    tiny = rng.integers(0, 0x04000000, n, dtype=np.uint32).view(np.float32)          # subnormals and their neighbours
    for c in (np.float32(4.0), np.float32(2685.3542), np.float32(1.0004), np.float32(42.96), np.float32(3.0), np.float32(1e-3)):
        for t in (anywhere, tiny, -tiny):
            with np.errstate(all="ignore"):
                q32 = (t / c).astype(np.float32)
                q64 = (t.astype(np.float64) / np.float64(c)).astype(np.float32)
            assert np.array_equal(q32.view(np.uint32), q64.view(np.uint32)), float(c)


def test_division_by_a_power_of_two_is_a_multiplication():
    """The pressure solve divides by 4: t / 4 and t * 0.25 are the same real number rounded once, subnormals included."""
    rng = np.random.default_rng(4)
    t = rng.integers(0, 0x7F800000, 2_000_000, dtype=np.uint32).view(np.float32)
    small = rng.integers(0, 0x02000000, 2_000_000, dtype=np.uint32).view(np.float32)
    for x in (t, -t, small, -small):
        for c, y in ((np.float32(4.0), np.float32(0.25)), (np.float32(2.0), np.float32(0.5)), (np.float32(64.0), np.float32(1 / 64))):
            with np.errstate(all="ignore"):
                assert np.array_equal((x / c).view(np.uint32), (x * y).view(np.uint32))
    one = np.float32(1.0)
    assert np.array_equal((one * t).view(np.uint32), t.view(np.uint32))  # a = 1: a * s is s


def test_bracketed_reciprocal_product_decides_the_correctly_rounded_quotient():
    """div_bracket (csrc/grid.cuh): q = RN64(t * RN64(1/c)) is within 2^-52 of t/c, so t/c lies strictly between
    q*(1-2^-50) and q*(1+2^-50); rounding to single is monotone, so when both ends round to the same float that float
    is RN32(t/c) -- for subnormal quotients too.  Checked here on the CPU over the operands it is used for (the
    subnormal fringe of a field); undecided cases (a rounding boundary or an exact tie near q) fall back to division."""
    rng = np.random.default_rng(0)
    for c in (np.float32(42.94), np.float32(3.0), np.float32(1.6553601), np.float32(4.0), np.float32(1e-3), np.float32(1.0000001)):
        t = rng.integers(0, 0x00ffffff * 4, 1_000_000, dtype=np.uint32).view(np.float32)
        t = np.concatenate([t, -t, (rng.standard_normal(200_000) * 1e-30).astype(np.float32), (rng.standard_normal(200_000)).astype(np.float32),
                            np.array([0.0, -0.0, 3 * 2.0 ** -149, 2.0 ** -149, 2.0 ** -126, np.inf, -np.inf], np.float32)])
        with np.errstate(all="ignore"):
            want = (t / c).astype(np.float32)
            q = t.astype(np.float64) * (np.float64(1.0) / np.float64(c))
            lo = (q * (1 - 2.0 ** -50)).astype(np.float32)
            hi = (q * (1 + 2.0 ** -50)).astype(np.float32)
        decided = lo.view(np.uint32) == hi.view(np.uint32)
        assert decided.mean() > 0.8
        assert np.array_equal(lo.view(np.uint32)[decided], want.view(np.uint32)[decided]), float(c)


def test_scaled_short_division_is_exact_on_the_subnormal_fringe(tmp_path):
    """div_scaled (csrc/grid.cuh), restated operation for operation in tests/helpers/div_scaled_check.c: the short
    reciprocal sequence on the dividend scaled by 2^64, the quotient scaled back, ties of subnormal quotients settled
    by the sign of the exact residual.  Every dividend whose quotient can be subnormal (all bit patterns below 2^-118,
    both signs, zeros included) and a stride through the rest of [0, 2^64), against the machine's IEEE division.
    (The GPU proves the same per divisor over all of [0, 2^64) before it uses the sequence: k_verify_division_full.)"""
    import os
    import subprocess
    src = os.path.join(os.path.dirname(os.path.abspath(__file__)), "helpers", "div_scaled_check.c")
    exe = str(tmp_path / "div_scaled_check")
    flags = ["-mfma"] if "fma" in open("/proc/cpuinfo").read() else []
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-ffp-contract=off"] + flags + ["-o", exe, src, "-lm"])
    divisors = ["3.0", "5.0", "11.4857597", "1.0004", "108.374184", "2685.3542", "1.0000001", "1000000"]
    for first, last, stride in ((0, 0x04000000, 1), (0, 0x5F800000, 997)):
        out = subprocess.check_output([exe, str(first), str(last), str(stride)] + (divisors if stride > 1 else divisors[:3]), text=True)
        for line in out.strip().splitlines():
            c, tried, bad, adjusted = line.split()
            assert int(bad) == 0, line
            assert int(tried) > 1_000_000
            if stride == 1:
                assert int(adjusted) > 0, "the tie rule was never exercised: " + line


def test_short_division_commutes_with_power_of_two_scaling(tmp_path):
    """The quick proof (k_verify_division_quick, csrc/grid.cuh) tries ONE binade of dividends per divisor instead of all
    of them, on the strength of this: from 2^(ec-100) up, every step of div_seq commutes with a scaling of the dividend by
    a power of two.  Checked here for whole binades at the lower end of that range, in the middle and at the top: the
    quotient's significand is the same in every binade, its exponent moves with the dividend's, and it is t / c."""
    import os
    import subprocess
    src = os.path.join(os.path.dirname(os.path.abspath(__file__)), "helpers", "div_scaled_check.c")
    exe = str(tmp_path / "div_scaled_check")
    flags = ["-mfma"] if "fma" in open("/proc/cpuinfo").read() else []
    subprocess.check_call(["gcc", "-O2", "-std=c11", "-ffp-contract=off"] + flags + ["-o", exe, src, "-lm"])
    out = subprocess.check_output([exe, "seq", "1.0004", "3.0", "11.4857597", "42.96", "2685.3542", "1.9999999", "30000.5"], text=True)
    lines = out.strip().splitlines()
    assert len(lines) == 7
    for line in lines:
        c, ec, bad, moved = line.split()
        assert int(bad) == 0 and int(moved) == 0, line
