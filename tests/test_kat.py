"""Component known-answer tests, following the reference's own unit tests (opus-fix/celt/tests/test_unit_cwrs32.c,
test_unit_mathops.c, test_unit_entropy.c, test_unit_laplace.c, test_unit_rotation.c): the leaf routines of the kernel
sources (host build, tests/emu/emu_kat.cpp) against the procedure of each unit test and -- where the compiled
reference exports the routine -- against the reference's own answer, value for value."""
import ctypes
import math
import os
import subprocess

import numpy as np
import pytest

from conftest import REF_SO, ROOT, needs_ref


@pytest.fixture(scope="module")
def kat():
    src = os.path.join(ROOT, "tests", "emu", "emu_kat.cpp")
    so = os.path.join(ROOT, "tests", "emu", "libemu_kat.so")
    csrc = os.path.join(ROOT, "concentus_b200", "csrc")
    newest = max([os.path.getmtime(src)] + [os.path.getmtime(os.path.join(csrc, f)) for f in os.listdir(csrc)])
    if not os.path.exists(so) or os.path.getmtime(so) < newest:
        subprocess.run(["g++", "-std=c++17", "-O2", "-fPIC", "-shared", "-Wall", "-Wno-unused-function",
                        "-Wno-unknown-pragmas", "-Wno-maybe-uninitialized", "-o", so, src], check=True)
    lib = ctypes.CDLL(so)
    lib.kat_pvq_v.restype = ctypes.c_uint
    lib.kat_icwrs.restype = ctypes.c_uint
    lib.kat_isqrt32.restype = ctypes.c_uint
    lib.kat_isqrt32.argtypes = [ctypes.c_uint]
    lib.kat_lcg_skip.restype = ctypes.c_uint
    lib.kat_lcg_skip.argtypes = [ctypes.c_uint, ctypes.c_uint]
    lib.kat_cwrsi.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_uint, ctypes.c_void_p]
    lib.kat_icwrs.argtypes = [ctypes.c_int, ctypes.c_void_p]
    lib.kat_alg_unquant.restype = ctypes.c_uint
    lib.kat_alg_unquant.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_uint, ctypes.c_int, ctypes.c_int,
                                    ctypes.c_int]
    return lib


def ref_lib():
    lib = ctypes.CDLL(REF_SO)
    lib.isqrt32.restype = ctypes.c_uint
    lib.isqrt32.argtypes = [ctypes.c_uint]
    lib.celt_sqrt.argtypes = [ctypes.c_int32]
    lib.celt_rcp.argtypes = [ctypes.c_int32]
    lib.celt_rsqrt_norm.argtypes = [ctypes.c_int32]
    lib.celt_rsqrt_norm.restype = ctypes.c_int16
    lib.celt_cos_norm.argtypes = [ctypes.c_int32]
    lib.celt_cos_norm.restype = ctypes.c_int16
    lib.frac_div32.argtypes = [ctypes.c_int32, ctypes.c_int32]
    lib.celt_lcg_rand.restype = ctypes.c_uint
    lib.celt_lcg_rand.argtypes = [ctypes.c_uint]
    return lib


# the (N, K) cells of the static 48 kHz mode that test_unit_cwrs32.c walks (its non-custom-modes table)
PN = [2, 3, 4, 6, 8, 9, 11, 12, 16, 18, 22, 24, 32, 36, 44, 48, 64, 72, 88, 96, 144, 176]
PKMAX = [128, 128, 128, 88, 36, 26, 18, 16, 12, 11, 9, 9, 7, 7, 6, 6, 5, 5, 5, 5, 4, 4]


def test_cwrs_bijection_over_the_mode_table(kat):
    """test_unit_cwrs32.c:88-160: for every (N, K) of the mode, indices 0, inc, 2 inc, ... (inc = V/20000): cwrsi gives a
    vector with K pulses and icwrs maps it back to the same index.  Both directions are the kernels' own routines
    (b3_cwrsi on the decode side, enc_icwrs on the encode side)."""
    y = np.zeros(256, dtype=np.int16)
    cells = 0
    for n, kmax in zip(PN, PKMAX):
        for pseudo in range(1, 41):
            k = kat.kat_get_pulses(pseudo)
            if k > kmax:
                break
            nc = kat.kat_pvq_v(n, k)
            inc = max(1, nc // 2000)                  # (a tenth of the reference's density keeps the CPU suite quick)
            for idx in range(0, nc, inc):
                yy = kat.kat_cwrsi(n, k, idx, y.ctypes.data)
                v = y[:n].astype(np.int64)
                assert int(np.abs(v).sum()) == k, (n, k, idx)
                assert int((v * v).sum()) == yy, (n, k, idx)
                assert kat.kat_icwrs(n, y.ctypes.data) == idx, (n, k, idx)
            cells += 1
    assert cells > 150


@needs_ref
def test_mathops_match_the_reference_value_for_value(kat):
    """Ranges of test_unit_mathops.c (testdiv :84-101, testsqrt :103-119, ...), but held to the reference's own
    fixed-point answers bit for bit instead of to a float tolerance."""
    ref = ref_lib()
    for i in list(range(1, 70000)) + list(range(70000, 327671, 7)):
        assert kat.kat_rcp(i) == ref.celt_rcp(i), i
    i = 1
    while i <= 1000000000:
        assert kat.kat_sqrt(i) == ref.celt_sqrt(i), i
        assert kat.kat_isqrt32(i) == ref.isqrt32(i) == math.isqrt(i), i
        i += 1 + (i >> 10)
    for x in range(16384, 65536, 3):                  # celt_rsqrt_norm: Q16 in [0.25, 1)
        assert kat.kat_rsqrt_norm(x) == ref.celt_rsqrt_norm(x), x
    for x in list(range(0, 140000, 7)) + [0, 32767, 32768, 65535, 65536, 65537, 98304, 131071, 131072]:
        assert kat.kat_cos_norm(x) == ref.celt_cos_norm(x), x
    rng = np.random.default_rng(5)
    for _ in range(20000):
        b = int(rng.integers(1 << 16, 1 << 31))
        a = int(rng.integers(0, b))
        assert kat.kat_frac_div32(a, b) == ref.frac_div32(a, b), (a, b)
    s = 12345
    for n in (1, 2, 3, 17, 120, 960, 100000):         # the O(log n) skip-ahead against n steps of celt_lcg_rand
        r = s
        for _ in range(n):
            r = ref.celt_lcg_rand(r)
        assert kat.kat_lcg_skip(s, n) == r, n


def test_log2_exp2_accuracy(kat):
    """test_unit_mathops.c testlog2 / testexp2 / testexp2log2 (FIXED_POINT branches): celt_log2 and celt_exp2 are inline
    in the reference (nothing exported to compare with), so the unit test's own tolerances are the bar."""
    x = 1
    while x <= 1073741824:
        err = abs(1.442695040888963387 * math.log(x / 16384.0) - kat.kat_log2(x) / 1024.0)
        assert err <= 0.003, (x, err)
        x += 1 + (x >> 8)
    for x in range(-32768, 15360):                    # testexp2, FIXED_POINT (:237-250)
        e = kat.kat_exp2(x)
        err1 = abs(x / 1024.0 - 1.442695040888963387 * math.log(e / 65536.0)) if e > 0 else 1e9
        err2 = abs(math.exp(0.6931471805599453094 * x / 1024.0) - e / 65536.0)
        assert err1 <= 0.0002 or err2 <= 0.00004, (x, e, err1, err2)
    x = 8
    while x < 65536:                                  # testexp2log2, FIXED_POINT (:252-264)
        err = abs(x - 0.25 * kat.kat_exp2(kat.kat_log2(x))) / 16384
        assert err <= 0.004, (x, err)
        x += x >> 3


def test_derived_rate_tables_equal_the_bisection(kat):
    """The parse walk reads bits2pulses (rate.h:48-70) and the split threshold (bands.c:893) from tables the
    generator derives (tools/gen_tables.py); every entry must equal what the reference's own search gives -- all
    (LM, band) rows of the mode, every budget up to well past the largest cache entry."""
    for LM in range(0, 4):
        for band in range(21):
            assert kat.kat_split_thresh(band, LM) == kat.kat_split_thresh_ref(band, LM)
            for bits in list(range(0, 420)) + [1000, 8191, 16383]:
                assert kat.kat_bits2pulses_table(band, LM, bits) == kat.kat_bits2pulses_bisect(band, LM, bits), (LM, band, bits)
    for band in range(8, 21):                       # LM = -1 rows exist from band 8 on (a split of an LM = 0 band)
        for bits in range(0, 300):
            assert kat.kat_bits2pulses_table(band, -1, bits) == kat.kat_bits2pulses_bisect(band, -1, bits)


def test_compute_qn(kat):
    """compute_qn (bands.c:596-616) against a restatement that divides the way the reference does."""
    exp2_table8 = [16384, 17866, 19483, 21247, 23170, 25267, 27554, 30048]

    def ref_qn(N, b, offset, pulse_cap, stereo):
        N2 = 2 * N - 1
        if stereo and N == 2:
            N2 -= 1
        num = b + N2 * offset
        qb = abs(num) // N2 * (1 if num >= 0 else -1)          # C division truncates toward zero
        qb = min(b - pulse_cap - 32, qb)
        qb = min(64, qb)
        if qb < 4:
            return 1
        qn = exp2_table8[qb & 7] >> (14 - (qb >> 3))
        return (qn + 1) >> 1 << 1

    for N in (2, 3, 4, 6, 8, 9, 11, 12, 16, 18, 22, 24, 36, 44, 48, 72, 88, 96, 144, 176):
        for stereo in (0, 1):
            for pulse_cap in (0, 8, 20, 33, 48, 64):
                offset = (pulse_cap >> 1) - (16 if stereo and N == 2 else 4)
                for b in list(range(0, 600, 7)) + [1000, 5000, 16383, 81600]:
                    assert kat.kat_compute_qn(N, b, offset, pulse_cap, stereo) == ref_qn(N, b, offset, pulse_cap, stereo)


def test_bitexact_cos_checksum(kat):
    """test_unit_mathops.c testbitexactcos :121-147: checksum and monotonic step bounds over 64..16320."""
    chk, last, min_d, max_d = 0, 32767, 32767, 0
    for i in range(64, 16321):
        q = kat.kat_bitexact_cos(i)
        chk ^= q * i
        d = last - q
        max_d, min_d, last = max(max_d, d), min(min_d, d), q
    assert (chk, max_d, min_d) == (89408644, 5, 0)
    assert (kat.kat_bitexact_cos(64), kat.kat_bitexact_cos(16320), kat.kat_bitexact_cos(8192)) == (32767, 200, 23171)


def test_bitexact_log2tan_checksum(kat):
    """test_unit_mathops.c testbitexactlog2tan :149-182."""
    chk, last, min_d, max_d = 0, 15059, 15059, 0
    for i in range(64, 8193):
        mid = kat.kat_bitexact_cos(i)
        side = kat.kat_bitexact_cos(16384 - i)
        q = kat.kat_bitexact_log2tan(mid, side)
        chk ^= q * i
        d = last - q
        assert q == -kat.kat_bitexact_log2tan(side, mid), i
        max_d, min_d, last = max(max_d, d), min(min_d, d), q
    assert (chk, max_d, min_d) == (15821257, 61, -2)
    assert kat.kat_bitexact_log2tan(32767, 200) == 15059
    assert kat.kat_bitexact_log2tan(30274, 12540) == 2611
    assert kat.kat_bitexact_log2tan(23171, 23171) == 0


def _script(seed, count):
    """A random sequence of coder operations in the spirit of test_unit_entropy.c:89-250 (uints of every size, raw bits,
    binary symbols) plus Laplace symbols as in test_unit_laplace.c:56-89 and 8-bit icdf symbols."""
    rng = np.random.default_rng(seed)
    kind = rng.integers(0, 5, count).astype(np.int32)
    a = np.zeros(count, dtype=np.int32)
    b = np.zeros(count, dtype=np.uint32)
    c = np.zeros(count, dtype=np.int32)
    for i in range(count):
        k = kind[i]
        if k == 0:
            ft = int(rng.integers(2, 1 << int(rng.integers(2, 31))))
            b[i], a[i] = ft, int(rng.integers(0, ft))
        elif k == 1:
            nb = int(rng.integers(1, 26))
            b[i], a[i] = nb, int(rng.integers(0, 1 << nb))
        elif k == 2:
            logp = int(rng.integers(1, 16))
            b[i], a[i] = logp, int(rng.random() < 2.0 ** -logp * 4)
        elif k == 3:
            a[i] = int(rng.integers(-14, 15)) if rng.random() < 0.9 else int(rng.integers(-300, 300))
            b[i] = int(rng.integers(1, 16000)) if i else 32000      # fs; first one as test_unit_laplace.c:68
            c[i] = int(rng.integers(0, 11000)) + 5 if i else 21
        else:
            p0 = int(rng.integers(2, 255))
            p1 = int(rng.integers(1, p0))
            b[i], c[i], a[i] = p0, p1, int(rng.integers(0, 3))
    return kind, a, b, c


@pytest.mark.parametrize("seed", range(6))
def test_range_coder_round_trip_and_reference_bytes(kat, seed):
    """Our encoder -> our decoder returns the script's values, tell_frac agrees symbol for symbol on both sides
    (test_unit_entropy.c:215-250), and -- with the compiled reference present -- the reference encoder fed the same
    script produces the same bytes and the reference decoder reads our bytes back."""
    count, size = 400, 4000
    kind, a, b, c = _script(seed, count)
    buf = np.zeros(size, dtype=np.uint8)
    coded = np.zeros(count, dtype=np.int32)
    tell_e = np.zeros(count, dtype=np.uint32)
    rng_e = ctypes.c_uint(0)
    p = lambda x: x.ctypes.data_as(ctypes.c_void_p)
    used = kat.kat_encode(p(kind), p(a), p(b), p(c), count, p(buf), size, p(coded), p(tell_e), ctypes.byref(rng_e))
    assert used > 0
    out = np.zeros(count, dtype=np.int32)
    tell_d = np.zeros(count, dtype=np.uint32)
    rng_d = ctypes.c_uint(0)
    err = kat.kat_decode(p(kind), p(b), p(c), count, p(buf), size, p(out), p(tell_d), ctypes.byref(rng_d))
    assert err == 0
    assert np.array_equal(out, coded)
    assert np.array_equal(tell_e, tell_d)
    assert rng_e.value == rng_d.value
    if not os.path.exists(REF_SO):
        return
    ref = ctypes.CDLL(REF_SO)
    ctx = (ctypes.c_ubyte * 256)()                       # room for struct ec_ctx (entcode.h:63-94)
    rbuf = np.zeros(size, dtype=np.uint8)
    ref.ec_enc_init(ctx, p(rbuf), ctypes.c_uint(size))
    for i in range(count):
        k = int(kind[i])
        if k == 0:
            ref.ec_enc_uint(ctx, ctypes.c_uint(int(a[i])), ctypes.c_uint(int(b[i])))
        elif k == 1:
            ref.ec_enc_bits(ctx, ctypes.c_uint(int(a[i])), ctypes.c_uint(int(b[i])))
        elif k == 2:
            ref.ec_enc_bit_logp(ctx, int(a[i]), ctypes.c_uint(int(b[i])))
        elif k == 3:
            v = ctypes.c_int(int(a[i]))
            ref.ec_laplace_encode(ctx, ctypes.byref(v), ctypes.c_uint(int(b[i])), int(c[i]))
            assert v.value == coded[i]
        else:
            t = (ctypes.c_ubyte * 3)(int(b[i]), int(c[i]), 0)
            ref.ec_enc_icdf(ctx, int(a[i]), t, ctypes.c_uint(8))
        ref.ec_tell_frac.restype = ctypes.c_uint
        assert ref.ec_tell_frac(ctx) == tell_e[i], i
    ref.ec_enc_done(ctx)
    assert np.array_equal(rbuf, buf), "our encoder's bytes differ from the reference encoder's"
    ref.ec_dec_init(ctx, p(buf), ctypes.c_uint(size))
    ref.ec_dec_uint.restype = ctypes.c_uint
    ref.ec_dec_bits.restype = ctypes.c_uint
    for i in range(count):
        k = int(kind[i])
        if k == 0:
            v = ref.ec_dec_uint(ctx, ctypes.c_uint(int(b[i])))
        elif k == 1:
            v = ref.ec_dec_bits(ctx, ctypes.c_uint(int(b[i])))
        elif k == 2:
            v = ref.ec_dec_bit_logp(ctx, ctypes.c_uint(int(b[i])))
        elif k == 3:
            v = ref.ec_laplace_decode(ctx, ctypes.c_uint(int(b[i])), int(c[i]))
        else:
            t = (ctypes.c_ubyte * 3)(int(b[i]), int(c[i]), 0)
            v = ref.ec_dec_icdf(ctx, t, ctypes.c_uint(8))
        assert v == coded[i], i


@needs_ref
def test_alg_unquant_against_the_reference(kat):
    """test_unit_rotation.c exercises exp_rotation through its SNR; here the whole of alg_unquant (cwrsi +
    normalise_residual + exp_rotation with dir = -1 + extract_collapse_mask, vq.c:329-347) is held to the reference's
    exported alg_unquant, vector for vector, over the mode's (N, K) cells, every spread and block count."""
    ref = ctypes.CDLL(REF_SO)
    ref.alg_unquant.restype = ctypes.c_uint
    rng = np.random.default_rng(11)
    ctx = (ctypes.c_ubyte * 256)()
    p = lambda x: x.ctypes.data_as(ctypes.c_void_p)
    checked = 0
    for n, kmax in zip(PN, PKMAX):
        for pseudo in (1, 2, 3, 5, 8, 12, 17, 23, 30, 40):
            k = kat.kat_get_pulses(pseudo)
            if k > kmax:
                break
            v = kat.kat_pvq_v(n, k)
            for B in (1, 2, 4, 8):
                if n % B or n // B < 1:
                    continue
                for spread in (0, 1, 2, 3):
                    idx = int(rng.integers(0, v))
                    gain = int(rng.integers(1000, 32768))
                    buf = np.zeros(64, dtype=np.uint8)
                    ref.ec_enc_init(ctx, p(buf), ctypes.c_uint(64))
                    ref.ec_enc_uint(ctx, ctypes.c_uint(idx), ctypes.c_uint(v))
                    ref.ec_enc_done(ctx)
                    ref.ec_dec_init(ctx, p(buf), ctypes.c_uint(64))
                    xr = np.zeros(n, dtype=np.int16)
                    mr = ref.alg_unquant(p(xr), n, k, spread, B, ctx, ctypes.c_int16(gain))
                    xo = np.zeros(n + 8, dtype=np.int16)
                    mo = kat.kat_alg_unquant(p(xo), n, k, idx, gain, B, spread)
                    assert np.array_equal(xr, xo[:n]), (n, k, B, spread, idx)
                    assert mr == mo, (n, k, B, spread, idx)
                    checked += 1
    assert checked > 1200


@pytest.mark.gpu
def test_gpu_cwrs_bijection_over_the_mode_table(kat):
    """The same walk as test_cwrs_bijection_over_the_mode_table on the device (opus_b200_selftest_cwrs): b3_cwrsi with a
    lane per vector, enc_icwrs with a warp per vector; the vectors must also equal the host build's."""
    import torch
    import concentus_b200 as cb
    lib = cb.load()
    ns, ks, idxs = [], [], []
    for n, kmax in zip(PN, PKMAX):
        for pseudo in range(1, 41):
            k = kat.kat_get_pulses(pseudo)
            if k > kmax:
                break
            nc = kat.kat_pvq_v(n, k)
            inc = max(1, nc // 400)
            for idx in range(0, nc, inc):
                ns.append(n); ks.append(k); idxs.append(idx)
    count = len(ns)
    assert count > 40000
    d_n = torch.tensor(ns, dtype=torch.int32, device="cuda")
    d_k = torch.tensor(ks, dtype=torch.int32, device="cuda")
    d_i = torch.tensor(np.array(idxs, dtype=np.uint32).view(np.int32), dtype=torch.int32, device="cuda")
    d_y = torch.zeros((count, 176), dtype=torch.int16, device="cuda")
    d_yy = torch.zeros(count, dtype=torch.int32, device="cuda")
    d_b = torch.zeros(count, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    vp = ctypes.c_void_p
    lib.opus_b200_selftest_cwrs.argtypes = [vp, vp, vp, ctypes.c_int, vp, vp, vp]
    assert lib.opus_b200_selftest_cwrs(d_n.data_ptr(), d_k.data_ptr(), d_i.data_ptr(), count, d_y.data_ptr(), d_yy.data_ptr(),
                                       d_b.data_ptr()) == 0
    y = d_y.cpu().numpy().astype(np.int64)
    back = d_b.cpu().numpy().view(np.uint32)
    assert np.array_equal(back, np.array(idxs, dtype=np.uint32))
    assert np.array_equal(np.abs(y).sum(axis=1), np.array(ks))
    assert np.array_equal((y * y).sum(axis=1), d_yy.cpu().numpy())
    ref = np.zeros(256, dtype=np.int16)
    for j in range(0, count, 37):
        kat.kat_cwrsi(ns[j], ks[j], idxs[j], ref.ctypes.data)
        assert np.array_equal(ref[:ns[j]], y[j, :ns[j]]), j


@pytest.mark.gpu
def test_gpu_isqrt_exact():
    """The device form of isqrt32 (float estimate + 64-bit correction) is exact: every value up to 2^20, the
    neighbourhood of every perfect square up to 2^32, and a random sample of the whole 32-bit range."""
    import torch
    import concentus_b200 as cb
    lib = cb.load()
    rng = np.random.default_rng(3)
    k = np.arange(1, 65536, dtype=np.uint64)
    near = np.concatenate([k * k - 1, k * k, k * k + 1, np.array([0xFFFFFFFF, 0xFFFE0001, 0xFFFE0000], dtype=np.uint64)])
    vals = np.concatenate([np.arange(1, 1 << 20, dtype=np.uint64), near[near < (1 << 32)],
                           rng.integers(1, 1 << 32, 2000000, dtype=np.uint64)]).astype(np.uint32)
    d_in = torch.from_numpy(vals.view(np.int32)).cuda()
    d_out = torch.zeros_like(d_in)
    vp = ctypes.c_void_p
    lib.opus_b200_selftest_isqrt.argtypes = [vp, ctypes.c_int, vp]
    assert lib.opus_b200_selftest_isqrt(d_in.data_ptr(), len(vals), d_out.data_ptr()) == 0
    got = d_out.cpu().numpy().view(np.uint32).astype(np.uint64)
    v = vals.astype(np.uint64)
    assert bool((got * got <= v).all()) and bool(((got + 1) * (got + 1) > v).all())
