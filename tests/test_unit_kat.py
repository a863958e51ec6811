"""The reference's unit known answers for the CELT building blocks, replayed on three
implementations of each block through one word protocol:
  oracle  oracle/celt_oracle.cpp::orc_kat        (the checker itself gets pinned)
  emul    tests/emul (mb_kat.cuh, team of one)   the kernel bodies on the CPU
  gpu     bd_debug_kat (mb_kat.cuh, one warp)    the kernel bodies on the B200         [-m gpu]
Each test cites the reference test it restates. Expected values are the reference's literals where it holds literals,
and its own straight-loop / naive definitions (restated here in numpy) where its tests compare against those."""
import ctypes as C
import math

import numpy as np
import pytest

import helpers

BACKENDS = ["oracle", "emul", pytest.param("gpu", marks=pytest.mark.gpu)]


def f2w(x):
    return np.asarray(x, dtype=np.float32).view(np.uint32)


def w2f(w):
    return np.asarray(w, dtype=np.uint32).view(np.float32)


def kat(backend, what, words, n_out):
    words = np.ascontiguousarray(words, dtype=np.uint32)
    out = np.zeros(max(n_out, 1), dtype=np.uint32)
    if backend == "oracle":
        lib = helpers.oracle_lib()
        rc = lib.orc_kat(what, helpers.ptr(words), len(words), helpers.ptr(out), n_out)
    elif backend == "emul":
        lib = helpers.emul_lib()
        rc = lib.emu_kat(what, helpers.ptr(words), len(words), helpers.ptr(out), n_out)
    else:
        import mousiki_b200
        lib = mousiki_b200.load_library()
        lib.bd_debug_kat.argtypes = [C.c_int, C.c_int, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint32]
        rc = lib.bd_debug_kat(0, what, helpers.ptr(words), len(words), helpers.ptr(out), n_out)
    assert rc == 0, (backend, what, rc)
    return out[:n_out]


def s32(w):
    return np.asarray(w, dtype=np.uint32).view(np.int32).astype(np.int64)


# ------------------------------------------------------------------------------------------------ bands.rs:4805-4915
@pytest.mark.parametrize("backend", BACKENDS)
def test_bitexact_cos_checksum(backend):
    """src/celt/bands.rs:4805-4831 (port of testbitexactcos): checksum, monotonicity bounds, sample points."""
    xs = np.arange(64, 16321, dtype=np.int64)
    q = s32(kat(backend, 1, xs.astype(np.uint32), len(xs)))
    chk = 0
    for qi, i in zip(q.tolist(), xs.tolist()):
        chk ^= qi * i
    d = np.concatenate([[32767], q[:-1]]) - q
    assert chk == 89408644 and d.max() == 5 and d.min() == 0
    assert q[0] == 32767 and q[-1] == 200 and q[8192 - 64] == 23171
    # bands.rs:4886-4896 sample points (negative and zero arguments)
    inputs = np.array([-16383, -12000, -6000, -1, 0, 1, 6000, 12000, 16383], dtype=np.int32)
    got = s32(kat(backend, 1, inputs.view(np.uint32), len(inputs)))
    assert got.tolist() == [3, 13371, 27494, -32768, -32768, -32768, 27494, 13371, 3]


@pytest.mark.parametrize("backend", BACKENDS)
def test_bitexact_log2tan_checksum(backend):
    """src/celt/bands.rs:4836-4884 (port of testbitexactlog2tan) and :4898-4915."""
    xs = np.arange(64, 8193, dtype=np.int64)
    mid = s32(kat(backend, 1, xs.astype(np.uint32), len(xs)))
    side = s32(kat(backend, 1, (16384 - xs).astype(np.uint32), len(xs)))
    pairs = np.stack([mid, side], axis=1).reshape(-1).astype(np.uint32)
    q = s32(kat(backend, 2, pairs, len(xs)))
    anti = s32(kat(backend, 2, np.stack([side, mid], axis=1).reshape(-1).astype(np.uint32), len(xs)))
    chk = 0
    for qi, i in zip(q.tolist(), xs.tolist()):
        chk ^= qi * i
    d = np.concatenate([[15059], q[:-1]]) - q
    assert chk == 15821257 and d.max() == 61 and d.min() == -2
    assert np.array_equal(q, -anti)
    cases = [(32767, 200, 15059), (30274, 12540, 2611), (23171, 23171, 0), (23170, 32767, -1025), (11585, 32767, -3073),
             (16384, 23170, -993), (30000, 12345, 2631), (12345, 30000, -2631), (1, 32767, -30690), (32767, 1, 30690)]
    got = s32(kat(backend, 2, np.array([v for c in cases for v in c[:2]], dtype=np.uint32), len(cases)))
    assert got.tolist() == [c[2] for c in cases]


# ------------------------------------------------------------------------------------------------ cwrs.rs:681-786
def _u_table(nmax, kmax):
    """U(n, k) of RFC 6716 section 4.3.4.2 by its recurrence U(n,k) = U(n-1,k) + U(n,k-1) + U(n-1,k-1)."""
    U = [[0] * (kmax + 2) for _ in range(nmax + 1)]
    U[0][0] = 1
    for n in range(1, nmax + 1):
        for k in range(1, kmax + 2):
            U[n][k] = U[n - 1][k] + U[n][k - 1] + U[n - 1][k - 1]
    return U


_PVQ_ROW_LEN = [209, 208, 207, 206, 205, 204, 104, 54, 33, 21, 15, 10, 7, 4, 1]


def _in_static_table(n, k):
    """The decode path reads CELT_PVQ_U_DATA (src/celt/cwrs_pvq.rs:12-54): rows min(n,k) <= 14, bounded columns."""
    row, col = min(n, k + 1), max(n, k + 1)
    return row <= 14 and col <= row + _PVQ_ROW_LEN[row] - 1


def _icwrs(y, U):
    n = len(y)
    j = n - 1
    i = 1 if y[j] < 0 else 0
    k = abs(y[j])
    while j > 0:
        j -= 1
        i += U[n - j][k]
        k += abs(y[j])
        if y[j] < 0:
            i += U[n - j][k + 1]
    return i


@pytest.mark.parametrize("backend", BACKENDS)
def test_cwrs_roundtrip_comprehensive(backend):
    """src/celt/cwrs.rs:717-786 (port of test_unit_cwrs32.c): for the (n, k) grid of the reference test that the decode
    path's static table covers, index -> pulses (the implementation under test) -> index (icwrs restated from RFC 6716
    section 4.3.4.2 on an independently generated U table) is the identity, sum |y| == k, V(n,k) matches."""
    PN = [2, 3, 4, 6, 8, 9, 11, 12, 16, 18, 22, 24, 32, 36, 44, 48]
    PKMAX = [128, 128, 128, 88, 36, 26, 18, 16, 12, 11, 9, 9, 7, 7, 6, 6]
    U = _u_table(48, 130)
    tested = 0
    for n, kmax in zip(PN, PKMAX):
        for k in range(1, min(kmax, 32) + 1):
            if not _in_static_table(n, k):
                continue
            nc = U[n][k] + U[n][k + 1]
            if nc >= 1 << 32:
                continue
            cap = 20000 if backend == "oracle" else 1500   # the reference walks ~20 000 indices per (n, k)
            inc = max(nc // cap, 1)
            count = (nc + inc - 1) // inc
            y = s32(kat(backend, 3, np.array([n, k, count, inc, 0], dtype=np.uint32), count * n)).reshape(count, n)
            assert (np.abs(y).sum(axis=1) == k).all(), (n, k)
            step = max(count // 200, 1)
            for j in range(0, count, step):
                assert _icwrs(y[j].tolist(), U) == j * inc, (n, k, j)
            tested += 1
    assert tested > 150


@pytest.mark.parametrize("backend", BACKENDS)
def test_cwrs_small_vectors_exhaustive(backend):
    """src/celt/cwrs.rs:681-713: every pulse vector of (n, k) in {(2,1), (2,2), (3,2)} is produced exactly once."""
    U = _u_table(4, 4)
    for n, k in [(2, 1), (2, 2), (3, 2)]:
        nc = U[n][k] + U[n][k + 1]
        y = s32(kat(backend, 3, np.array([n, k, nc, 1, 0], dtype=np.uint32), nc * n)).reshape(nc, n)
        assert len({tuple(v) for v in y.tolist()}) == nc and (np.abs(y).sum(axis=1) == k).all()
        for j in range(nc):
            assert _icwrs(y[j].tolist(), U) == j


# ------------------------------------------------------------------------------------------------ celt.rs:868-1190
TAPSET_GAINS = np.array([[0.3066406250, 0.2170410156, 0.1296386719], [0.4638671875, 0.2680664062, 0.0], [0.7998046875, 0.1000976562, 0.0]],
                        dtype=np.float32)


def _celt_window120():
    i = np.arange(120, dtype=np.float64)
    return np.sin(0.5 * np.pi * np.sin(0.5 * np.pi * (i + 0.5) / 120) ** 2).astype(np.float32)


def comb_filter_reference(x, x_start, n, t0, t1, g0, g1, tapset0, tapset1, window, overlap):
    """The reference test's own straight loops (src/celt/celt.rs:905-986 + :876-903), in float32."""
    f32 = np.float32
    x = np.asarray(x, dtype=np.float32)
    if g0 == 0 and g1 == 0:
        return x[x_start:x_start + n].copy()
    t0, t1 = max(t0, 15), max(t1, 15)
    g00, g01, g02 = [f32(g0) * TAPSET_GAINS[tapset0][i] for i in range(3)]
    g10, g11, g12 = [f32(g1) * TAPSET_GAINS[tapset1][i] for i in range(3)]
    out = np.zeros(n, dtype=np.float32)
    overlap = min(overlap, n)
    if g0 == g1 and t0 == t1 and tapset0 == tapset1:
        overlap = 0
    for i in range(overlap):
        f = window[i] * window[i]
        omf = f32(1.0) - f
        s = x_start + i
        out[i] = (x[s] + omf * g00 * x[s - t0] + omf * g01 * (x[s - t0 + 1] + x[s - t0 - 1]) + omf * g02 * (x[s - t0 + 2] + x[s - t0 - 2])
                  + f * g10 * x[s - t1] + f * g11 * (x[s - t1 + 1] + x[s - t1 - 1]) + f * g12 * (x[s - t1 + 2] + x[s - t1 - 2]))
    if g1 == 0:
        out[overlap:] = x[x_start + overlap:x_start + n]
        return out
    for i in range(overlap, n):
        s = x_start + i
        out[i] = x[s] + g10 * x[s - t1] + g11 * (x[s - t1 + 1] + x[s - t1 - 1]) + g12 * (x[s - t1 + 2] + x[s - t1 - 2])
    return out


def _comb_case(backend, x, hist, n, t0, t1, g0, g1, tap0, tap1):
    words = np.concatenate([np.array([n, t0, t1, tap0, tap1], dtype=np.uint32), f2w([g0, g1]), np.array([hist], dtype=np.uint32),
                            f2w(x[:hist + n])])
    return w2f(kat(backend, 4, words, n))


@pytest.mark.parametrize("backend", BACKENDS)
def test_comb_filter_matches_reference_loops(backend):
    """src/celt/celt.rs:1091-1121 / :1138-1176 (comb_filter / comb_filter_in_place vs the straight loops, 1e-6): the
    reference's signal sin(0.13 i), periods 21 -> 27, gains 0.6 -> 0.35, tapsets 0 -> 2 — here with the static mode's
    window (the product has no other), over 12 samples and over a full 20 ms frame part (overlap 120, then the constant
    section). In place: the filter reads its own output T-2..T+2 back only through samples it has not rewritten yet
    when T >= n; the long case (n > T) is the in-place semantics the decoder relies on."""
    win = _celt_window120()
    for n, hist in [(12, 30), (480, 1024), (840, 1024)]:
        x = np.sin(np.arange(hist + n + 6, dtype=np.float32) * np.float32(0.13)).astype(np.float32)
        got = _comb_case(backend, x, hist, n, 21, 27, 0.6, 0.35, 0, 2)
        if n <= 19:   # no sample depends on a rewritten one: the out-of-place reference loops apply as they are
            want = comb_filter_reference(x, hist, n, 21, 27, np.float32(0.6), np.float32(0.35), 0, 2, win, 120)
        else:         # in place: run the same loops sample by sample on a buffer that is updated as it goes
            buf = x.copy()
            for i in range(n):
                buf[hist + i] = comb_filter_reference(buf, hist + i, 1, 21, 27, np.float32(0.6), np.float32(0.35), 0, 2, win[i:], 120 - i)[0] \
                    if i < 120 else comb_filter_reference(buf, hist + i, 1, 27, 27, np.float32(0.35), np.float32(0.35), 2, 2, win, 0)[0]
            want = buf[hist:hist + n]
        assert np.abs(got - want).max() <= 1e-6, (n, float(np.abs(got - want).max()))


@pytest.mark.parametrize("backend", BACKENDS)
def test_comb_filter_const_and_zero_gain(backend):
    """src/celt/celt.rs:1061-1089 (constant section: ramp signal, T = 15) and :1123-1136 / :1178-1190 (zero gains copy)."""
    hist, n = 17, 6
    x = (np.arange(hist + n + 4, dtype=np.float32) * np.float32(0.25)).astype(np.float32)
    # equal old and new parameters skip the cross-fade: the whole run is comb_filter_const with g1's taps
    got = _comb_case(backend, x, hist, n, 15, 15, 0.5, 0.5, 1, 1)
    g = np.float32(0.5) * TAPSET_GAINS[1]
    want = np.array([x[hist + i] + g[0] * x[hist + i - 15] + g[1] * (x[hist + i - 14] + x[hist + i - 16]) + g[2] * (x[hist + i - 13] + x[hist + i - 17])
                     for i in range(n)], dtype=np.float32)
    assert np.abs(got - want).max() <= 1e-6
    hist, n = 20, 8
    x = (np.arange(hist + n, dtype=np.float32) * np.float32(0.5)).astype(np.float32)
    got = _comb_case(backend, x, hist, n, 10, 12, 0.0, 0.0, 0, 1)
    assert np.array_equal(got, x[hist:hist + n])


# ------------------------------------------------------------------------------------------------ kiss_fft.rs:1364-1440
@pytest.mark.parametrize("backend", BACKENDS)
def test_ifft_snr_across_static_sizes(backend):
    """src/celt/kiss_fft.rs:1371-1440: LCG input (seed 1, scaled by 32768/nfft as in generate_input(inverse)), inverse
    transform of the four sizes the 48 kHz mode uses (480, 240, 120, 60) against the f64 DFT definition: SNR >= 60 dB."""
    for shift in range(4):
        nfft = 480 >> shift
        seed = 1
        vals = []
        for _ in range(2 * nfft):
            seed = (seed * 1664525 + 1013904223) & 0xFFFFFFFF
            vals.append(np.float32((seed & 0x7FFF) - 16384.0) * np.float32(32768.0))
        x = (np.array(vals, dtype=np.float32) * np.float32(1.0 / nfft)).astype(np.float32)
        out = w2f(kat(backend, 5, np.concatenate([np.array([shift], dtype=np.uint32), f2w(x)]), 2 * nfft)).astype(np.float64)
        z = x[0::2].astype(np.float64) + 1j * x[1::2].astype(np.float64)
        k = np.arange(nfft)
        ans = (np.exp(2j * np.pi * np.outer(k, k) / nfft) * z[None, :]).sum(axis=1)
        got = out[0::2] + 1j * out[1::2]
        snr = 10 * np.log10((np.abs(ans) ** 2).sum() / (np.abs(ans - got) ** 2).sum())
        assert snr >= 60.0, (nfft, snr)


# ------------------------------------------------------------------------------------------------ mdct.rs:1746-1843
def naive_frame_imdct(freq, lm, transient):
    """clt_mdct_backward by its definition (src/celt/mdct.rs:1641-1656 reference_backward: pre-rotation, DFT by the
    formula, post-rotation, window mirror), f64, twiddles and window from their closed forms — plus the frame
    layout of celt_synthesis (short blocks interleaved, each transform run on a zero-filled tail, src/celt/mdct.rs:373)."""
    M = 1 << lm
    N = 120 * M
    nbk, nb = (M, 120) if transient else (1, N)
    n2, n4 = nb, nb // 2
    n = 2 * nb
    i = np.arange(n4)
    tw = np.cos(2 * np.pi * (np.arange(n2) + 0.125) / n)      # trig[i] = cos(2 pi (i + 1/8) / N), N/2 entries
    t0, t1 = tw[:n4], tw[n4:]
    win = np.sin(0.5 * np.pi * np.sin(0.5 * np.pi * (np.arange(120) + 0.5) / 120) ** 2)
    out = np.zeros(N + 120)
    for b in range(nbk):
        x = np.asarray(freq, dtype=np.float64)[b::nbk][:nb]
        x1, x2 = x[2 * i], x[n2 - 1 - 2 * i]
        pre = (x2 * t0 + x1 * t1) + 1j * (x1 * t0 - x2 * t1)
        k = np.arange(n4)
        f = (np.exp(2j * np.pi * np.outer(k, k) / n4) * pre[None, :]).sum(axis=1)
        yr = f.real * t0 + f.imag * t1
        yi = f.real * t1 - f.imag * t0
        temp = np.zeros(n2)
        temp[2 * i] = yr
        temp[n2 - 1 - 2 * i] = yi
        o = out[b * nb:]
        o[:] = 0.0
        o[60:60 + n2] = temp
        for j in range(60):
            a, c = o[119 - j], o[j]
            o[j] = c * win[119 - j] - a * win[j]
            o[119 - j] = c * win[j] + a * win[119 - j]
    return out[:N]


@pytest.mark.parametrize("backend", BACKENDS)
def test_frame_imdct_matches_naive_definition(backend):
    """src/celt/mdct.rs:1746-1764, :1771-1826 (backward transform vs the naive reference, input cos(0.19 i), tolerance
    1e-3 + n * 1e-5) on every transform the 48 kHz mode runs: long blocks of 120/240/480/960 bins and 2/4/8 short blocks."""
    for lm in range(4):
        for transient in ([0] if lm == 0 else [0, 1]):
            N = 120 << lm
            freq = np.cos(np.arange(N, dtype=np.float32) * np.float32(0.19)).astype(np.float32)
            want = naive_frame_imdct(freq, lm, transient)
            n = 2 * (120 if transient else N)
            tol = 1e-3 + n * 1e-5
            for case in ((6,) if backend == "oracle" else (6, 9)):   # the team form over shared memory, and the warp-resident form k_imdct runs
                got = w2f(kat(backend, case, np.concatenate([np.array([lm, transient], dtype=np.uint32), f2w(freq)]), N)).astype(np.float64)
                assert np.abs(got - want).max() < tol, (case, lm, transient, float(np.abs(got - want).max()))


@pytest.mark.parametrize("backend", [b for b in BACKENDS if b != "oracle"])
def test_warp_resident_imdct_is_bit_identical_to_the_team_form(backend):
    """k_imdct's register / shuffle layout computes every butterfly of src/celt/kiss_fft.rs:900-1340 and the rotations of
    src/celt/mdct.rs:197-388 with the same operands in the same order as the shared-memory team form: the two agree bit for
    bit, the 60 samples behind the frame (the overlap region a pitch concealment reads) included."""
    rng = np.random.default_rng(77)
    for lm in range(4):
        for transient in ([0] if lm == 0 else [0, 1]):
            N = 120 << lm
            for scale in (1.0, 1e-3, 3e4):
                freq = (rng.standard_normal(N) * scale).astype(np.float32)
                freq[rng.integers(0, N, N // 8)] = 0.0
                words = np.concatenate([np.array([lm, transient], dtype=np.uint32), f2w(freq)])
                a = kat(backend, 6, words, N + 60)
                b = kat(backend, 9, words, N + 60)
                assert np.array_equal(a, b), (lm, transient, scale)


# ------------------------------------------------------------------------------------------------ entdec.rs / range.rs
def _script(backend, data, ops, tables=()):
    """ops: (op, a, b); an icdf op names its table by index into `tables` (appended to the input, one entry per word)."""
    words = [len(data)] + list(data)
    base = len(words) + 3 * len(ops)
    offs = []
    for t in tables:
        offs.append(base)
        base += len(t)
    for op, a, b in ops:
        words += [op, offs[a] if op == 1 else a, b]
    for t in tables:
        words += list(t)
    return kat(backend, 7, np.array(words, dtype=np.uint32), len(ops))


@pytest.mark.parametrize("backend", BACKENDS)
def test_range_decoder_initial_state(backend):
    """src/celt/entdec.rs:287-337: a fresh decoder on AA BB CC DD against the initialisation written out step by step."""
    buf = [0xAA, 0xBB, 0xCC, 0xDD]
    rng, offs = 1 << 7, 0
    nbits = 32 + 1 - ((32 - 7) // 8) * 8
    rem = buf[offs]; offs += 1
    val = rng - 1 - (rem >> 1)
    while rng <= (1 << 23):
        nbits += 8
        rng <<= 8
        sym = rem
        rem = buf[offs] if offs < len(buf) else 0
        offs += 1 if offs < len(buf) else 0
        sym = ((sym << 8) | rem) >> 1
        val = ((val << 8) + (255 & ~sym)) & 0x7FFFFFFF
    got = _script(backend, buf, [(8, 0, 0), (6, 0, 0), (7, 0, 0), (9, 0, 0), (10, 0, 0), (4, 0, 0)])
    assert got.tolist() == [nbits, rng, val, offs, rem, 1]     # ec_tell of a fresh decoder is 1


@pytest.mark.parametrize("backend", BACKENDS)
def test_range_decoder_raw_bits_and_uint(backend):
    """src/celt/entdec.rs:339-362: raw bits come from the end of the buffer, LSB first; dec_uint stays below ft."""
    got = _script(backend, [0, 0, 0, 0b10110101], [(8, 0, 0), (3, 4, 0), (8, 0, 0), (12, 0, 0), (3, 4, 0)])
    assert got[1] == 0b0101 and got[2] == got[0] + 4 and got[3] == 32 - 4 and got[4] == 0b1011
    got = _script(backend, [0xFF] * 4, [(2, 16, 0), (11, 0, 0)])
    assert got[0] < 16 and got[1] == 0


def _cdf_to_icdf(cdf):
    return [256 - c for c in cdf]


@pytest.mark.parametrize("backend", BACKENDS)
def test_range_decoder_symbol_walk(backend):
    """src/range.rs:659-695: the SILK symbol walk over 0b e4 c1 36 ec c5 80 — two logp(1) flags, then the frame type,
    gain, LSF, interpolation, seed, rate and pulse-count contexts (cumulative tables of src/range.rs:552-655, turned into
    the inverse-CDF form the path's decoder reads)."""
    T = [_cdf_to_icdf(c) for c in [
        [26, 256],                                                            # 0 SILK_FRAME_TYPE_INACTIVE
        [32, 144, 212, 241, 253, 254, 255, 256],                              # 1 SILK_GAIN_HIGH_BITS[0]
        [32, 64, 96, 128, 160, 192, 224, 256],                                # 2 SILK_GAIN_LOW_BITS
        [6, 11, 22, 53, 185, 206, 214, 218, 221, 223, 225, 227, 228, 229, 230, 231, 232, 233, 234, 235, 236, 237, 238, 239, 240, 241,
         242, 243, 244, 245, 246, 247, 248, 249, 250, 251, 252, 253, 254, 255, 256],   # 3 SILK_GAIN_DELTA
        [31, 52, 55, 72, 73, 81, 98, 102, 103, 121, 137, 141, 143, 146, 147, 157, 158, 161, 177, 188, 204, 206, 208, 211, 213, 224,
         225, 229, 238, 246, 253, 256],                                       # 4 SILK_LSF_S1[1][0]
        [1, 2, 5, 47, 199, 252, 254, 255, 256],                               # 5 SILK_LSF_S2[10]
        [1, 2, 4, 32, 218, 253, 254, 255, 256],                               # 6 SILK_LSF_S2[9]
        [1, 2, 3, 12, 244, 253, 254, 255, 256],                               # 7 SILK_LSF_S2[8]
        [13, 35, 64, 75, 256],                                                # 8 SILK_LSF_INTERPOLATION_OFFSET
        [64, 128, 192, 256],                                                  # 9 SILK_LCG_SEED
        [15, 66, 78, 124, 169, 182, 215, 242, 256],                           # 10 SILK_EXC_RATE[0]
        [131, 205, 230, 238, 241, 244, 245, 246, 247, 248, 249, 250, 251, 252, 253, 254, 255, 256],   # 11 SILK_PULSE_COUNT[0]
    ]]
    ops = [(0, 1, 0), (0, 1, 0), (1, 0, 8), (1, 1, 8), (1, 2, 8), (1, 3, 8), (1, 3, 8), (1, 3, 8), (1, 4, 8), (1, 5, 8), (1, 6, 8)]
    ops += [(1, 7, 8)] * 14 + [(1, 8, 8), (1, 9, 8), (1, 10, 8)] + [(1, 11, 8)] * 20
    # decode_symbol_logp returns 1 when the low-probability branch is taken; the path's bit_logp returns the same bit
    want = [0, 0, 1, 0, 6, 0, 3, 4, 9, 5, 4] + [4] * 14 + [4, 2, 0] + [0] * 20
    got = _script(backend, [0x0B, 0xE4, 0xC1, 0x36, 0xEC, 0xC5, 0x80], ops, T)
    assert got.tolist() == want


# ------------------------------------------------------------------------------------------------ rate.rs:1424-1465
@pytest.mark.parametrize("backend", BACKENDS)
def test_bits2pulses_pulses2bits_round_trip(backend):
    """src/celt/rate.rs:1424-1465 on the real 48 kHz mode: for every band, every LM (-1..3) and every pulse count of the
    cache row whose entry differs from its predecessor's, pulses2bits >= 1 and bits2pulses(pulses2bits(p)) == p;
    pulses2bits(0) == 0."""
    trip = []
    for lm in range(-1, 4):
        for band in range(21):
            for p in range(0, 41):
                trip += [band, lm & 0xFFFFFFFF, p]
    out = s32(kat(backend, 8, np.array(trip, dtype=np.uint32), len(trip))).reshape(-1, 3)
    by = {tuple(k): tuple(v) for k, v in zip(np.array(trip, dtype=np.uint32).view(np.int32).reshape(-1, 3).tolist(), out.tolist())}
    checked = 0
    for (band, lm, p), (bits, back, maxp) in by.items():
        if p == 0:
            assert bits == 0
            continue
        if p > maxp:      # beyond the row (or the band has no row at this LM)
            continue
        prev = by[(band, lm, p - 1)][0] if p > 1 else maxp + 1   # the reference compares table[p] with table[p - 1]; table[0] is the row's limit
        if bits == prev:
            continue
        assert bits >= 1 and back == p, (band, lm, p, bits, back)
        checked += 1
    assert checked > 1000, checked
