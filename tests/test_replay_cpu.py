"""Device code replayed on the CPU (tests/emu) against the oracle: the same ckks_core.cuh phases the
kernels run, thread by thread.  Bit-exact for integer work and — because the FFT butterflies are unfused
and follow the reference's order — for encode/decode as well."""
import ctypes as C
import math

import numpy as np
import pytest

N = 8192


def vp(a):
    return C.c_void_p(a.ctypes.data)


def test_reduction_schedule(emu, emu_ctx):
    # 60-bit q0 (bounds grow 2q per stage, 16q fit in 64 bits): one mid-transform reduction (end of pass 1),
    # shift-type quick reduce;
    # 52-bit q1: no mid reduction, "just above a power of two" quick reduce
    h = emu_ctx["pybind"]
    assert (emu.L.emu_fwd_mask(h, 0), emu.L.emu_rmode(h, 0)) == (2, 0)
    assert (emu.L.emu_fwd_mask(h, 1), emu.L.emu_rmode(h, 1)) == (0, 1)


@pytest.mark.parametrize("pattern", ["random", "max", "zero", "delta", "alternating"])
def test_ntt_matches_oracle(emu, emu_ctx, oracles, pattern):
    o, h = oracles["pybind"], emu_ctx["pybind"]
    rng = np.random.default_rng(7)
    for l in range(2):
        q = o.modulus(l)
        a = {"random": rng.integers(0, q, N, dtype=np.uint64), "max": np.full(N, q - 1, np.uint64),
             "zero": np.zeros(N, np.uint64), "delta": np.eye(1, N, 5, dtype=np.uint64)[0] * np.uint64(q - 1),
             "alternating": np.where(np.arange(N) % 2 == 0, q - 1, 0).astype(np.uint64)}[pattern]
        f = a.copy(); emu.L.emu_ntt_forward(h, l, vp(f))
        assert np.array_equal(f, o.ntt_forward(l, a))
        b = a.copy(); emu.L.emu_ntt_inverse(h, l, vp(b))
        assert np.array_equal(b, o.ntt_inverse(l, a))


def test_generic_modulus_uses_barrett_path(emu, oracle_mod):
    """A prime far from a power of two (quick-reduce mode 2) transforms correctly: inverse(forward) = identity."""
    M = 2 * N
    q = oracle_mod.previous_prime((3 << 43) + 1, M)      # 1.5 * 2^44: nowhere near a power of two
    psi = oracle_mod.min_root_of_unity(M, q)
    h = emu.context([q], [psi])
    assert emu.L.emu_rmode(h, 0) == 2
    a = np.random.default_rng(0).integers(0, q, N, dtype=np.uint64)
    b = a.copy()
    emu.L.emu_ntt_forward(h, 0, vp(b))
    assert not np.array_equal(a, b)
    emu.L.emu_ntt_inverse(h, 0, vp(b))
    assert np.array_equal(a, b)


@pytest.mark.parametrize("n_valid,scale", [(4096, 1.0), (4096, 100.0), (1, 1.0), (1000, 0.05), (4095, 7.0)])
def test_encode_matches_oracle(emu, emu_ctx, oracles, n_valid, scale):
    o, h = oracles["code"], emu_ctx["code"]
    x = (np.random.default_rng(n_valid).random(n_valid) - 0.5) * 2 * scale
    m = np.empty(N, np.int64)
    emu.L.emu_encode(h, vp(x), n_valid, vp(m))
    assert np.array_equal(m, o.encode_coeffs(x))


def test_sampler_matches_oracle_and_distribution(emu, emu_ctx, oracles):
    o, h = oracles["pybind"], emu_ctx["pybind"]
    x = np.zeros(4096)
    vs, es = [], []
    for ct in (0, 1, 2**32 + 5):
        m_e0, ve = np.empty(N, np.int64), np.empty(N, np.uint8)
        emu.L.emu_encode_sample(h, vp(x), 4096, 99, ct, vp(m_e0), vp(ve))
        v, e0, e1 = o.sample_noise(99, ct)
        assert np.array_equal(m_e0, e0)                      # x = 0 -> m = 0
        assert np.array_equal((ve & 3).astype(np.int64) - 1, v)          # pack_ve: bits 1..0 = v + 1, bits 7..2 = e1 + 32
        assert np.array_equal((ve >> 2).astype(np.int64) - 32, e1)
        vs.append(v); es += [e0, e1]
    v, e = np.concatenate(vs), np.concatenate(es)
    assert set(np.unique(v)) == {-1, 0, 1}
    assert abs(np.mean(v == 0) - 1 / 3) < 0.02
    assert abs(e.std() - 3.19) < 0.08 and abs(e.mean()) < 0.08 and np.abs(e).max() <= 31


def test_encrypt_decrypt_match_oracle(emu, emu_ctx, oracles, fixture_name):
    o, h = oracles[fixture_name], emu_ctx[fixture_name]
    x = np.random.default_rng(3).normal(0, 0.05, 4096)
    m_e0, ve = np.empty(N, np.int64), np.empty(N, np.uint8)
    emu.L.emu_encode_sample(h, vp(x), 4096, 1234, 7, vp(m_e0), vp(ve))
    b, a = o.public_key()
    ct = np.empty((2, 2, N), np.uint64)
    emu.L.emu_encrypt(h, vp(m_e0), vp(ve), vp(b), vp(a), vp(ct))
    ref = o.encrypt_ct(x, 1234, 7)
    assert np.array_equal(ct, ref)
    s = o.secret_key()
    sf = o.scaling_factor
    out_ref, hi_ref, lo_ref = o.decrypt_ct(ref, sf, want_coeffs=True)
    out, hi, lo = np.zeros(4096), np.empty(N, np.int64), np.empty(N, np.uint64)
    emu.L.emu_decrypt(h, vp(ct), vp(s), math.pow(sf, -1) * math.pow(2, 52), math.pow(2, -52), vp(out), 4096, vp(hi),
                      vp(lo), None)
    assert np.array_equal(hi, hi_ref) and np.array_equal(lo, lo_ref)
    assert np.array_equal(out, out_ref)
    assert np.abs(out - x).max() < 1e-9


def test_weighted_sum_arithmetic_matches_oracle(emu, emu_ctx, oracles):
    o, h = oracles["pybind"], emu_ctx["pybind"]
    rng = np.random.default_rng(11)
    for l in range(2):
        q = o.modulus(l)
        k = 16
        ins = [rng.integers(0, q, 512, dtype=np.uint64) for _ in range(k)]
        ins[0][:] = q - 1
        w = [np.float32(v) for v in rng.random(k)]
        w[0] = np.float32(1.0)
        c = np.array([o.weight_constant(float(v)) % q for v in w], dtype=np.uint64)
        ptrs = (C.c_void_p * k)(*[a.ctypes.data for a in ins])
        out = np.empty(512, np.uint64)
        emu.L.emu_weighted_sum(h, l, ptrs, vp(c), k, vp(out), 512)
        ref = [sum(int(ci) * int(a[j]) for ci, a in zip(c, ins)) % q for j in range(512)]
        assert out.tolist() == ref


# ---- ring dimensions above one NTT tile, chains deeper than two limbs, slot counts above one FFT tile ----
DEEP = ["n16k_l4", "n32k_l8", "n8k_l3", "n16k_sparse"]


@pytest.mark.parametrize("name", DEEP)
def test_deep_ntt_matches_oracle(emu, deep_emu_ctx, deep_oracles, name):
    o, h = deep_oracles[name], deep_emu_ctx[name]
    n = o.ring_dim
    rng = np.random.default_rng(17)
    for l in sorted({0, 1, o.num_limbs - 1}):
        q = o.modulus(l)
        for a in (rng.integers(0, q, n, dtype=np.uint64), np.full(n, q - 1, np.uint64)):
            f = a.copy(); emu.L.emu_ntt_forward(h, l, vp(f))
            assert np.array_equal(f, o.ntt_forward(l, a))
            b = a.copy(); emu.L.emu_ntt_inverse(h, l, vp(b))
            assert np.array_equal(b, o.ntt_inverse(l, a))


@pytest.mark.parametrize("name", DEEP)
def test_deep_encode_matches_oracle(emu, deep_emu_ctx, deep_oracles, name):
    o, h = deep_oracles[name], deep_emu_ctx[name]
    n, slots = o.ring_dim, o.slots
    for n_valid, scale in ((slots, 1.0), (slots - 3, 50.0), (1, 0.01)):
        x = (np.random.default_rng(n_valid).random(n_valid) - 0.5) * 2 * scale
        m = np.full(n, -7, np.int64)
        emu.L.emu_encode(h, vp(x), n_valid, vp(m))
        assert np.array_equal(m, o.encode_coeffs(x))


@pytest.mark.parametrize("name", DEEP)
def test_deep_encrypt_decrypt_match_oracle(emu, deep_emu_ctx, deep_oracles, name, oracle_mod):
    o, h = deep_oracles[name], deep_emu_ctx[name]
    n, slots, L, bits = o.ring_dim, o.slots, o.num_limbs, o.scaleFactorBits
    x = np.random.default_rng(5).normal(0, 0.05, slots)
    m_e0, ve = np.empty(n, np.int64), np.empty(n, np.uint8)
    emu.L.emu_encode_sample(h, vp(x), slots, 4321, 9, vp(m_e0), vp(ve))
    v, e0, e1 = o.sample_noise(4321, 9)
    assert np.array_equal(m_e0, o.encode_coeffs(x) + e0)
    assert np.array_equal((ve & 3).astype(np.int64) - 1, v) and np.array_equal((ve >> 2).astype(np.int64) - 32, e1)
    b, a = o.public_key()
    ct = np.empty((2, L, n), np.uint64)
    emu.L.emu_encrypt(h, vp(m_e0), vp(ve), vp(b), vp(a), vp(ct))
    ref = o.encrypt_ct(x, 4321, 9)
    assert np.array_equal(ct, ref)
    s = o.secret_key()
    sf = o.scaling_factor
    out_ref, words_ref = o.decrypt_ct_words(ref, sf)
    out, words = np.zeros(slots), np.empty((n, 8), np.uint64)
    emu.L.emu_decrypt(h, vp(ct), vp(s), math.pow(sf, -1) * math.pow(2, bits), math.pow(2, -bits), vp(out), slots, None,
                      None, vp(words))
    assert np.array_equal(words, words_ref)
    assert np.array_equal(out, out_ref)
    assert np.abs(out - x).max() < 2.0 ** (24 - bits)   # fresh-encryption noise scales with 1 / Delta


def test_deep_parameter_rules_match_oracle(emu, oracle_mod):
    for bits, depth in ((52, 1), (52, 3), (52, 7), (40, 2), (45, 3), (30, 5), (59, 1), (14, 1)):
        n = emu.L.emu_select_ring_dim_ex(bits, depth)
        assert n == oracle_mod.select_ring_dim_ex(bits, depth)
        if n > 32768:
            continue
        q = np.zeros(8, np.uint64)
        L = emu.L.emu_modulus_chain(bits, depth, n, vp(q))
        assert [int(v) for v in q[:L]] == oracle_mod.modulus_chain(bits, depth, n)
    assert (emu.L.emu_select_ring_dim_ex(52, 3), emu.L.emu_select_ring_dim_ex(52, 7)) == (16384, 32768)


# ---- FP64-pipe forward transform (limbs with q <= 1.125 * 2^52): same bits as the integer path and the oracle ----
def test_fp64_path_selected_for_the_scaling_limb(emu, emu_ctx, deep_emu_ctx):
    h = emu_ctx["pybind"]
    assert (emu.L.emu_limb_fp64(h, 0), emu.L.emu_limb_fp64(h, 1)) == (0, 1)      # 60-bit q0: integer pipe
    assert [emu.L.emu_limb_fp64(deep_emu_ctx["n16k_l4"], l) for l in range(4)] == [0, 1, 1, 1]
    assert [emu.L.emu_limb_fp64(deep_emu_ctx["n32k_l8"], l) for l in range(8)] == [0] + [1] * 7
    assert [emu.L.emu_limb_fp64(deep_emu_ctx["n8k_l3"], l) for l in range(3)] == [0, 1, 1]


@pytest.mark.parametrize("bits", [14, 20, 33, 40, 45, 50, 51, 52])
def test_fp64_ntt_equals_integer_ntt_and_oracle(emu, oracle_mod, bits):
    """Forward transform on doubles vs the integer path, for scaling primes of every size the reference sweeps
    (code/benchmark_crypto.py:123-124) and adversarial inputs: all q-1, values around q/2, deltas, random."""
    M = 2 * N
    qs = [oracle_mod.first_prime(bits, M), oracle_mod.previous_prime(oracle_mod.first_prime(bits, M) + M * 64, M)]
    if bits == 52:   # the largest modulus the FP64 path admits (1.125 * 2^52): the lazy-stage bounds are tightest here
        qs.append(oracle_mod.previous_prime((9 << 49) + 1, M))
        assert (9 << 49) - M * 4096 < qs[-1] <= 9 << 49
    rng = np.random.default_rng(bits)
    for q in qs:
        psi = oracle_mod.min_root_of_unity(M, q)
        h = emu.context([q], [psi])
        assert emu.L.emu_limb_fp64(h, 0) == 1
        half = q // 2
        rows = [np.full(N, q - 1, np.uint64), np.full(N, half, np.uint64), np.full(N, half + 1, np.uint64),
                np.where(np.arange(N) % 2 == 0, half, half + 1).astype(np.uint64),
                np.eye(1, N, N - 1, dtype=np.uint64)[0] * np.uint64(q - 1)]
        rows += [rng.integers(0, q, N, dtype=np.uint64) for _ in range(6)]
        rows += [(half + rng.integers(-3, 4, N)).astype(np.uint64)]
        for a in rows:
            f = a.copy(); emu.L.emu_ntt_forward(h, 0, vp(f))
            emu.L.emu_set_fp64(h, 0)
            g = a.copy(); emu.L.emu_ntt_forward(h, 0, vp(g))
            emu.L.emu_set_fp64(h, 1)
            assert np.array_equal(f, g)
            assert f.max() < q
        emu.L.emu_ntt_inverse(h, 0, vp(f))
        assert np.array_equal(f, rows[-1])
        emu.L.emu_destroy(h)


@pytest.mark.parametrize("bits", [14, 33, 45, 51, 52])
def test_fp64_inverse_ntt_equals_integer_inverse_and_oracle(emu, oracle_mod, bits):
    """The inverse transform on doubles (decrypt, rescale) against the integer path and the oracle, on the same
    adversarial rows as the forward one; and forward followed by inverse is the identity."""
    M = 2 * N
    qs = [oracle_mod.first_prime(bits, M)]
    if bits == 52:
        qs.append(oracle_mod.previous_prime((9 << 49) + 1, M))      # the largest modulus the FP64 path admits
    rng = np.random.default_rng(100 + bits)
    for q in qs:
        psi = oracle_mod.min_root_of_unity(M, q)
        h = emu.context([q], [psi])
        half = q // 2
        rows = [np.full(N, q - 1, np.uint64), np.full(N, half, np.uint64), np.full(N, half + 1, np.uint64),
                np.where(np.arange(N) % 2 == 0, half, half + 1).astype(np.uint64), np.zeros(N, np.uint64),
                np.eye(1, N, 5, dtype=np.uint64)[0] * np.uint64(q - 1)]
        rows += [rng.integers(0, q, N, dtype=np.uint64) for _ in range(5)]
        o = None
        for a in rows:
            f = a.copy(); emu.L.emu_ntt_inverse(h, 0, vp(f))
            emu.L.emu_set_fp64(h, 0)
            g = a.copy(); emu.L.emu_ntt_inverse(h, 0, vp(g))
            emu.L.emu_set_fp64(h, 1)
            assert np.array_equal(f, g) and int(f.max()) < q
            back = f.copy(); emu.L.emu_ntt_forward(h, 0, vp(back))
            assert np.array_equal(back, a)
        emu.L.emu_destroy(h)


def test_fp64_inverse_on_the_fixture_and_deep_contexts(emu, emu_ctx, deep_emu_ctx, oracles, deep_oracles):
    for name, o in list(oracles.items()) + list(deep_oracles.items()):
        h = emu_ctx[name] if name in emu_ctx else deep_emu_ctx[name]
        n = o.ring_dim
        rng = np.random.default_rng(len(name))
        for l in range(o.num_limbs):
            a = rng.integers(0, o.modulus(l), n, dtype=np.uint64)
            f = a.copy(); emu.L.emu_ntt_inverse(h, l, vp(f))
            assert np.array_equal(f, o.ntt_inverse(l, a)), (name, l)


def test_fp64_path_rejected_above_its_exactness_bound(emu, oracle_mod):
    M = 2 * N
    for bits, want in ((53, 0), (59, 0)):
        q = oracle_mod.first_prime(bits, M)
        h = emu.context([q], [oracle_mod.min_root_of_unity(M, q)])
        assert emu.L.emu_limb_fp64(h, 0) == want
        emu.L.emu_destroy(h)


def test_fp64_encrypt_equals_integer_encrypt(emu, emu_ctx, oracles):
    o, h = oracles["code"], emu_ctx["code"]
    x = np.random.default_rng(31).normal(0, 30.0, 4096)          # large messages: m + e0 near 2^58
    m_e0, ve = np.empty(N, np.int64), np.empty(N, np.uint8)
    emu.L.emu_encode_sample(h, vp(x), 4096, 77, 3, vp(m_e0), vp(ve))
    b, a = o.public_key()
    ct1, ct2 = np.empty((2, 2, N), np.uint64), np.empty((2, 2, N), np.uint64)
    emu.L.emu_encrypt(h, vp(m_e0), vp(ve), vp(b), vp(a), vp(ct1))
    emu.L.emu_set_fp64(h, 0)
    emu.L.emu_encrypt(h, vp(m_e0), vp(ve), vp(b), vp(a), vp(ct2))
    emu.L.emu_set_fp64(h, 1)
    assert np.array_equal(ct1, ct2) and np.array_equal(ct1, o.encrypt_ct(x, 77, 3))


@pytest.mark.parametrize("which", ["code", "n16k_l4"])
def test_selection_gather_on_encode_and_scatter_on_decode(emu, emu_ctx, deep_emu_ctx, oracles, deep_oracles, which):
    """Selective encryption: with an index list the encode load reads x[sel[i]] and the decode store writes
    out[sel[i]]; everything in between is the dense path, so the result equals the oracle's on the gathered copy."""
    o = oracles[which] if which in oracles else deep_oracles[which]
    h = emu_ctx[which] if which in emu_ctx else deep_emu_ctx[which]
    n, slots, L, bits = o.ring_dim, o.slots, o.num_limbs, o.scaleFactorBits
    rng = np.random.default_rng(21)
    x = rng.normal(0, 0.05, 3 * slots)
    n_valid = slots - 5
    sel = rng.permutation(x.size)[:n_valid].astype(np.uint32)
    m = np.empty(n, np.int64)
    emu.L.emu_set_selection(vp(sel))
    try:
        emu.L.emu_encode(h, vp(x), n_valid, vp(m))
        assert np.array_equal(m, o.encode_coeffs(x[sel]))
        ct = o.encrypt_ct(x[sel], 5, 0)
        s, sf = o.secret_key(), o.scaling_factor
        out = np.full(x.size, 7.0)
        emu.L.emu_decrypt(h, vp(ct), vp(s), math.pow(sf, -1) * math.pow(2, bits), math.pow(2, -bits), vp(out), n_valid,
                          None, None, None)
    finally:
        emu.L.emu_set_selection(None)
    ref = o.decrypt_ct(ct, sf)[:n_valid] if L <= 2 else o.decrypt_ct_words(ct, sf)[0][:n_valid]
    assert np.array_equal(out[sel], ref)
    untouched = np.ones(x.size, bool)
    untouched[sel] = False
    assert np.all(out[untouched] == 7.0)
