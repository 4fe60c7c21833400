"""The sampler: ChaCha20 in counter mode (RFC 8439), 256-bit keys.

Known-answer tests against the RFC's own vectors for the two independent implementations (the oracle's C, written
from the RFC, and the device code's `chacha20_block` replayed on the CPU through tests/emu), the test-key layout, and
the key-generation draws.  The GPU run of the same function is tests/test_gpu_sampler.py.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

# RFC 8439 section 2.3.2: key 00 01 .. 1f, nonce 00 00 00 09 00 00 00 4a 00 00 00 00, block counter 1
RFC_KEY = np.frombuffer(bytes(range(32)), dtype="<u4")
RFC_NONCE = np.frombuffer(bytes.fromhex("000000090000004a00000000"), dtype="<u4")
RFC_BLOCK = np.array([0xe4e7f110, 0x15593bd1, 0x1fdd0f50, 0xc47120a3, 0xc7f4d1c7, 0x0368c033, 0x9aaa2204, 0x4e6cd4c3,
                      0x466482d2, 0x09aa9f07, 0x05d7c214, 0xa2028bd9, 0xd19c12b5, 0xb94e16de, 0xe883d0cb, 0x4e3c50a2],
                     dtype=np.uint32)
# RFC 8439 appendix A.1 test vector #1: all-zero key and nonce, counter 0
ZERO_BLOCK = np.frombuffer(bytes.fromhex(
    "76b8e0ada0f13d90405d6ae55386bd28bdd219b8a08ded1aa836efcc8b770dc7"
    "da41597c5157488d7724e03fb8d84a376a43b8f41518a11cc387b669b2ee6586"), dtype="<u4")


def vp(a):
    return a.ctypes.data_as(C.c_void_p)


def test_oracle_chacha20_matches_rfc8439(oracle_mod):
    assert np.array_equal(oracle_mod.chacha20_block(RFC_KEY, 1, RFC_NONCE), RFC_BLOCK)
    assert np.array_equal(oracle_mod.chacha20_block(np.zeros(8, np.uint32), 0, np.zeros(3, np.uint32)), ZERO_BLOCK)


def test_device_code_chacha20_matches_rfc8439(emu):
    for key, ctr, nonce, want in ((RFC_KEY, 1, RFC_NONCE, RFC_BLOCK),
                                  (np.zeros(8, np.uint32), 0, np.zeros(3, np.uint32), ZERO_BLOCK)):
        key, nonce = np.ascontiguousarray(key), np.ascontiguousarray(nonce)
        out = np.empty(16, np.uint32)
        emu.L.emu_chacha20_block(vp(key), ctr, vp(nonce), vp(out))
        assert np.array_equal(out, want)


def test_seed_key_layout_separates_domains(oracle_mod):
    k0, k1, k2 = (oracle_mod.seed_key(0x1122334455667788, d) for d in (0, 1, 2))
    assert k0[0] == 0x55667788 and k0[1] == 0x11223344
    assert bytes(k0[2:6].tobytes()) == b"SHELFI-B200 seed"
    assert k0[6] == 0 and k1[6] == 1 and k2[6] == 2
    blocks = [oracle_mod.chacha20_block(k, 0, [0, 0, 0]) for k in (k0, k1, k2)]
    assert not np.array_equal(blocks[0], blocks[1]) and not np.array_equal(blocks[1], blocks[2])


def _ternary_ref(h16, retry):
    for a in range(8):
        c = (h16 >> (2 * a)) & 3
        if c != 3:
            return c - 1
    return retry


def test_ternary_sampler_is_exact_and_takes_the_retry_path_only_after_eight_rejections(emu, oracle_mod):
    used = C.c_int()
    seen = {-1: 0, 0: 0, 1: 0}
    for h16 in range(1 << 16):                      # every possible half word
        got = emu.L.emu_ternary16(h16, 7, C.byref(used))
        assert got == _ternary_ref(h16, 7)
        assert used.value == (1 if h16 == 0xFFFF else 0)
        if got != 7:
            seen[got] += 1
    assert seen[-1] == seen[0] == seen[1] == (2**16 - 1) // 3        # exactly uniform over the accepted half words
    # the retry stream: first accepted 2-bit attempt of block j, the same in the device code and in the oracle
    key = oracle_mod.seed_key(5, 0)
    for j in (0, 1, 77, 8191):
        blk = oracle_mod.chacha20_block(key, j, [9, 0, 8])
        want = next(int((w >> (2 * a)) & 3) - 1 for w in blk for a in range(16) if (w >> (2 * a)) & 3 != 3)
        assert emu.L.emu_retry_ternary(vp(key), 9, 8, j) == want
        nonce = np.array([9, 0, 8], np.uint32)
        assert oracle_mod.lib().orc_ternary_from_half(0xFFFF, vp(key), vp(nonce), j) == want
        assert oracle_mod.lib().orc_ternary_from_half(0xFFF7, vp(key), vp(nonce), j) == 0       # attempts 0 = 3, 1 = 1


def test_lazy_gaussian_equals_the_full_precision_inversion(emu, emu_ctx, oracles):
    """The device decides from the top 31 bits and looks at the low word only on a tie with a table entry; the oracle
    always inverts the full 63-bit number.  Same value on every tie (all table entries, low words around the entry's
    own low word) and on random draws; the low word is requested exactly on ties."""
    o, h = oracles["pybind"], emu_ctx["pybind"]
    cdf = np.empty(32, np.uint64)
    emu.L.emu_gauss_table(h, vp(cdf))
    assert all(int(cdf[k]) <= int(cdf[k + 1]) for k in range(31)) and int(cdf[0]) < int(cdf[10])
    his = sorted({int(c) >> 32 for c in cdf[:31]})
    used = C.c_int()
    for k in range(31):
        hi, lo = int(cdf[k]) >> 32, int(cdf[k]) & 0xFFFFFFFF
        for sign in (0, 1):
            hword = (hi << 1) | sign
            for l in {lo, max(lo - 1, 0), min(lo + 1, 2**32 - 1), 0, 2**32 - 1}:
                got = emu.L.emu_gauss_lazy(h, hword, l, C.byref(used))
                assert used.value == 1
                assert got == o.gauss_from_words(hword, l), (k, sign, l)
    rng = np.random.default_rng(9)
    draws = rng.integers(0, 2**32, 20000, dtype=np.uint64)
    vals = []
    for hw in draws:
        hw = int(hw)
        got = emu.L.emu_gauss_lazy(h, hw, 0x12345678, C.byref(used))
        assert used.value == (1 if (hw >> 1) in his else 0)
        assert got == o.gauss_from_words(hw, 0x12345678)
        vals.append(got)
    vals = np.array(vals)
    assert abs(vals.std() - 3.19) < 0.1 and abs(vals.mean()) < 0.1 and np.abs(vals).max() <= 31


def test_keygen_draws_follow_the_spec(emu, emu_ctx, oracles, oracle_mod):
    """s, e under the secret test key and a under the public one, drawn by the device code (replayed), against the
    spec restated with the oracle's block function; s ternary, e Gaussian(3.19), a below q and well spread."""
    o, h = oracles["pybind"], emu_ctx["pybind"]
    n, L, seed = o.ring_dim, o.num_limbs, 424242
    s, e, a = np.empty(n, np.int64), np.empty(n, np.int64), np.empty((L, n), np.uint64)
    emu.L.emu_keygen_draws(h, seed, vp(s), vp(e), vp(a))
    ks, kp = oracle_mod.seed_key(seed, 1), oracle_mod.seed_key(seed, 2)
    blk = oracle_mod.chacha20_block(ks, 3, [0, 0, 2])                 # ternary block 3: coefficients 96..127
    halves = [(int(blk[i >> 1]) >> (16 * (i & 1))) & 0xFFFF for i in range(32)]
    assert [int(v) for v in s[96:128]] == [_ternary_ref(hh, None) for hh in halves]
    blk = oracle_mod.chacha20_block(ks, 5, [0, 0, 3])                 # Gaussian block 5: coefficients 80..95
    assert [int(v) for v in e[80:96]] == [o.gauss_from_words(int(w), 0) for w in blk]
    assert set(np.unique(s)) == {-1, 0, 1} and abs(np.mean(s == 0) - 1 / 3) < 0.03
    assert abs(e.std() - 3.19) < 0.15 and np.abs(e).max() <= 31
    for l in range(L):
        q = o.modulus(l)
        assert int(a[l].max()) < q
        bits = q.bit_length()
        j, t = 1234, 0
        while True:   # first accepted attempt for coefficient j of limb l
            blk = oracle_mod.chacha20_block(kp, j >> 3, [l, t, 4])
            r = (int(blk[2 * (j & 7) + 1]) << 32) | int(blk[2 * (j & 7)])
            cand = r >> (64 - bits)
            if cand < q:
                break
            t += 1
        assert int(a[l, j]) == cand
        assert abs(float(np.mean(a[l].astype(np.float64))) / q - 0.5) < 0.02
    # public and secret streams come from different keys: the same counters give unrelated words
    assert not np.array_equal(oracle_mod.chacha20_block(ks, 0, [0, 0, 2]), oracle_mod.chacha20_block(kp, 0, [0, 0, 2]))
