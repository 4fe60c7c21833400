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


def test_keygen_draws_follow_the_spec(emu, emu_ctx, oracles, oracle_mod):
    """s, e under the secret test key and a under the public one, drawn by the device code (replayed), against the
    spec restated with the oracle's block function; s ternary, e Gaussian(3.19), a below q and well spread."""
    o, h = oracles["pybind"], emu_ctx["pybind"]
    n, L, seed = o.ring_dim, o.num_limbs, 424242
    s, e, a = np.empty(n, np.int64), np.empty(n, np.int64), np.empty((L, n), np.uint64)
    emu.L.emu_keygen_draws(h, seed, vp(s), vp(e), vp(a))
    ks, kp = oracle_mod.seed_key(seed, 1), oracle_mod.seed_key(seed, 2)
    blk = oracle_mod.chacha20_block(ks, 3, [0, 0, 2])
    assert np.array_equal(s[48:64], ((blk.astype(np.uint64) * 3) >> 32).astype(np.int64) - 1)
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
