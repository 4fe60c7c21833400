"""The device sampler on the GPU: RFC 8439 known answers through the C ABI, key hygiene of the engine."""
from __future__ import annotations

import ctypes as C

import numpy as np
import pytest

from test_sampler_cpu import RFC_BLOCK, RFC_KEY, RFC_NONCE, ZERO_BLOCK

pytestmark = pytest.mark.gpu


def test_device_chacha20_block_matches_rfc8439(gpu_ckks, shelfi):
    h = gpu_ckks["pybind"]
    L = shelfi.lib()
    for key, ctr, nonce, want in ((RFC_KEY, 1, RFC_NONCE, RFC_BLOCK),
                                  (np.zeros(8, np.uint32), 0, np.zeros(3, np.uint32), ZERO_BLOCK)):
        key, nonce = np.ascontiguousarray(key), np.ascontiguousarray(nonce)
        out = np.zeros(16, np.uint32)
        assert L.shelfi_chacha20_block_device(h.handle, key.ctypes.data, ctr, nonce.ctypes.data, out.ctypes.data) == 0
        assert np.array_equal(out, want)


def test_unseeded_calls_draw_fresh_keys(gpu_ckks):
    """Without a test seed every encrypt call takes a new 256-bit key from the OS: same input, different ciphertexts,
    same decryption; a seeded call is reproducible and differs from both."""
    h = gpu_ckks["code"]
    x = np.random.default_rng(0).normal(0, 0.05, 5000)
    a, b = h.encrypt(x), h.encrypt(x)
    assert a != b and len(a) == len(b)
    h.set_sampler_seed(1)
    c, d = h.encrypt(x), h.encrypt(x)
    h.set_sampler_seed(None)
    assert c == d and c != a
    for blob in (a, b, c):
        assert np.abs(h.decrypt(blob, x.size) - x).max() < 1e-9


def test_generated_keys_use_independent_public_and_secret_streams(shelfi, tmp_path, oracle_mod):
    """genCryptoContextAndKeyGen without a seed: two runs give unrelated keys, each key pair satisfies
    INTT(b + a s) = small e (the K2 relation of the reference's fixtures), and the files load and round-trip."""
    tags = []
    for i in range(2):
        d = str(tmp_path / f"k{i}") + "/"
        (tmp_path / f"k{i}").mkdir()
        g = shelfi.CKKS("ckks", 4096, 52, d, device=0)
        assert g.genCryptoContextAndKeyGen() == 1
        o = oracle_mod.Oracle("ckks", 4096, 52, d)
        assert o.loadCryptoParams() == 0
        b, a = o.public_key()
        s = o.secret_key()
        for l in range(o.num_limbs):
            q = o.modulus(l)
            t = (b[l].astype(object) + a[l].astype(object) * s[l].astype(object)) % q
            e = o.ntt_inverse(l, np.array(t, dtype=np.uint64)).astype(object)
            e = np.where(e > q // 2, e - q, e)
            assert max(abs(int(v)) for v in e) <= 31
            sc = o.ntt_inverse(l, s[l]).astype(object)
            sc = np.where(sc > q // 2, sc - q, sc)
            assert set(int(v) for v in sc) <= {-1, 0, 1}
        tags.append(o.key_tag())
        g.loadCryptoParams()
        x = np.random.default_rng(i).random(6000)
        assert np.abs(g.decrypt(g.encrypt(x), x.size) - x).max() < 1e-9
    assert tags[0] != tags[1] and all(len(t) == 32 for t in tags)
