"""Parity of the CUDA path on contexts beyond the reference's hard-coded multDepth = 1 (ckks.cpp:26):
ring dimensions 16384 / 32768 (2 / 4 NTT tiles per limb polynomial), chains of 3-8 limbs (Garner CRT in
decode), slot counts above one FFT tile and sparse packing in a large ring — BASELINE configs 3 and 4.
Same tiers as test_gpu_parity.py: integers, bytes and (because the FFT order is the reference's) decoded
doubles bit-exact against the CPU oracle; decoded weighted averages within RTOL = 1e-6 of the plaintext.
"""
import ctypes as C
import hashlib

import numpy as np
import pytest

from conftest import DEEP_CONTEXTS

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
from gpu_util import check, dev_u64, host_u64, ptr, stream  # noqa: E402

RTOL = 1e-6
DEEP = list(DEEP_CONTEXTS)


def rel_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-3)


@pytest.fixture(scope="module")
def gpu_deep(shelfi, deep_oracles):
    out = {}
    for name, o in deep_oracles.items():
        h = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0)
        h.loadCryptoParams()
        assert h.load_status == 0, name
        out[name] = h
    return out


@pytest.mark.parametrize("name", DEEP)
def test_context_loaded(gpu_deep, deep_oracles, name):
    h, o = gpu_deep[name], deep_oracles[name]
    L = o.num_limbs
    assert (h.ring_dim, h.num_limbs, h.slots) == (o.ring_dim, L, o.slots)
    assert [h.modulus(l) for l in range(L)] == [o.modulus(l) for l in range(L)]
    assert [h.root_of_unity(l) for l in range(L)] == [o.psi(l) for l in range(L)]
    assert h.key_tag() == o.key_tag() and h.scaling_factor == o.scaling_factor
    assert h.ciphertext_words == 2 * L * o.ring_dim
    for w in (0.5, 0.1, -0.25):
        assert h.weight_constant(w) == o.weight_constant(w)


@pytest.mark.parametrize("name", DEEP)
def test_ntt_bit_exact(shelfi, gpu_deep, deep_oracles, name):
    h, o = gpu_deep[name], deep_oracles[name]
    n, L = o.ring_dim, o.num_limbs
    rows = 3 * L + 1
    rng = np.random.default_rng(rows)
    a = np.empty((rows, n), np.uint64)
    for r in range(rows):
        a[r] = rng.integers(0, o.modulus(r % L), n, dtype=np.uint64)
    a[0] = o.modulus(0) - 1
    d = dev_u64(a)
    check(h, shelfi.lib().shelfi_ntt_forward_device(h.handle, ptr(d), rows, stream()))
    f = host_u64(d)
    for r in range(rows):
        assert np.array_equal(f[r], o.ntt_forward(r % L, a[r])), r
    check(h, shelfi.lib().shelfi_ntt_inverse_device(h.handle, ptr(d), rows, stream()))
    assert np.array_equal(host_u64(d), a)
    d = dev_u64(a)
    check(h, shelfi.lib().shelfi_ntt_inverse_device(h.handle, ptr(d), rows, stream()))
    b = host_u64(d)
    for r in range(L + 1):
        assert np.array_equal(b[r], o.ntt_inverse(r % L, a[r])), r


@pytest.mark.parametrize("name", DEEP)
def test_encode_bit_exact(shelfi, gpu_deep, deep_oracles, name):
    h, o = gpu_deep[name], deep_oracles[name]
    n_ring, batch = o.ring_dim, o.batchSize
    for n, scale in ((batch, 1.0), (2 * batch + 5, 100.0), (1, 1.0)):
        x = (np.random.default_rng(n).random(n) - 0.5) * 2 * scale
        n_ct = -(-n // batch)
        dx = torch.from_numpy(x).cuda()
        dm = torch.full((n_ct, n_ring), -3, dtype=torch.int64, device="cuda")
        check(h, shelfi.lib().shelfi_encode_device(h.handle, ptr(dx), n, ptr(dm), stream()))
        m = dm.cpu().numpy()
        for i in range(n_ct):
            assert np.array_equal(m[i], o.encode_coeffs(x[i * batch:(i + 1) * batch])), (n, i)


@pytest.mark.parametrize("name", DEEP)
def test_encrypt_wavg_decrypt_bit_exact(shelfi, gpu_deep, deep_oracles, oracle_mod, name):
    """encrypt blob, weighted-average blob, CRT integers and decoded doubles all equal the oracle's."""
    h, o = gpu_deep[name], deep_oracles[name]
    batch, n_ring, L = o.batchSize, o.ring_dim, o.num_limbs
    n = 2 * batch + 7
    rng = np.random.default_rng(11)
    xs = [rng.normal(0, 0.05, n) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    blobs = []
    for i, x in enumerate(xs):
        h.set_sampler_seed(900 + i)
        blob = h.encrypt(x)
        h.set_sampler_seed(None)
        ref = o.encrypt(x, seed=900 + i)
        assert len(blob) == len(ref) == h.blob_size(3)
        assert hashlib.sha256(blob).digest() == hashlib.sha256(ref).digest(), i
        blobs.append(blob)
    avg = h.computeWeightedAverage(blobs, w)
    assert hashlib.sha256(avg).digest() == hashlib.sha256(o.computeWeightedAverage(blobs, w)).digest()
    expect = sum(float(np.float32(wi)) * x for wi, x in zip(w, xs))
    for blob, want in ((blobs[0], xs[0]), (avg, expect)):
        got, ref = h.decrypt(blob, n), o.decrypt(blob, n)
        assert np.array_equal(got, ref)
        tol = max(RTOL, 2.0 ** (26 - o.scaleFactorBits))       # fresh noise / Delta dominates at small scale bits
        assert rel_err(got, want) < tol
    info, payload = oracle_mod.blob_parse(avg)
    d = dev_u64(payload)
    words = torch.empty((info.n_ct, n_ring, 8), dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_decrypt_words_device(h.handle, ptr(d), info.n_ct, ptr(words), stream()))
    got_words = host_u64(words)
    for j in range(info.n_ct):
        _, ref_words = o.decrypt_ct_words(payload[j], info.scaling_factor)
        assert np.array_equal(got_words[j], ref_words), j
    if L <= 2:
        return
    with pytest.raises(AssertionError):                        # the 128-bit tap is for 1-2 limbs
        hi = torch.empty((info.n_ct, n_ring), dtype=torch.int64, device="cuda")
        check(h, shelfi.lib().shelfi_decrypt_integer_device(h.handle, ptr(d), info.n_ct, ptr(hi), ptr(hi), stream()))


@pytest.mark.parametrize("name", DEEP)
def test_device_api_matches_blob_api(shelfi, gpu_deep, deep_oracles, name):
    """encrypt_many / weighted_sum / decrypt on dense device buffers == the host-blob calls."""
    h, o = gpu_deep[name], deep_oracles[name]
    batch, n_ring, L = o.batchSize, o.ring_dim, o.num_limbs
    k, n_ct = 3, 2
    n = n_ct * batch - 11
    rng = np.random.default_rng(4)
    xs = [rng.normal(0, 0.05, n) for _ in range(k)]
    w = [0.25, 0.5, 0.25]
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dct = [torch.empty((n_ct, 2, L, n_ring), dtype=torch.int64, device="cuda") for _ in range(k)]
    xp = (C.c_void_p * k)(*[t.data_ptr() for t in dx])
    cp = (C.c_void_p * k)(*[t.data_ptr() for t in dct])
    h.set_sampler_seed(55)
    check(h, shelfi.lib().shelfi_encrypt_many_device(h.handle, xp, k, n, 0, cp, stream()))
    h.set_sampler_seed(None)
    for i in range(k):
        got = host_u64(dct[i])
        for j in range(n_ct):
            assert np.array_equal(got[j], o.encrypt_ct(xs[i][j * batch:(j + 1) * batch], 55, j)), (i, j)
    out = torch.empty_like(dct[0])
    wa = (C.c_float * k)(*w)
    check(h, shelfi.lib().shelfi_weighted_sum_device(h.handle, cp, wa, k, n_ct, ptr(out), stream()))
    dec = torch.empty(n, dtype=torch.float64, device="cuda")
    sf = o.scaling_factor * o.scaling_factor
    check(h, shelfi.lib().shelfi_decrypt_device(h.handle, ptr(out), n_ct, sf, ptr(dec), n, stream()))
    expect = sum(float(np.float32(wi)) * x for wi, x in zip(w, xs))
    assert rel_err(dec.cpu().numpy(), expect) < max(RTOL, 2.0 ** (26 - o.scaleFactorBits))


def test_keygen_ex_matches_oracle(shelfi, oracle_mod, tmp_path):
    """genCryptoContextAndKeyGen with multDepth 3 / 7 (configs 3 / 4): files equal the oracle's under one seed."""
    for bits, batch, depth, want in ((52, 8192, 3, (16384, 4)), (52, 16384, 7, (32768, 8)), (40, 2048, 2, (16384, 3))):
        d_gpu, d_cpu = tmp_path / f"gpu{depth}", tmp_path / f"cpu{depth}"
        d_gpu.mkdir(); d_cpu.mkdir()
        g = shelfi.CKKS("ckks", batch, bits, str(d_gpu) + "/", device=0)
        g.set_sampler_seed(271828)
        assert g.genCryptoContextAndKeyGen(mult_depth=depth) == 1
        o = oracle_mod.Oracle("ckks", batch, bits, str(d_cpu) + "/")
        assert o.genCryptoContextAndKeyGen(seed=271828, nine_int=True, mult_depth=depth) == 1
        for fname in ("cryptocontext.txt", "key-public.txt", "key-private.txt"):
            assert (d_gpu / fname).read_bytes() == (d_cpu / fname).read_bytes(), (depth, fname)
        g.set_sampler_seed(None)
        g.loadCryptoParams()
        assert g.load_status == 0 and (g.ring_dim, g.num_limbs) == want
        x = np.random.default_rng(depth).normal(0, 0.05, batch + 3)
        out = g.decrypt(g.computeWeightedAverage([g.encrypt(x), g.encrypt(-x)], [0.25, 0.75]), x.size)
        assert np.abs(out - (-0.5 * x)).max() < max(RTOL, 2.0 ** (26 - bits))


def test_config3_size_properties(shelfi, tmp_path):
    """BASELINE config 3 shape at reduced learner count: N = 16384, 4 limbs, ResNet-18-sized vector
    (11,689,512 params -> 1,427 ciphertexts), 3 learners: round trip and linearity through the blob API."""
    d = tmp_path / "c3"
    d.mkdir()
    h = shelfi.CKKS("ckks", 8192, 52, str(d) + "/", device=0)
    assert h.genCryptoContextAndKeyGen(mult_depth=3) == 1
    h.loadCryptoParams()
    P, K = 11689512, 3
    rng = np.random.default_rng(3)
    xs = [rng.normal(0, 0.05, P).astype(np.float32).astype(np.float64) for _ in range(K)]
    blobs = [h.encrypt(x) for x in xs]
    assert h.num_ciphertexts(P) == 1427 and len(blobs[0]) == h.blob_size(1427)
    w = [0.5, 0.25, 0.25]
    out = h.decrypt(h.computeWeightedAverage(blobs, w), P)
    assert rel_err(out, sum(wi * x for wi, x in zip(w, xs))) < RTOL
    a1 = h.computeWeightedAverage(blobs, w)
    a2 = h.computeWeightedAverage([blobs[2], blobs[0], blobs[1]], [0.25, 0.5, 0.25])
    assert a1 == a2


def test_fp64_pipe_equals_integer_pipe_on_deep_contexts(shelfi, gpu_deep, deep_oracles, monkeypatch):
    """Multi-tile rings: head loader and tile passes on doubles vs the integer path (SHELFI_FP64_NTT=0), bit for bit."""
    for name in ("n16k_l4", "n32k_l8"):
        o, hf = deep_oracles[name], gpu_deep[name]
        monkeypatch.setenv("SHELFI_FP64_NTT", "0")
        hi = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0)
        hi.loadCryptoParams()
        monkeypatch.delenv("SHELFI_FP64_NTT")
        x = np.random.default_rng(2).normal(0, 10.0, 5 * o.batchSize + 3)
        hf.set_sampler_seed(8); hi.set_sampler_seed(8)
        b1, b2 = hf.encrypt(x), hi.encrypt(x)
        hf.set_sampler_seed(None)
        assert hashlib.sha256(b1).digest() == hashlib.sha256(b2).digest(), name


def test_random_soak_against_oracle():
    """tools/soak_gpu.py: random (context, learners, length, amplitude, weights) cases through the public API,
    blobs bit-exact and decoded doubles identical to the oracle; 12 contexts incl. sparse packing and 3 ring sizes."""
    import subprocess
    import sys
    from pathlib import Path
    root = Path(__file__).resolve().parent.parent
    r = subprocess.run([sys.executable, str(root / "tools" / "soak_gpu.py"), "36", "3"], capture_output=True, text=True,
                       timeout=600)
    assert r.returncode == 0 and "soak ok: 36 cases" in r.stdout, r.stdout[-2000:] + r.stderr[-2000:]
