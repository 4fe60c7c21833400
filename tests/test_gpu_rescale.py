"""Rescale-before-transport on the GPU, through the C ABI, against the CPU oracle's restatement of
PALISADE's ModReduce under EXACTRESCALE (DCRTPoly::DropLastElementAndScale).

Tiers: the rescaled dense batch and the rescaled serialized blob are bit-exact; the centred CRT integers of the
rescaled ciphertexts' decryption are bit-exact; decoded doubles are identical to the oracle's and within RTOL = 1e-6
of the plaintext weighted average.  Opt-in: the reference itself leaves the average at depth 2 (ckks.cpp:287-303)."""
import ctypes as C
import hashlib

import numpy as np
import pytest

from conftest import DEEP_CONTEXTS

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
from gpu_util import check, dev_u64, host_u64, ptr, stream  # noqa: E402

RTOL = 1e-6


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


def _learners(o, n, k, seed, amp=0.05):
    rng = np.random.default_rng(seed)
    xs = [rng.normal(0, amp, n) for _ in range(k)]
    w = [float(v) for v in rng.dirichlet(np.ones(k)).astype(np.float32)]
    blobs = [o.encrypt(x, seed=70 + i) for i, x in enumerate(xs)]
    ref = sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, xs))
    return blobs, w, ref


@pytest.mark.parametrize("name", ["pybind", "code"])
def test_rescaled_average_blob_bit_exact(gpu_ckks, oracles, oracle_mod, name):
    h, o = gpu_ckks[name], oracles[name]
    for n, k in ((1, 2), (4096, 3), (3 * 4096 + 17, 5)):
        blobs, w, ref = _learners(o, n, k, seed=n)
        want = o.rescale(o.computeWeightedAverage(blobs, w))
        got = h.computeWeightedAverage(blobs, w, rescale=True)
        assert len(got) == len(want)
        assert hashlib.sha256(got).digest() == hashlib.sha256(want).digest(), (n, k)
        # the two-step form gives the same bytes
        assert h.rescaleCiphertexts(h.computeWeightedAverage(blobs, w)) == want
        info, _ = oracle_mod.blob_parse(bytes(got), want_payload=False)
        assert (info.n_limbs, info.depth, info.level) == (1, 1, 1)
        out = h.decrypt(got, n)
        assert np.array_equal(out, o.decrypt(want, n))          # same doubles as the oracle
        assert rel_err(out, ref) < RTOL
        assert rel_err(out, h.decrypt(h.computeWeightedAverage(blobs, w), n)) < 1e-9
    assert len(got) < 0.53 * len(h.computeWeightedAverage(blobs, w))


def test_rescale_is_off_by_default_and_settable_per_handle(shelfi, gpu_ckks, oracles):
    o = oracles["pybind"]
    blobs, w, _ = _learners(o, 5000, 2, seed=1)
    plain = gpu_ckks["pybind"].computeWeightedAverage(blobs, w)
    assert plain == o.computeWeightedAverage(blobs, w)           # the reference's depth-2 result, untouched
    h = shelfi.CKKS("ckks", 4096, 52, o.cryptodir, device=0, rescale=True)
    h.loadCryptoParams()
    assert h.computeWeightedAverage(blobs, w) == o.rescale(plain)
    assert h.computeWeightedAverage(blobs, w, rescale=False) == plain


def test_rescale_error_paths(shelfi, gpu_ckks, oracles):
    h, o = gpu_ckks["pybind"], oracles["pybind"]
    blobs, w, _ = _learners(o, 100, 2, seed=2)
    fresh = blobs[0]
    with pytest.raises(shelfi.ShelfiError):                      # depth 1: nothing to rescale
        h.rescaleCiphertexts(fresh)
    r = h.computeWeightedAverage(blobs, w, rescale=True)
    with pytest.raises(shelfi.ShelfiError):                      # one tower left
        h.rescaleCiphertexts(r)
    with pytest.raises(shelfi.ShelfiError):                      # rescaled inputs are not fresh ciphertexts
        h.computeWeightedAverage([r, r], w)
    with pytest.raises(shelfi.ShelfiError):
        h.rescaleCiphertexts(bytes(r)[:-3])


@pytest.mark.parametrize("name", ["pybind", "code"])
def test_dense_rescale_and_crt_integers_bit_exact(shelfi, gpu_ckks, oracles, oracle_mod, name):
    h, o = gpu_ckks[name], oracles[name]
    N = o.ring_dim
    blobs, w, ref = _learners(o, 2 * 4096 + 9, 3, seed=3, amp=10.0)
    _, payload = oracle_mod.blob_parse(o.computeWeightedAverage(blobs, w))
    n_ct = payload.shape[0]
    rng = np.random.default_rng(4)
    q1 = o.modulus(1)
    payload[2, 0, 1, :4] = [0, q1 - 1, q1 >> 1, (q1 >> 1) + 1]  # the centring boundary of SwitchModulus
    payload[2, 1, 0] = rng.integers(0, o.modulus(0), N, dtype=np.uint64)
    d_in = dev_u64(payload)
    d_out = torch.full((n_ct, 2, 1, N), -1, dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_rescale_device(h.handle, ptr(d_in), n_ct, 0, ptr(d_out), stream()))
    got = host_u64(d_out)
    want = np.stack([o.rescale_ct(payload[j]) for j in range(n_ct)])
    assert np.array_equal(got, want)
    assert np.array_equal(host_u64(d_in), payload)               # the input batch is not touched
    # decrypt of the rescaled batch: centred CRT integers, then the doubles
    sf = o.scaling_factor * o.scaling_factor / q1
    d_words = torch.zeros((n_ct, N, 8), dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_decrypt_words_level_device(h.handle, ptr(d_out), n_ct, 1, ptr(d_words), stream()))
    words = host_u64(d_words)
    n = (n_ct - 1) * 4096 + 9
    d_vals = torch.zeros(n, dtype=torch.float64, device="cuda")
    check(h, shelfi.lib().shelfi_decrypt_level_device(h.handle, ptr(d_out), n_ct, 1, sf, ptr(d_vals), n, stream()))
    vals = d_vals.cpu().numpy()
    for j in range(n_ct):
        dv, dw = o.decrypt_ct_level(want[j], sf)
        assert np.array_equal(words[j], dw), j
        lo, hi = j * 4096, min(n, (j + 1) * 4096)
        assert np.array_equal(vals[lo:hi], dv[:hi - lo]), j
    assert rel_err(vals[:2 * 4096], ref[:2 * 4096]) < RTOL


@pytest.mark.parametrize("name", list(DEEP_CONTEXTS))
def test_deep_contexts_bit_exact(shelfi, gpu_deep, deep_oracles, oracle_mod, name):
    """Rings of 2 and 4 tiles (head stages / cluster head / tail pass), 3-8 towers, FP64- and integer-pipe limbs."""
    h, o = gpu_deep[name], deep_oracles[name]
    L, N, batch = o.num_limbs, o.ring_dim, o.batchSize
    n = batch + 3
    blobs, w, ref = _learners(o, n, 2, seed=5)
    avg = o.computeWeightedAverage(blobs, w)
    want = o.rescale(avg)
    got = h.computeWeightedAverage(blobs, w, rescale=True)
    assert hashlib.sha256(got).digest() == hashlib.sha256(want).digest()
    out = h.decrypt(got, n)
    assert np.array_equal(out, o.decrypt(want, n))
    assert rel_err(out, ref) < RTOL
    # dense form on random residues, integer pipe against FP64 pipe is covered by the oracle equality on both
    rng = np.random.default_rng(6)
    ct = np.stack([np.stack([np.stack([rng.integers(0, o.modulus(l), N, dtype=np.uint64) for l in range(L)])
                             for _ in range(2)]) for _ in range(3)])
    d_in = dev_u64(ct)
    d_out = torch.zeros((3, 2, L - 1, N), dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_rescale_device(h.handle, ptr(d_in), 3, L, ptr(d_out), stream()))
    got = host_u64(d_out)
    for j in range(3):
        assert np.array_equal(got[j], o.rescale_ct(ct[j])), j
    # a second level: drop one more tower of the already rescaled batch (L >= 3)
    if L >= 3:
        d_out2 = torch.zeros((3, 2, L - 2, N), dtype=torch.int64, device="cuda")
        check(h, shelfi.lib().shelfi_rescale_device(h.handle, ptr(d_out), 3, L - 1, ptr(d_out2), stream()))
        got2 = host_u64(d_out2)
        for j in range(3):
            assert np.array_equal(got2[j], o.rescale_ct(got[j])), j


def test_config2_size_property(shelfi, gpu_ckks):
    """BASELINE configs[1] size (10 learners x 407 ciphertexts), device-resident: decrypt(rescale(sum)) equals
    decrypt(sum) to 1e-9 and the plaintext average to RTOL — no oracle at this size."""
    h = gpu_ckks["pybind"]
    lib = shelfi.lib()
    K, P = 10, 1_663_370
    n_ct = h.num_ciphertexts(P)
    words = h.ciphertext_words
    rng = np.random.default_rng(2)
    xs = [torch.from_numpy(rng.normal(0, 0.05, P).astype(np.float32).astype(np.float64)).cuda() for _ in range(K)]
    cts = [torch.empty(n_ct * words, dtype=torch.int64, device="cuda") for _ in range(K)]
    xp = (C.c_void_p * K)(*[x.data_ptr() for x in xs])
    cp = (C.c_void_p * K)(*[c.data_ptr() for c in cts])
    check(h, lib.shelfi_encrypt_many_device(h.handle, xp, K, P, 0, cp, stream()))
    w = (C.c_float * K)(*[1.0 / K] * K)
    acc = torch.empty(n_ct * words, dtype=torch.int64, device="cuda")
    check(h, lib.shelfi_weighted_sum_device(h.handle, cp, w, K, n_ct, ptr(acc), stream()))
    red = torch.empty(n_ct * words // 2, dtype=torch.int64, device="cuda")
    check(h, lib.shelfi_rescale_device(h.handle, ptr(acc), n_ct, 0, ptr(red), stream()))
    sf2 = h.scaling_factor * h.scaling_factor
    a = torch.empty(P, dtype=torch.float64, device="cuda")
    b = torch.empty(P, dtype=torch.float64, device="cuda")
    check(h, lib.shelfi_decrypt_device(h.handle, ptr(acc), n_ct, sf2, ptr(a), P, stream()))
    check(h, lib.shelfi_decrypt_level_device(h.handle, ptr(red), n_ct, 1, sf2 / h.modulus(1), ptr(b), P, stream()))
    torch.cuda.synchronize()
    ref = sum(float(np.float32(1.0 / K)) * x for x in xs)
    assert float((a - b).abs().max()) < 1e-9
    assert float((b - ref).abs().max()) / max(float(ref.abs().max()), 1e-3) < RTOL


def test_layerwise_rescaled_average_equals_per_layer_calls(shelfi, gpu_ckks, oracles):
    """shelfi_compute_weighted_average_layers_rescaled[_into]: blob s == the single-vector rescaled call on layer s ==
    the oracle's ModReduce of the oracle's average; decrypt_layers reads the shorter blobs."""
    o = oracles["pybind"]
    d = str(o.cryptodir)
    sizes = [800, 32, 4096, 4097, 1, 9000]
    rng = np.random.default_rng(18)
    K, w = 3, [0.5, 0.2, 0.3]
    models = [[rng.normal(0, 0.05, n) for n in sizes] for _ in range(K)]
    enc = [[o.encrypt(a, seed=300 + 10 * i + s) for s, a in enumerate(layers)] for i, layers in enumerate(models)]
    want = [o.rescale(o.computeWeightedAverage([enc[i][s] for i in range(K)], w)) for s in range(len(sizes))]
    for pinned in (False, True):   # False: the shim loops over layers; True: one C call for all layers
        h = shelfi.CKKS("ckks", 4096, 52, d, device=0, pinned=pinned)
        h.loadCryptoParams()
        assert h.load_status == 0
        got = h.computeWeightedAverage_layers(enc, w, rescale=True)
        assert [bytes(b) for b in got] == want
        assert [bytes(b) for b in got] == [bytes(h.computeWeightedAverage([enc[i][s] for i in range(K)], w, rescale=True))
                                           for s in range(len(sizes))]
        dec = h.decrypt_layers(got, sizes)
        for s, n in enumerate(sizes):
            assert np.array_equal(dec[s], o.decrypt(want[s], n))
            expect = sum(float(np.float32(wi)) * models[i][s] for i, wi in enumerate(w))
            assert rel_err(dec[s], expect) < RTOL
        # default stays the reference's depth-2 result
        plain = h.computeWeightedAverage_layers(enc, w)
        assert [bytes(b) for b in plain] == [o.computeWeightedAverage([enc[i][s] for i in range(K)], w) for s in range(len(sizes))]
    # the allocating C entry point
    lib = shelfi.lib()
    h = gpu_ckks["pybind"]
    flat = [enc[i][s] for i in range(K) for s in range(len(sizes))]
    n = len(sizes)
    ptrs = (C.c_void_p * (K * n))(*[C.cast(C.c_char_p(b), C.c_void_p).value for b in flat])
    lens_in = (C.c_uint64 * (K * n))(*[len(b) for b in flat])
    ww = (C.c_float * K)(*w)
    out, offs, lens, total = C.c_void_p(), (C.c_uint64 * n)(), (C.c_uint64 * n)(), C.c_uint64()
    check(h, lib.shelfi_compute_weighted_average_layers_rescaled(h.handle, ptrs, lens_in, ww, K, n, C.byref(out), offs, lens,
                                                                 C.byref(total)))
    try:
        for s in range(n):
            assert C.string_at(out.value + offs[s], lens[s]) == want[s]
    finally:
        lib.shelfi_free(out)
