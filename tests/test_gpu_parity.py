"""Parity of the CUDA path, called through the C ABI, against the CPU oracle on identical inputs.

Tiers: bit-exact for every integer / byte stage (NTT, encrypt with the shared sampler
spec, weighted average, wire bytes, the CRT integers of decrypt); decoded doubles are compared both
exactly against the oracle (same unfused FFT order) and within RTOL = 1e-6 relative against the plaintext
weighted average — the tolerance BASELINE.json's north_star states for PALISADE's own decrypt.
"""
import ctypes as C
import gc
import hashlib
import math

import numpy as np
import pytest

from conftest import FIXTURE_DIRS, GOLDEN

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
from gpu_util import check, dev_u64, host_u64, ptr, stream  # noqa: E402

RTOL = 1e-6
N = 8192


def rel_err(a, b):
    return np.abs(a - b).max() / max(np.abs(b).max(), 1e-3)


def test_context_loaded(gpu_ckks, oracles, fixture_name):
    h, o = gpu_ckks[fixture_name], oracles[fixture_name]
    assert (h.ring_dim, h.num_limbs, h.slots) == (o.ring_dim, o.num_limbs, o.slots)
    assert [h.modulus(l) for l in range(2)] == [o.modulus(l) for l in range(2)]
    assert [h.root_of_unity(l) for l in range(2)] == [o.psi(l) for l in range(2)]
    assert h.key_tag() == o.key_tag() and h.scaling_factor == o.scaling_factor
    for w in (0.5, 0.2, 0.3, 1.0, 0.1, -0.25, 1e-3):
        assert h.weight_constant(w) == o.weight_constant(w)


@pytest.mark.parametrize("rows", [2, 6, 1000])
def test_ntt_bit_exact(shelfi, gpu_ckks, oracles, rows):
    h, o = gpu_ckks["pybind"], oracles["pybind"]
    rng = np.random.default_rng(rows)
    a = np.empty((rows, N), np.uint64)
    for r in range(rows):
        a[r] = rng.integers(0, o.modulus(r % 2), N, dtype=np.uint64)
    a[0] = o.modulus(0) - 1
    a[1] = o.modulus(1) - 1
    d = dev_u64(a)
    check(h, shelfi.lib().shelfi_ntt_forward_device(h.handle, ptr(d), rows, stream()))
    f = host_u64(d)
    for r in list(range(min(rows, 6))) + [rows - 1]:
        assert np.array_equal(f[r], o.ntt_forward(r % 2, a[r])), r
    check(h, shelfi.lib().shelfi_ntt_inverse_device(h.handle, ptr(d), rows, stream()))
    assert np.array_equal(host_u64(d), a)            # round trip on every row
    d = dev_u64(a)
    check(h, shelfi.lib().shelfi_ntt_inverse_device(h.handle, ptr(d), rows, stream()))
    b = host_u64(d)
    for r in range(min(rows, 4)):
        assert np.array_equal(b[r], o.ntt_inverse(r % 2, a[r])), r


def test_lane_exchange_ntt_variant_is_bit_identical(shelfi, gpu_ckks):
    """The measured-and-rejected variant of the forward transform (last stage by warp shuffle) gives the same words."""
    h = gpu_ckks["pybind"]
    rows = 64
    rng = np.random.default_rng(5)
    a = np.stack([rng.integers(0, h.modulus(r % 2), N, dtype=np.uint64) for r in range(rows)])
    a[0] = h.modulus(0) - 1
    d1, d2 = dev_u64(a), dev_u64(a)
    check(h, shelfi.lib().shelfi_ntt_forward_device(h.handle, ptr(d1), rows, stream()))
    check(h, shelfi.lib().shelfi_ntt_forward_shfl_device(h.handle, ptr(d2), rows, stream()))
    assert torch.equal(d1, d2)


@pytest.mark.parametrize("n,scale", [(4096, 1.0), (10000, 100.0), (1, 1.0), (4097, 0.05)])
def test_encode_bit_exact(shelfi, gpu_ckks, oracles, n, scale):
    h, o = gpu_ckks["code"], oracles["code"]
    x = (np.random.default_rng(n).random(n) - 0.5) * 2 * scale
    n_ct = -(-n // 4096)
    dx = torch.from_numpy(x).cuda()
    dm = torch.empty((n_ct, N), dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_encode_device(h.handle, ptr(dx), n, ptr(dm), stream()))
    m = dm.cpu().numpy()
    for i in range(n_ct):
        assert np.array_equal(m[i], o.encode_coeffs(x[i * 4096:(i + 1) * 4096])), i


@pytest.mark.parametrize("n", [1, 4096, 4097, 10000, 100000])
def test_encrypt_blob_bit_exact(gpu_ckks, oracles, fixture_name, n):
    """Same sampler spec on both sides => the serialized bytes are identical (ckks.cpp:61-104)."""
    h, o = gpu_ckks[fixture_name], oracles[fixture_name]
    x = np.random.default_rng(n).normal(0, 1.0, n)
    h.set_sampler_seed(4242)
    blob = h.encrypt(x)
    h.set_sampler_seed(None)
    ref = o.encrypt(x, seed=4242)
    assert len(blob) == len(ref) == h.blob_size(h.num_ciphertexts(n))
    assert hashlib.sha256(blob).digest() == hashlib.sha256(ref).digest()


def test_encrypt_device_sharded_draws_the_same_stream(shelfi, gpu_ckks, oracles):
    """A shard that starts at ciphertext 3 produces ciphertexts 3.. of the unsharded call."""
    h, o = gpu_ckks["pybind"], oracles["pybind"]
    x = np.random.default_rng(0).random(5 * 4096)
    h.set_sampler_seed(77)
    dx = torch.from_numpy(x[3 * 4096:]).cuda()
    dct = torch.empty((2, 2, 2, N), dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_encrypt_device(h.handle, ptr(dx), 2 * 4096, 3, ptr(dct), stream()))
    h.set_sampler_seed(None)
    got = host_u64(dct)
    for j in range(2):
        assert np.array_equal(got[j], o.encrypt_ct(x[(3 + j) * 4096:(4 + j) * 4096], 77, 3 + j))


def test_encrypt_many_equals_separate_calls(shelfi, gpu_ckks, oracles):
    """One batched launch over k learners == k single-learner launches == the oracle, bit for bit."""
    h, o = gpu_ckks["pybind"], oracles["pybind"]
    rng = np.random.default_rng(12)
    k, n = 3, 2 * 4096 + 100
    xs = [rng.normal(0, 0.05, n) for _ in range(k)]
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dct = [torch.empty((3, 2, 2, N), dtype=torch.int64, device="cuda") for _ in range(k)]
    xp = (C.c_void_p * k)(*[t.data_ptr() for t in dx])
    cp = (C.c_void_p * k)(*[t.data_ptr() for t in dct])
    h.set_sampler_seed(99)
    check(h, shelfi.lib().shelfi_encrypt_many_device(h.handle, xp, k, n, 0, cp, stream()))
    h.set_sampler_seed(None)
    for i in range(k):
        got = host_u64(dct[i])
        for j in range(3):
            assert np.array_equal(got[j], o.encrypt_ct(xs[i][j * 4096:(j + 1) * 4096], 99, j)), (i, j)


def test_unseeded_encryptions_differ_but_decrypt_equal(gpu_ckks):
    h = gpu_ckks["pybind"]
    x = np.random.default_rng(9).random(5000)
    b1, b2 = h.encrypt(x), h.encrypt(x)
    assert b1 != b2
    assert rel_err(h.decrypt(b1, 5000), x) < RTOL and rel_err(h.decrypt(b2, 5000), x) < RTOL


@pytest.mark.parametrize("k,n", [(3, 10000), (1, 4096), (20, 9000)])
def test_weighted_average_blob_bit_exact(gpu_ckks, oracles, fixture_name, k, n):
    """Tier 1: identical input blobs -> identical output blob (ckks.cpp:264-320)."""
    h, o = gpu_ckks[fixture_name], oracles[fixture_name]
    rng = np.random.default_rng(k * 1000 + n)
    w = [0.5, 0.2, 0.3] if k == 3 else list(rng.random(k) / k)
    blobs = [o.encrypt(rng.normal(0, 0.05, n), seed=100 + i) for i in range(k)]
    ref = o.computeWeightedAverage(blobs, w)
    got = h.computeWeightedAverage(blobs, w)
    assert hashlib.sha256(got).digest() == hashlib.sha256(ref).digest()


def test_weighted_sum_device_bit_exact(shelfi, gpu_ckks, oracles, oracle_mod):
    h, o = gpu_ckks["pybind"], oracles["pybind"]
    rng = np.random.default_rng(21)
    k, n = 19, 3 * 4096                                  # > 16 learners: two chained launches
    w = [float(np.float32(v)) for v in rng.random(k)]
    blobs = [o.encrypt(rng.random(n), seed=i) for i in range(k)]
    payloads = [oracle_mod.blob_parse(b)[1] for b in blobs]
    _, ref = oracle_mod.blob_parse(o.computeWeightedAverage(blobs, w))
    dev = [dev_u64(p) for p in payloads]
    ptrs = (C.c_void_p * k)(*[t.data_ptr() for t in dev])
    wa = (C.c_float * k)(*w)
    out = torch.empty_like(dev[0])
    check(h, shelfi.lib().shelfi_weighted_sum_device(h.handle, ptrs, wa, k, 3, ptr(out), stream()))
    assert np.array_equal(host_u64(out), ref)


@pytest.mark.parametrize("n", [1, 4096, 10000])
def test_decrypt_matches_oracle(shelfi, gpu_ckks, oracles, oracle_mod, fixture_name, n):
    h, o = gpu_ckks[fixture_name], oracles[fixture_name]
    rng = np.random.default_rng(n + 1)
    xs = [rng.normal(0, 1.0, n) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    fresh = o.encrypt(xs[0], seed=5)
    avg = o.computeWeightedAverage([o.encrypt(x, seed=6 + i) for i, x in enumerate(xs)], w)
    for blob, expect in ((fresh, xs[0]), (avg, sum(float(np.float32(wi)) * x for wi, x in zip(w, xs)))):
        got, ref = h.decrypt(blob, n), o.decrypt(blob, n)
        assert rel_err(got, ref) < RTOL                   # the tier north_star states
        assert np.array_equal(got, ref)                   # and in fact identical: same unfused FFT order
        assert rel_err(got, expect) < RTOL
    # integer stage: centred CRT value of every coefficient (bit-exact tier)
    info, payload = oracle_mod.blob_parse(avg)
    d = dev_u64(payload)
    hi = torch.empty((info.n_ct, N), dtype=torch.int64, device="cuda")
    lo = torch.empty((info.n_ct, N), dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_decrypt_integer_device(h.handle, ptr(d), info.n_ct, ptr(hi), ptr(lo), stream()))
    for j in range(info.n_ct):
        _, rhi, rlo = o.decrypt_ct(payload[j], info.scaling_factor, want_coeffs=True)
        assert np.array_equal(hi[j].cpu().numpy(), rhi) and np.array_equal(host_u64(lo[j]), rlo)


def test_pythonapi_example(gpu_ckks):
    """pythonApi/ckks_example.py:8-10,54-112 as an asserting test: 3 learners x 100,000, weights .5/.2/.3."""
    h = gpu_ckks["pybind"]
    rng = np.random.default_rng(1)
    n = 100000
    xs = [rng.random(n) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    blobs = [h.encrypt(x) for x in xs]
    out = h.decrypt(h.computeWeightedAverage(blobs, w), n)
    assert out.dtype == np.float64 and out.shape == (n,)
    assert rel_err(out, sum(float(np.float32(wi)) * x for wi, x in zip(w, xs))) < RTOL


def test_main_cpp_example(gpu_ckks):
    """src/main.cpp:8-23,28-78: 100 values in [0,100), three copies, weights .5/.3/.5 (sum 1.3)."""
    h = gpu_ckks["pybind"]
    x = np.random.default_rng(2).random(100) * 100
    blob = h.encrypt_cpp(x)
    out = h.decrypt_cpp(h.computeWeightedAverage_cpp([blob] * 3, [0.5, 0.3, 0.5]), 100)
    assert rel_err(out, x * (0.5 + float(np.float32(0.3)) + 0.5)) < RTOL


def test_float32_input_and_str_blobs(gpu_ckks):
    """benchmark.py passes float32 tensors (function_helper.py:7-26); pybind11 2.6.1 also takes str blobs."""
    h = gpu_ckks["code"]
    x = np.random.default_rng(3).normal(0, 0.05, (37, 11)).astype(np.float32)
    blob = h.encrypt(x)
    out = h.decrypt(blob.decode("latin-1"), x.size)
    assert rel_err(out, x.ravel().astype(np.float64)) < RTOL


def test_pack_unpack_device_roundtrip(shelfi, gpu_ckks, oracles, oracle_mod):
    h, o = gpu_ckks["code"], oracles["code"]
    blob = o.encrypt(np.random.default_rng(4).random(3 * 4096), seed=1)
    _, payload = oracle_mod.blob_parse(blob)
    size = h.blob_size(3)
    image = torch.zeros(((size + 7) // 8 + 1,), dtype=torch.int64, device="cuda")
    d = dev_u64(payload)
    check(h, shelfi.lib().shelfi_pack_device(h.handle, ptr(d), 3, ptr(image), stream()))
    back = torch.zeros_like(d)
    check(h, shelfi.lib().shelfi_unpack_device(h.handle, ptr(image), 3, ptr(back), stream()))
    assert np.array_equal(host_u64(back), payload)
    raw = image.cpu().numpy().tobytes()[:size]
    info, _ = oracle_mod.blob_parse(blob, want_payload=False)
    off = info.first_data_offset
    assert raw[off:off + 8 * N] == blob[off:off + 8 * N]


def test_error_paths(shelfi, gpu_ckks):
    h = gpu_ckks["pybind"]
    with pytest.raises(shelfi.ShelfiError):                    # n == 0 is UB in the reference (ckks.cpp:66,95)
        h.encrypt(np.zeros(0))
    with pytest.raises(shelfi.ShelfiError) as e:
        h.decrypt(b"\x01garbage", 10)
    assert e.value.code == -3
    blob = h.encrypt(np.ones(5000))
    with pytest.raises(shelfi.ShelfiError):
        h.decrypt(blob, 3 * 4096 + 1)                          # more values than two ciphertexts hold
    with pytest.raises(shelfi.ShelfiError):
        h.decrypt(blob[:-7], 5000)
    assert h.computeWeightedAverage([blob], [0.5, 0.5]) == b""  # ckks.cpp:265-268
    with pytest.raises(shelfi.ShelfiError) as e:               # a ciphertext of another context (q1 = 2^50 + ...)
        ct = (GOLDEN / "mkhe" / "CT1.txt").read_bytes()
        vec = b"\x01" + (1).to_bytes(8, "little") + ct[1:]
        h.decrypt(vec, 10)
    assert e.value.code == -4
    other = shelfi.CKKS("ckks", 4096, 52, "/nonexistent/dir/", device=0)
    other.loadCryptoParams()
    assert other.load_status == 7
    with pytest.raises(shelfi.ShelfiError) as e:
        other.encrypt(np.ones(4))
    assert e.value.code == -2


def test_keygen_matches_oracle_and_roundtrips(shelfi, oracle_mod, tmp_path):
    """ckks.cpp:25-59 on the device engine; with a fixed seed the three files equal the oracle's, byte for byte."""
    for bits, batch in ((52, 4096), (40, 2048), (20, 1024)):     # code/benchmark_crypto.py:123-124
        d_gpu, d_cpu = tmp_path / f"gpu{bits}", tmp_path / f"cpu{bits}"
        d_gpu.mkdir(); d_cpu.mkdir()
        g = shelfi.CKKS("ckks", batch, bits, str(d_gpu) + "/", device=0)
        g.set_sampler_seed(31337)
        assert g.genCryptoContextAndKeyGen() == 1
        assert g.ring_dim == 0                                   # generating does not load (ckks.cpp:28)
        o = oracle_mod.Oracle("ckks", batch, bits, str(d_cpu) + "/")
        assert o.genCryptoContextAndKeyGen(seed=31337, nine_int=True) == 1
        for name in ("cryptocontext.txt", "key-public.txt", "key-private.txt"):
            assert (d_gpu / name).read_bytes() == (d_cpu / name).read_bytes(), (bits, name)
        g.set_sampler_seed(None)
        g.loadCryptoParams()
        assert g.load_status == 0 and g.slots == batch
        x = np.random.default_rng(bits).normal(0, 0.05, 3 * batch + 17)
        w = [0.25, 0.75]
        out = g.decrypt(g.computeWeightedAverage([g.encrypt(x), g.encrypt(-x)], w), x.size)
        tol = max(RTOL, 2.0 ** (18 - bits))                      # params_results.csv: low bit counts are lossy
        assert np.abs(out - (0.25 * x - 0.75 * x)).max() < tol * max(1.0, np.abs(x).max())
        assert o.loadCryptoParams() == 0                         # sparse packing parity (batch < N/2)
        g.set_sampler_seed(5)
        assert g.encrypt(x) == o.encrypt(x, seed=5)
        g.set_sampler_seed(None)


def test_golden_vectors(gpu_ckks):
    """Committed oracle outputs (tests/golden/make_vectors.py): the GPU path reproduces them."""
    for name in ("pybind", "code"):
        v = np.load(GOLDEN / f"vectors_{name}.npz")
        h = gpu_ckks[name]
        h.set_sampler_seed(int(v["seed"]))
        blobs = [h.encrypt(x) for x in v["inputs"]]
        h.set_sampler_seed(None)
        assert [hashlib.sha256(b).hexdigest() for b in blobs] == [str(s) for s in v["encrypt_sha256"]]
        avg = h.computeWeightedAverage(blobs, [float(t) for t in v["weights"]])
        assert hashlib.sha256(avg).hexdigest() == str(v["wavg_sha256"])
        out = h.decrypt(avg, v["inputs"].shape[1])
        assert np.array_equal(out, v["decrypted"])


def test_full_size_properties(gpu_ckks):
    """BASELINE config 2 size (10 learners x 1,663,370 params): round trip, linearity, checksum of checksums."""
    h = gpu_ckks["pybind"]
    P, K = 1663370, 10
    rng = np.random.default_rng(2)
    xs = [rng.normal(0, 0.05, P).astype(np.float32).astype(np.float64) for _ in range(K)]
    w = [1.0 / K] * K
    blobs = [h.encrypt(x) for x in xs]
    n_ct = h.num_ciphertexts(P)
    assert n_ct == 407 and all(len(b) == h.blob_size(n_ct) for b in blobs)
    avg = h.computeWeightedAverage(blobs, w)
    out = h.decrypt(avg, P)
    expect = sum(float(np.float32(wi)) * x for wi, x in zip(w, xs))
    assert rel_err(out, expect) < RTOL
    # linearity: wavg of (A, B) with weights (1, 1) decrypts to dec(A) + dec(B)
    s = h.decrypt(h.computeWeightedAverage(blobs[:2], [1.0, 1.0]), P)
    assert rel_err(s, xs[0] + xs[1]) < RTOL
    # order independence of the accumulation, bit for bit
    a1 = h.computeWeightedAverage(blobs[:3], [0.5, 0.25, 0.25])
    a2 = h.computeWeightedAverage([blobs[2], blobs[0], blobs[1]], [0.25, 0.5, 0.25])
    assert a1 == a2


# ---- layer-wise callers: one call for all state_dict entries (benchmark.py:490-493, benchmark_selection.py:152-193) ----
CNN_LAYERS = [800, 32, 51200, 64, 1605632, 512, 5120, 10]      # CNN_OriginalFedAvg (benchmark.py:152-219)


def test_layers_api_equals_per_layer_calls(gpu_ckks, oracles):
    """encrypt_layers / computeWeightedAverage_layers / decrypt_layers == the reference's per-layer loop, byte for byte."""
    h, o = gpu_ckks["pybind"], oracles["pybind"]
    sizes = [800, 32, 4096, 4097, 1, 9000]
    rng = np.random.default_rng(8)
    K, w = 3, [0.5, 0.2, 0.3]
    models = [[rng.normal(0, 0.05, n) for n in sizes] for _ in range(K)]
    enc = []
    for i, layers in enumerate(models):
        h.set_sampler_seed(40 + i)
        blobs = h.encrypt_layers(layers)
        single = [h.encrypt(a) for a in layers]
        h.set_sampler_seed(None)
        assert blobs == single
        assert blobs == [o.encrypt(a, seed=40 + i) for a in layers]
        enc.append(blobs)
    avg = h.computeWeightedAverage_layers(enc, w)
    assert avg == [h.computeWeightedAverage([enc[i][s] for i in range(K)], w) for s in range(len(sizes))]
    assert avg == [o.computeWeightedAverage([enc[i][s] for i in range(K)], w) for s in range(len(sizes))]
    dec = h.decrypt_layers(avg, sizes)
    for s, n in enumerate(sizes):
        assert np.array_equal(dec[s], h.decrypt(avg[s], n)) and np.array_equal(dec[s], o.decrypt(avg[s], n))
        expect = sum(float(np.float32(wi)) * models[i][s] for i, wi in enumerate(w))
        assert rel_err(dec[s], expect) < RTOL
    assert h.computeWeightedAverage_layers(enc, w[:2]) == []        # ckks.cpp:265-268
    with pytest.raises(Exception):
        h.decrypt_layers(avg, [n + 5000 for n in sizes])


def test_layers_api_full_cnn(gpu_ckks):
    """The 8 state_dict entries of the config-2 CNN, 10 learners: 412 ciphertexts per learner instead of 407."""
    h = gpu_ckks["code"]
    rng = np.random.default_rng(6)
    K = 10
    models = [[rng.normal(0, 0.05, n).astype(np.float32) for n in CNN_LAYERS] for _ in range(K)]
    enc = [h.encrypt_layers(m) for m in models]
    assert sum(h.num_ciphertexts(n) for n in CNN_LAYERS) == 412
    avg = h.computeWeightedAverage_layers(enc, [1.0 / K] * K)
    dec = h.decrypt_layers(avg, CNN_LAYERS)
    for s in range(len(CNN_LAYERS)):
        expect = sum(np.float64(np.float32(1.0 / K)) * models[i][s].astype(np.float64) for i in range(K))
        assert rel_err(dec[s], expect) < RTOL


def test_pinned_blob_mode_equals_bytes_mode(shelfi, gpu_ckks):
    """CKKS(..., pinned=True) returns bytes-like PinnedBlob objects (page-locked host memory, pooled): same bytes,
    accepted wherever a blob is, and interchangeable with real bytes."""
    h = gpu_ckks["pybind"]
    d = str(FIXTURE_DIRS["pybind"]) + "/"
    p = shelfi.CKKS("ckks", 4096, 52, d, device=0, pinned=True)
    p.loadCryptoParams()
    rng = np.random.default_rng(14)
    xs = [rng.normal(0, 0.05, 9000) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    blobs_b, blobs_p = [], []
    for i, x in enumerate(xs):
        for hh, out in ((h, blobs_b), (p, blobs_p)):
            hh.set_sampler_seed(70 + i)
            out.append(hh.encrypt(x))
            hh.set_sampler_seed(None)
    assert all(isinstance(b, shelfi.PinnedBlob) for b in blobs_p) and all(isinstance(b, bytes) for b in blobs_b)
    assert [bytes(b) for b in blobs_p] == blobs_b and blobs_p[0] == blobs_b[0] and len(blobs_p[0]) == len(blobs_b[0])
    assert bytes(memoryview(blobs_p[1])) == blobs_b[1]           # buffer protocol (what sock.sendall / f.write use)
    import copy
    import pickle
    assert pickle.loads(pickle.dumps(blobs_p[2])) == blobs_b[2] and copy.deepcopy({"k": blobs_p[0]})["k"] == blobs_b[0]
    avg_b = h.computeWeightedAverage(blobs_b, w)
    avg_p = p.computeWeightedAverage(blobs_p, w)
    assert bytes(avg_p) == avg_b == bytes(p.computeWeightedAverage([blobs_p[0], blobs_b[1], memoryview(blobs_b[2])], w))
    assert np.array_equal(p.decrypt(avg_p, 9000), h.decrypt(avg_b, 9000))
    assert np.array_equal(h.decrypt(avg_p, 9000), h.decrypt(avg_b, 9000))
    layers = [xs[0][:100], xs[1][:5000], xs[2]]
    p.set_sampler_seed(3); h.set_sampler_seed(3)
    lp, lb = p.encrypt_layers(layers), h.encrypt_layers(layers)
    p.set_sampler_seed(None); h.set_sampler_seed(None)
    assert [bytes(b) for b in lp] == lb
    ap = p.computeWeightedAverage_layers([lp, lp], [0.5, 0.5])
    assert [bytes(b) for b in ap] == h.computeWeightedAverage_layers([lb, lb], [0.5, 0.5])
    for got, want in zip(p.decrypt_layers(ap, [100, 5000, 9000]), layers):
        assert rel_err(got, want) < RTOL
    del blobs_p, lp, ap, avg_p                                   # buffers go back to the pool and are reused
    again = p.encrypt(xs[0])
    assert rel_err(p.decrypt(again, 9000), xs[0]) < RTOL


def test_fp64_pipe_ntt_equals_integer_pipe(shelfi, gpu_ckks, monkeypatch):
    """The scaling limb's forward transform runs on the FP64 pipe (exact double arithmetic); a handle loaded with
    SHELFI_FP64_NTT=0 keeps it on the integer pipe.  Same bits on 4000 rows and on whole encryptions."""
    d = str(FIXTURE_DIRS["pybind"]) + "/"
    monkeypatch.setenv("SHELFI_FP64_NTT", "0")
    hi = shelfi.CKKS("ckks", 4096, 52, d, device=0)
    hi.loadCryptoParams()
    monkeypatch.delenv("SHELFI_FP64_NTT")
    hf = gpu_ckks["pybind"]
    rows = 4000
    rng = np.random.default_rng(77)
    q = [hf.modulus(0), hf.modulus(1)]
    a = np.empty((rows, N), np.uint64)
    for r in range(rows):
        a[r] = rng.integers(0, q[r % 2], N, dtype=np.uint64)
    a[1] = q[1] - 1
    a[3] = q[1] // 2
    a[5] = q[1] // 2 + 1
    a[7, ::2] = q[1] - 1
    d1, d2 = dev_u64(a), dev_u64(a)
    check(hf, shelfi.lib().shelfi_ntt_forward_device(hf.handle, ptr(d1), rows, stream()))
    check(hi, shelfi.lib().shelfi_ntt_forward_device(hi.handle, ptr(d2), rows, stream()))
    assert torch.equal(d1, d2)
    # ... and back: the inverse transform of the scaling limb runs on the FP64 pipe as well (round 2)
    e1, e2 = dev_u64(a), dev_u64(a)
    check(hf, shelfi.lib().shelfi_ntt_inverse_device(hf.handle, ptr(e1), rows, stream()))
    check(hi, shelfi.lib().shelfi_ntt_inverse_device(hi.handle, ptr(e2), rows, stream()))
    assert torch.equal(e1, e2)
    check(hf, shelfi.lib().shelfi_ntt_inverse_device(hf.handle, ptr(d1), rows, stream()))
    assert np.array_equal(host_u64(d1).reshape(rows, N), a)
    x = rng.normal(0, 20.0, 300000)
    hf.set_sampler_seed(5); hi.set_sampler_seed(5)
    b1, b2 = hf.encrypt(x), hi.encrypt(x)
    hf.set_sampler_seed(None)
    assert hashlib.sha256(b1).digest() == hashlib.sha256(b2).digest()


def test_one_launch_decrypt_equals_the_two_stages(shelfi, gpu_ckks, monkeypatch):
    """From 256 ciphertexts the inverse transforms and the decode of a ciphertext run in one launch
    (decrypt_fused_kernel); a handle made under SHELFI_DECRYPT_UNFUSED=1 keeps the two kernels.  Same doubles, bit for
    bit, on a ragged vector of 300 ciphertexts, on a short one (both take the two-kernel route) and with a selection."""
    d = str(FIXTURE_DIRS["pybind"]) + "/"
    monkeypatch.setenv("SHELFI_DECRYPT_UNFUSED", "1")
    h2 = shelfi.CKKS("ckks", 4096, 52, d, device=0)
    h2.loadCryptoParams()
    monkeypatch.delenv("SHELFI_DECRYPT_UNFUSED")
    h1 = gpu_ckks["pybind"]
    L = shelfi.lib()
    rng = np.random.default_rng(41)
    for n in (299 * 4096 + 1234, 3 * 4096 + 5):
        x = rng.normal(0, 3.0, n)
        n_ct = h1.num_ciphertexts(n)
        dct = torch.empty(n_ct * h1.ciphertext_words, dtype=torch.int64, device="cuda")
        check(h1, L.shelfi_encrypt_device(h1.handle, ptr(torch.from_numpy(x).cuda()), n, 0, ptr(dct), stream()))
        outs = []
        for h in (h1, h2):
            o = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
            check(h, L.shelfi_decrypt_device(h.handle, ptr(dct), n_ct, h.scaling_factor, ptr(o), n, stream()))
            torch.cuda.synchronize()
            outs.append(o.cpu().numpy())
        assert np.array_equal(outs[0].view(np.uint64), outs[1].view(np.uint64)), n
        assert np.abs(outs[0] - x).max() < 1e-6
    blob = h1.encrypt(x := rng.normal(0, 1.0, 280 * 4096))
    assert np.array_equal(np.asarray(h1.decrypt(blob, x.size)).view(np.uint64), np.asarray(h2.decrypt(blob, x.size)).view(np.uint64))


def test_pipelined_host_paths_equal_the_plain_ones(shelfi, gpu_ckks):
    """Vectors of >= 128 / 256 ciphertexts into page-locked results take the group-pipelined encrypt and the split
    weighted average (copy back while the next part computes / arrives): same bytes as the unsplit pageable path."""
    h = gpu_ckks["pybind"]
    d = str(FIXTURE_DIRS["pybind"]) + "/"
    p = shelfi.CKKS("ckks", 4096, 52, d, device=0, pinned=True)
    p.loadCryptoParams()
    rng = np.random.default_rng(15)
    for n in (260 * 4096 - 7, 515 * 4096 + 1, 130 * 4096):
        xs = [rng.normal(0, 0.05, n) for _ in range(3)]
        w = [0.5, 0.2, 0.3]
        bb, bp = [], []
        for i, x in enumerate(xs):
            for hh, out in ((h, bb), (p, bp)):
                hh.set_sampler_seed(80 + i)
                out.append(hh.encrypt(x))
                hh.set_sampler_seed(None)
        for a, b in zip(bb, bp):
            assert len(a) == len(b) and np.array_equal(np.frombuffer(a, np.uint8), b.array)
        avg_b, avg_p = h.computeWeightedAverage(bb, w), p.computeWeightedAverage(bp, w)
        assert np.array_equal(np.frombuffer(avg_b, np.uint8), avg_p.array)
        got = p.decrypt(avg_p, n)
        assert rel_err(got, sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, xs))) < RTOL
        # pinned mode hands back an ndarray over pooled page-locked memory: a view must keep that memory out of the pool
        assert isinstance(got, np.ndarray) and got.dtype == np.float64 and got.shape == (n,)
        view, want_view = got[5:15], got[5:15].copy()
        del got
        gc.collect()
        other = p.decrypt(bp[0], n)                      # different values; would land in the same buffer if it were free
        assert np.array_equal(view, want_view) and not np.array_equal(other[5:15], want_view)
        got = p.decrypt(avg_p, n)
        del view, other
        # page-locked blob AND page-locked result: the group-pipelined decrypt (upload / decrypt / copy back overlapped)
        pin_out = torch.full((n,), np.nan, dtype=torch.float64).pin_memory()
        lib = shelfi.lib()
        assert lib.shelfi_decrypt(p.handle, C.c_void_p(avg_p.array.ctypes.data), len(avg_p), n, C.c_void_p(pin_out.data_ptr())) == 0
        assert np.array_equal(pin_out.numpy(), h.decrypt(avg_b, n))
        assert np.array_equal(pin_out.numpy(), got)
        del bb, bp, avg_b, avg_p


def test_layers_with_many_distinct_sizes(shelfi, gpu_ckks, oracles):
    """More distinct layer sizes in one call than any plan cache holds, large ones (headers written by the device
    kernel) next to small ones (written by the host): every blob equals the per-layer call and the oracle's."""
    o = oracles["pybind"]
    d = str(FIXTURE_DIRS["pybind"]) + "/"
    p = shelfi.CKKS("ckks", 4096, 52, d, device=0, pinned=True)
    p.loadCryptoParams()
    h = gpu_ckks["pybind"]
    sizes = [4096 * j + 5 for j in range(1, 41)] + [7, 4096, 300]
    rng = np.random.default_rng(23)
    models = [[rng.normal(0, 0.05, n) for n in sizes] for _ in range(2)]
    enc = []
    for i, layers in enumerate(models):
        p.set_sampler_seed(500 + i)
        blobs = p.encrypt_layers(layers)
        p.set_sampler_seed(None)
        for s in (0, 17, 39, 40, 42):
            assert bytes(blobs[s]) == o.encrypt(layers[s], seed=500 + i), s
        h.set_sampler_seed(500 + i)
        assert [bytes(b) for b in blobs] == [h.encrypt(a) for a in layers]
        h.set_sampler_seed(None)
        enc.append(blobs)
    w = [0.25, 0.75]
    avg = p.computeWeightedAverage_layers(enc, w)
    for s in (0, 31, 39, 41):
        assert bytes(avg[s]) == o.computeWeightedAverage([bytes(enc[i][s]) for i in range(2)], w), s
    assert [bytes(b) for b in avg] == [h.computeWeightedAverage([bytes(enc[i][s]) for i in range(2)], w) for s in range(len(sizes))]
    dec = p.decrypt_layers(avg, sizes)
    for s, n in enumerate(sizes):
        assert rel_err(dec[s], sum(float(np.float32(wi)) * models[i][s] for i, wi in enumerate(w))) < RTOL
