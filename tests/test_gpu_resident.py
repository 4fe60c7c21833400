"""Resident ciphertext vectors (CKKS(resident=True), shelfi_ct_*): the drop-in's three calls with the ciphertexts left in
HBM.  The bar: whatever a ResidentBlob materialises to is byte-identical to what the bytes-mode call returns and to the
CPU oracle's blob; decrypt gives the same doubles; nothing about the values depends on where the input lived (pageable,
page-locked, device; float64 or float32)."""
from __future__ import annotations

import hashlib
import pickle

import numpy as np
import pytest

from conftest import FIXTURE_DIRS

pytestmark = pytest.mark.gpu


def sha(b):
    return hashlib.sha256(bytes(b)).hexdigest()


@pytest.fixture(scope="module")
def res(shelfi, gpu_ckks):
    out = {}
    for name, d in FIXTURE_DIRS.items():
        r = shelfi.CKKS("ckks", 4096, 52, str(d) + "/", device=0, resident=True)
        r.loadCryptoParams()
        assert r.load_status == 0
        out[name] = r
    return out


@pytest.mark.parametrize("n", [1, 4096, 4097, 10000, 250_000])
def test_encrypt_materialises_to_the_oracle_blob(res, gpu_ckks, oracles, shelfi, fixture_name, n):
    r, h, o = res[fixture_name], gpu_ckks[fixture_name], oracles[fixture_name]
    x = np.random.default_rng(n).normal(0, 0.05, n)
    for g in (r, h):
        g.set_sampler_seed(1000 + n)
    try:
        rb, hb = r.encrypt(x), h.encrypt(x)
    finally:
        for g in (r, h):
            g.set_sampler_seed(None)
    assert isinstance(rb, shelfi.ResidentBlob) and isinstance(hb, bytes)
    assert len(rb) == len(hb)                      # known without producing the bytes
    assert rb._cache is None
    ref = o.encrypt(x, seed=1000 + n)
    assert sha(rb) == sha(hb) == sha(ref)
    assert rb == hb and hash(rb) == hash(hb)
    assert bytes(memoryview(rb)) == hb and pickle.loads(pickle.dumps(rb)) == hb
    assert np.array_equal(r.decrypt(rb, n), h.decrypt(hb, n))
    assert rb.n_ct == -(-n // 4096) and rb.depth == 1 and rb.level == 0 and rb.scaling_factor == h.scaling_factor


def test_input_placement_and_dtype_do_not_change_the_ciphertext(res, oracles, shelfi):
    import torch
    r, o = res["code"], oracles["code"]
    n = 3 * 4096 + 77
    x32 = np.random.default_rng(5).normal(0, 0.05, n).astype(np.float32)
    x64 = x32.astype(np.float64)
    ref = sha(o.encrypt(x64, seed=31))
    variants = {"pageable f64": x64, "pageable f32": x32, "pinned f64": shelfi.pinned_array(x64),
                "pinned f32": shelfi.pinned_array(x32), "device f64": torch.from_numpy(x64).cuda(),
                "device f32": torch.from_numpy(x32).cuda(), "list": list(x64[:10])}
    torch.cuda.synchronize()
    r.set_sampler_seed(31)
    try:
        for name, v in variants.items():
            if name == "list":
                assert sha(r.encrypt(v)) == sha(o.encrypt(x64[:10], seed=31)), name
            else:
                assert sha(r.encrypt(v)) == ref, name
    finally:
        r.set_sampler_seed(None)
    big = np.random.default_rng(6).normal(0, 0.05, 300 * 4096 + 5)     # above the staging threshold for pageable memory
    r.set_sampler_seed(32)
    try:
        a, b = r.encrypt(big), r.encrypt(shelfi.pinned_array(big))
        c = r.encrypt(big.astype(np.float32))
        d = r.encrypt(shelfi.pinned_array(big.astype(np.float32)))
    finally:
        r.set_sampler_seed(None)
    # (page-locked inputs of 256 ciphertexts and more go up and are encrypted in four groups on four streams)
    assert sha(a) == sha(b) == sha(o.encrypt(big, seed=32))
    assert sha(c) == sha(d) == sha(o.encrypt(big.astype(np.float32).astype(np.float64), seed=32))


@pytest.mark.parametrize("k,n", [(1, 5000), (3, 10000), (10, 40 * 4096 + 3), (20, 4096)])
def test_weighted_average_of_mixed_learner_types(res, gpu_ckks, oracles, shelfi, fixture_name, k, n):
    r, h, o = res[fixture_name], gpu_ckks[fixture_name], oracles[fixture_name]
    rng = np.random.default_rng(k * 1000 + n)
    xs = [rng.normal(0, 0.05, n) for _ in range(k)]
    w = list(rng.random(k) / k)
    blobs = [o.encrypt(x, seed=50 + i) for i, x in enumerate(xs)]
    ref = o.computeWeightedAverage(blobs, w)
    other = shelfi.CKKS("ckks", 4096, 52, str(FIXTURE_DIRS[fixture_name]) + "/", device=0, resident=True)
    other.loadCryptoParams()
    mixed = []
    for i, b in enumerate(blobs):          # resident handle, bytes, page-locked blob, a handle of ANOTHER CKKS object
        if i % 4 == 0:
            mixed.append(r.to_resident(b))
        elif i % 4 == 1:
            mixed.append(b)
        elif i % 4 == 2:
            p = shelfi.PinnedBlob(len(b))
            p.array[:] = np.frombuffer(b, np.uint8)
            mixed.append(p)
        else:
            mixed.append(other.to_resident(b))
    avg = r.computeWeightedAverage(mixed, w)
    assert isinstance(avg, shelfi.ResidentBlob) and avg.depth == 2
    assert sha(avg) == sha(ref) == sha(h.computeWeightedAverage(blobs, w))
    out = r.decrypt(avg, n)
    assert np.array_equal(out, o.decrypt(ref, n)) and np.array_equal(out, h.decrypt(ref, n))
    expect = sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, xs))
    assert np.abs(out - expect).max() / np.abs(expect).max() < 1e-6
    # a bytes-mode object given resident handles returns bytes
    assert h.computeWeightedAverage(mixed, w) == ref
    assert r.computeWeightedAverage(mixed, w[:-1]) == b""          # ckks.cpp:265-268


def test_rescaled_average_and_level_aware_decrypt(res, oracles):
    r, o = res["pybind"], oracles["pybind"]
    rng = np.random.default_rng(3)
    xs = [rng.normal(0, 0.05, 9000) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    enc = [r.encrypt(x) for x in xs]
    avg = r.computeWeightedAverage(enc, w, rescale=True)
    assert avg.limbs == 1 and avg.depth == 1 and avg.level == 1
    ref = o.rescale(o.computeWeightedAverage([bytes(e) for e in enc], w))
    assert sha(avg) == sha(ref)
    assert np.array_equal(r.decrypt(avg, 9000), o.decrypt(ref, 9000))


def test_many_vectors_in_flight_share_the_input_slots(res, oracles):
    """encrypt returns before its kernel runs; the calls rotate over eight input slots and streams, so with 24 vectors
    in flight every slot is reused twice while earlier encryptions may still be reading it."""
    r, o = res["code"], oracles["code"]
    rng = np.random.default_rng(8)
    xs = [rng.normal(0, 0.05, 4096 * (1 + i % 5) + i) for i in range(24)]
    r.set_sampler_seed(77)
    try:
        enc = [r.encrypt(x) for x in xs]
    finally:
        r.set_sampler_seed(None)
    for x, e in zip(xs, enc):
        assert sha(e) == sha(o.encrypt(x, seed=77))


def test_pool_accounting_and_limit(shelfi):
    r = shelfi.CKKS("ckks", 4096, 52, str(FIXTURE_DIRS["pybind"]) + "/", device=0, resident=True)
    r.loadCryptoParams()
    x = np.zeros(40 * 4096)
    a = r.encrypt(x)
    per = 40 * 2 * 2 * 8192 * 8
    st = r.resident_stats()
    assert st["in_use"] >= per and st["limit"] > st["in_use"]
    del a
    st2 = r.resident_stats()
    assert st2["in_use"] == 0 and st2["pooled"] >= per          # parked for reuse, not returned to the driver
    b = r.encrypt(x)
    assert r.resident_stats()["pooled"] == 0                     # ... and reused
    r.set_resident_limit(3 * per)
    keep = [b, r.encrypt(x)]
    with pytest.raises(shelfi.ShelfiError) as e:
        keep.append(r.encrypt(x))
        keep.append(r.encrypt(x))
    assert e.value.code == -8 and "device pool exhausted" in str(e.value) and "MiB" in str(e.value)
    keep.clear()
    del b
    assert np.abs(r.decrypt(r.encrypt(x + 1.0), x.size) - 1.0).max() < 1e-9   # usable again once vectors are freed


def test_resident_path_on_a_deep_context(shelfi, deep_oracles):
    o = deep_oracles["n16k_l4"]
    r = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0, resident=True)
    r.loadCryptoParams()
    assert r.load_status == 0
    rng = np.random.default_rng(4)
    xs = [rng.normal(0, 0.05, 3 * 8192 + 11) for _ in range(3)]
    w = [0.25, 0.5, 0.25]
    r.set_sampler_seed(9)
    try:
        enc = [r.encrypt(x) for x in xs]
    finally:
        r.set_sampler_seed(None)
    assert [sha(e) for e in enc] == [sha(o.encrypt(x, seed=9)) for x in xs]
    avg = r.computeWeightedAverage(enc, w)
    ref = o.computeWeightedAverage([bytes(e) for e in enc], w)
    assert sha(avg) == sha(ref)
    assert np.array_equal(r.decrypt(avg, xs[0].size), o.decrypt(ref, xs[0].size))


def test_config2_size_round_trip_through_handles(res, shelfi):
    """BASELINE configs[1] through the drop-in in resident mode: 10 learners x 1,663,370 float32 parameters."""
    r = res["pybind"]
    P, K = 1_663_370, 10
    rng = np.random.default_rng(2)
    xs = [shelfi.pinned_array(rng.normal(0, 0.05, P).astype(np.float32)) for _ in range(K)]
    w = [1.0 / K] * K
    out = r.decrypt(r.computeWeightedAverage([r.encrypt(x) for x in xs], w), P)
    expect = sum(np.float64(np.float32(1.0 / K)) * x.astype(np.float64) for x in xs)
    assert out.shape == (P,) and np.abs(out - expect).max() / np.abs(expect).max() < 1e-6
    assert r.resident_stats()["in_use"] == 0


def test_decrypt_under_the_weighted_sum_is_the_plain_decrypt(res, gpu_ckks, shelfi):
    """A large average is summed in four parts and each part is decrypted on a side stream while the later parts are
    still being summed (engine.cpp: shelfi_ct::part_ev).  Same doubles as the host-buffer decrypt of the same bytes,
    step after step with recycled buffers, also when the average is decrypted twice or serialised in between."""
    r, g = res["pybind"], gpu_ckks["pybind"]
    P, K = 300 * 4096 + 123, 4
    rng = np.random.default_rng(31)
    w = [0.1, 0.2, 0.3, 0.4]
    for step in range(4):
        xs = [shelfi.pinned_array(rng.normal(0, 0.05 * (step + 1), P).astype(np.float32)) for _ in range(K)]
        avg = r.computeWeightedAverage([r.encrypt(x) for x in xs], w)
        first = np.array(r.decrypt(avg, P))            # by part, under the sums
        if step % 2:
            blob = bytes(avg)
            again = np.array(r.decrypt(avg, P))        # the parts' events are long past
            assert np.array_equal(first.view(np.uint64), again.view(np.uint64))
            ref = np.asarray(g.decrypt(blob, P))
            assert np.array_equal(first.view(np.uint64), ref.view(np.uint64)), step
        expect = sum(np.float64(np.float32(wi)) * x.astype(np.float64) for wi, x in zip(w, xs))
        assert np.abs(first - expect).max() / np.abs(expect).max() < 1e-6, step
        del avg
    assert r.resident_stats()["in_use"] == 0
