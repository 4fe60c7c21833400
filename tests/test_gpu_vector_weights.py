"""-m gpu: per-slot plaintext weights through the C ABI — PALISADE EvalMult(ct, plaintext) + EvalAdd, the ct x
plaintext-vector aggregation of the TenSEAL/FLARE-style workload (code/benchmark_nvidia.py:439,461-465).
Bit-exact against the oracle's restatement, decoded result against sum_l w_l * x_l."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from gpu_util import check, host_u64, ptr, stream  # noqa: E402

pytestmark = pytest.mark.gpu
RTOL = 1e-6


def _ptrs(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


@pytest.fixture(scope="module")
def ctxs(shelfi, gpu_ckks, oracles, deep_oracles):
    out = {"pybind": (gpu_ckks["pybind"], oracles["pybind"])}
    for name in ("n16k_l4", "n32k_l8", "n16k_sparse"):
        o = deep_oracles[name]
        h = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0)
        h.loadCryptoParams()
        assert h.load_status == 0
        out[name] = (h, o)
    return out


@pytest.mark.parametrize("which", ["pybind", "n16k_l4", "n32k_l8", "n16k_sparse"])
def test_vector_weighted_sum_matches_oracle(shelfi, ctxs, which):
    h, o = ctxs[which]
    L, N, batch = o.num_limbs, o.ring_dim, o.batchSize
    rng = np.random.default_rng(41)
    k, n_ct = 3, 2
    n = batch + batch // 2 + 3                                   # second ciphertext short
    xs = [rng.normal(0, 0.05, n) for _ in range(k)]
    ws = [rng.uniform(-1, 1, n) for _ in range(k)]
    lib = shelfi.lib()
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dw = [torch.from_numpy(w).cuda() for w in ws]
    dct = [torch.empty((n_ct, 2, L, N), dtype=torch.int64, device="cuda") for _ in range(k)]
    dout = torch.empty((n_ct, 2, L, N), dtype=torch.int64, device="cuda")
    h.set_sampler_seed(9)
    check(h, lib.shelfi_encrypt_many_device(h.handle, _ptrs(dx), k, n, 0, _ptrs(dct), stream()))
    h.set_sampler_seed(None)
    check(h, lib.shelfi_weighted_sum_vector_device(h.handle, _ptrs(dct), _ptrs(dw), k, n, ptr(dout), stream()))
    got = host_u64(dout)
    cts = [host_u64(t) for t in dct]
    for j in range(n_ct):
        acc = None
        for i in range(k):
            acc = o.ct_mul_plain_accumulate(cts[i][j], ws[i][j * batch:(j + 1) * batch], acc)
        assert np.array_equal(got[j], acc), (which, j)
    dec = torch.empty(n, dtype=torch.float64, device="cuda")
    sf2 = h.scaling_factor * h.scaling_factor
    check(h, lib.shelfi_decrypt_device(h.handle, ptr(dout), n_ct, sf2, ptr(dec), n, stream()))
    expect = sum(w * x for w, x in zip(ws, xs))
    assert np.abs(dec.cpu().numpy() - expect).max() / np.abs(expect).max() < RTOL


def test_shared_weight_vector_and_scalar_equivalence(shelfi, ctxs):
    """All learners sharing one vector (the TenSEAL script's constant 0.3 vector) is encoded once; a constant vector c
    decodes to the same average as the scalar form with weight c (different ciphertext bits: the scalar form multiplies
    by round(c * Delta), the vector form by the encoding of the constant vector)."""
    h, o = ctxs["pybind"]
    lib = shelfi.lib()
    rng = np.random.default_rng(42)
    k, n = 4, 3 * 4096
    n_ct = 3
    xs = [rng.normal(0, 0.05, n) for _ in range(k)]
    w = np.full(n, 0.3)
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dw = torch.from_numpy(w).cuda()
    dct = [torch.empty((n_ct, 2, 2, 8192), dtype=torch.int64, device="cuda") for _ in range(k)]
    check(h, lib.shelfi_encrypt_many_device(h.handle, _ptrs(dx), k, n, 0, _ptrs(dct), stream()))
    out_v = torch.empty((n_ct, 2, 2, 8192), dtype=torch.int64, device="cuda")
    out_s = torch.empty_like(out_v)
    check(h, lib.shelfi_weighted_sum_vector_device(h.handle, _ptrs(dct), _ptrs([dw] * k), k, n, ptr(out_v), stream()))
    wf = (C.c_float * k)(*[0.3] * k)
    check(h, lib.shelfi_weighted_sum_device(h.handle, _ptrs(dct), wf, k, n_ct, ptr(out_s), stream()))
    got, cts = host_u64(out_v), [host_u64(t) for t in dct]
    acc = None
    for i in range(k):                                          # learner by learner, as PALISADE would
        acc = o.ct_mul_plain_accumulate(cts[i][1], w[4096:8192], acc)
    assert np.array_equal(got[1], acc)
    sf2 = h.scaling_factor ** 2
    dv, ds = torch.empty(n, dtype=torch.float64, device="cuda"), torch.empty(n, dtype=torch.float64, device="cuda")
    check(h, lib.shelfi_decrypt_device(h.handle, ptr(out_v), n_ct, sf2, ptr(dv), n, stream()))
    check(h, lib.shelfi_decrypt_device(h.handle, ptr(out_s), n_ct, sf2, ptr(ds), n, stream()))
    expect = 0.3 * sum(xs)
    assert np.abs(dv.cpu().numpy() - expect).max() / np.abs(expect).max() < RTOL
    assert np.abs(ds.cpu().numpy() - np.float64(np.float32(0.3)) * sum(xs)).max() / np.abs(expect).max() < RTOL
    assert lib.shelfi_weighted_sum_vector_device(h.handle, _ptrs(dct), _ptrs([dw] * k), 0, n, ptr(out_v), stream()) != 0
