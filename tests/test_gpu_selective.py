"""-m gpu: selective encryption through the C ABI (callers code/benchmark_selection.py:152-193,
benchmark_selection_rate.py:133-141): gather fused into the encode load, scatter fused into the decode store, plaintext
weighted sum for the parameters that stay unencrypted.  Checked against the oracle on the gathered copies."""
import ctypes as C

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from conftest import DEEP_CONTEXTS  # noqa: E402
from gpu_util import check, host_u64, ptr, stream  # noqa: E402

pytestmark = pytest.mark.gpu
RTOL = 1e-6


@pytest.fixture(scope="module")
def gpu_deep_one(shelfi, deep_oracles):
    o = deep_oracles["n16k_l4"]
    h = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0)
    h.loadCryptoParams()
    assert h.load_status == 0
    return h, o


def _ctx(which, gpu_ckks, oracles, gpu_deep_one):
    return (gpu_ckks[which], oracles[which]) if which in oracles else gpu_deep_one


def _ptrs(ts):
    return (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])


@pytest.mark.parametrize("which", ["pybind", "n16k_l4"])
def test_selected_encrypt_is_encrypt_of_the_gathered_vector(shelfi, gpu_ckks, oracles, gpu_deep_one, which):
    h, o = _ctx(which, gpu_ckks, oracles, gpu_deep_one)
    L, N, batch = o.num_limbs, o.ring_dim, o.batchSize
    rng = np.random.default_rng(31)
    k, P = 3, 5 * batch + 17
    n_sel = 2 * batch + 100                                   # three ciphertexts, the last one short
    xs = [rng.normal(0, 0.05, P) for _ in range(k)]
    sel = rng.permutation(P)[:n_sel].astype(np.uint32)
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dsel = torch.from_numpy(sel.view(np.int32)).cuda()
    dct = [torch.empty((3, 2, L, N), dtype=torch.int64, device="cuda") for _ in range(k)]
    ref = [torch.empty((3, 2, L, N), dtype=torch.int64, device="cuda") for _ in range(k)]
    dg = [torch.from_numpy(np.ascontiguousarray(x[sel])).cuda() for x in xs]
    lib = shelfi.lib()
    h.set_sampler_seed(77)
    check(h, lib.shelfi_encrypt_selected_many_device(h.handle, _ptrs(dx), k, ptr(dsel), n_sel, 4, _ptrs(dct), stream()))
    check(h, lib.shelfi_encrypt_many_device(h.handle, _ptrs(dg), k, n_sel, 4, _ptrs(ref), stream()))
    h.set_sampler_seed(None)
    for i in range(k):
        assert torch.equal(dct[i], ref[i]), i
    got = host_u64(dct[1])
    for j in range(3):
        assert np.array_equal(got[j], o.encrypt_ct(xs[1][sel][j * batch:(j + 1) * batch], 77, 4 + j)), j


@pytest.mark.parametrize("which", ["pybind", "n16k_l4"])
def test_selective_aggregation_end_to_end(shelfi, gpu_ckks, oracles, gpu_deep_one, which):
    """encrypt selected -> weighted sum -> decrypt + scatter over the plaintext weighted sum == the plain weighted
    average everywhere; the unselected entries are exactly numpy's sum of w_k * x_k."""
    h, o = _ctx(which, gpu_ckks, oracles, gpu_deep_one)
    L, N, batch = o.num_limbs, o.ring_dim, o.batchSize
    rng = np.random.default_rng(32)
    k, P = 4, 3 * batch + 999
    n_sel = batch + 321
    n_ct = 2
    xs = [rng.normal(0, 0.05, P) for _ in range(k)]
    w = [np.float32(v) for v in (0.4, 0.3, 0.2, 0.1)]
    sel = np.sort(rng.permutation(P)[:n_sel]).astype(np.uint32)
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dsel = torch.from_numpy(sel.view(np.int32)).cuda()
    dct = [torch.empty((n_ct, 2, L, N), dtype=torch.int64, device="cuda") for _ in range(k)]
    davg = torch.empty((n_ct, 2, L, N), dtype=torch.int64, device="cuda")
    dout = torch.full((P,), 123.0, dtype=torch.float64, device="cuda")
    wf = (C.c_float * k)(*[float(v) for v in w])
    lib = shelfi.lib()
    check(h, lib.shelfi_encrypt_selected_many_device(h.handle, _ptrs(dx), k, ptr(dsel), n_sel, 0, _ptrs(dct), stream()))
    check(h, lib.shelfi_weighted_sum_device(h.handle, _ptrs(dct), wf, k, n_ct, ptr(davg), stream()))
    check(h, lib.shelfi_plain_weighted_sum_device(h.handle, _ptrs(dx), wf, k, P, ptr(dout), stream()))
    plain = dout.cpu().numpy().copy()
    expect = sum(np.float64(wi) * x for wi, x in zip(w, xs))
    assert np.array_equal(plain, expect)                       # same operation order as numpy: bit-identical
    sf2 = h.scaling_factor * h.scaling_factor
    check(h, lib.shelfi_decrypt_scatter_device(h.handle, ptr(davg), n_ct, sf2, ptr(dout), ptr(dsel), n_sel, stream()))
    got = dout.cpu().numpy()
    mask = np.zeros(P, bool)
    mask[sel] = True
    assert np.array_equal(got[~mask], expect[~mask])          # untouched by the scatter
    assert not np.array_equal(got[mask], expect[mask])        # these came through CKKS (noise in the low bits)
    assert np.abs(got - expect).max() / np.abs(expect).max() < RTOL
    # the decrypted part equals a dense decrypt of the same ciphertexts
    dense = torch.empty(n_sel, dtype=torch.float64, device="cuda")
    check(h, lib.shelfi_decrypt_device(h.handle, ptr(davg), n_ct, sf2, ptr(dense), n_sel, stream()))
    assert np.array_equal(got[sel], dense.cpu().numpy())


def test_plain_weighted_sum_many_learners_and_argument_checks(shelfi, gpu_ckks):
    h = gpu_ckks["pybind"]
    lib = shelfi.lib()
    rng = np.random.default_rng(33)
    k, n = 70, 10_001                                          # more learners than one launch takes (chained launches)
    xs = [rng.normal(0, 1, n) for _ in range(k)]
    w = rng.random(k).astype(np.float32)
    dx = [torch.from_numpy(x).cuda() for x in xs]
    dout = torch.empty(n, dtype=torch.float64, device="cuda")
    wf = (C.c_float * k)(*[float(v) for v in w])
    check(h, lib.shelfi_plain_weighted_sum_device(h.handle, _ptrs(dx), wf, k, n, ptr(dout), stream()))
    acc = np.zeros(n)
    for wi, x in zip(w, xs):
        acc = acc + np.float64(wi) * x
    assert np.array_equal(dout.cpu().numpy(), acc)
    assert lib.shelfi_plain_weighted_sum_device(h.handle, _ptrs(dx), wf, 0, n, ptr(dout), stream()) != 0
    assert lib.shelfi_encrypt_selected_many_device(h.handle, _ptrs(dx), 1, None, 5, 0, _ptrs(dx), stream()) != 0
    assert lib.shelfi_decrypt_scatter_device(h.handle, ptr(dout), 1, 1.0, ptr(dout), ptr(dout), 2 * 4096 + 1,
                                             stream()) != 0
