"""shelfi_decrypt_static_device: the decrypt whose scratch is a file-scope device array (no allocation in the call).

The entry point has no counterpart in the reference; it was added for callers
that cannot call cudaMalloc.  Checked here through the C ABI against shelfi_decrypt_device on the same ciphertexts:
the two run the same kernels on the same inputs, so the decoded doubles are compared bit for bit (tolerance 0, the
bar of the other device-API tests for integer/byte work and for doubles produced by one FFT order), on two calls in a
row with different inputs — the array keeps the first call's contents — and once more after the existing entry point
has run in between.  This file sorts last on purpose: it passes alone in a fresh
process and after the rest of the suite.
"""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
torch = pytest.importorskip("torch")
from gpu_util import check, ptr, stream  # noqa: E402

STATIC_BYTES = 8 << 20   # SHELFI_STATIC_SCRATCH_BYTES, include/shelfi_b200.h


def per_ct_bytes(h):
    """Scratch bytes one ciphertext needs (header comment of shelfi_decrypt_static_device)."""
    N, L, slots = h.ring_dim, h.num_limbs, h.slots
    general = N > 8192 or L > 2
    words = L * N
    if general:
        words += N + (2 * slots if slots > 4096 else 0)
    return 8 * words


def encrypt(shelfi, h, x, seed):
    n = len(x)
    n_ct = -(-n // h.batchSize)
    dct = torch.empty(n_ct * h.ciphertext_words, dtype=torch.int64, device="cuda")
    h.set_sampler_seed(seed)
    check(h, shelfi.lib().shelfi_encrypt_device(h.handle, ptr(torch.from_numpy(x).cuda()), n, 0, ptr(dct), stream()))
    h.set_sampler_seed(None)
    return dct, n_ct


def decrypt_both(shelfi, h, dct, n_ct, n, order):
    """Decode dct through both entry points in the given order; returns {name: ndarray}."""
    L = shelfi.lib()
    fn = {"static": L.shelfi_decrypt_static_device, "heap": L.shelfi_decrypt_device}
    out = {}
    for name in order:
        d = torch.full((n,), float("nan"), dtype=torch.float64, device="cuda")
        check(h, fn[name](h.handle, ptr(dct), n_ct, h.scaling_factor, ptr(d), n, stream()))
        torch.cuda.synchronize()
        out[name] = d.cpu().numpy()
    return out


def two_calls_in_a_row(shelfi, h, sizes):
    rng = np.random.default_rng(2024)
    for call, n in enumerate(sizes):
        x = rng.normal(0, 1.0 + call, n)
        dct, n_ct = encrypt(shelfi, h, x, 900 + call)
        # the static entry point first on the first call (nothing primed by the other one), second on the next
        got = decrypt_both(shelfi, h, dct, n_ct, n, ("static", "heap") if call % 2 == 0 else ("heap", "static"))
        assert np.array_equal(got["static"].view(np.uint64), got["heap"].view(np.uint64)), (call, n)
        assert np.abs(got["static"] - x).max() < 1e-6, (call, n)


def test_static_decrypt_matches_decrypt_device_twice(shelfi, gpu_ckks):
    """N = 8192, two limbs (BASELINE config 1/2 context): the fused decode route."""
    h = gpu_ckks["code"]
    assert per_ct_bytes(h) == 8 * 2 * 8192
    cap = STATIC_BYTES // per_ct_bytes(h)
    assert cap == 64
    # full array, then a short ragged call that leaves most of the previous contents in place, then full again
    two_calls_in_a_row(shelfi, h, (cap * h.batchSize, 2 * h.batchSize + 17, cap * h.batchSize - 1, 1))


def test_static_decrypt_general_route_twice(shelfi, deep_oracles):
    """Three limbs at N = 8192 and the two-tile ring of config 3: CRT + FFT scratch also comes from the array."""
    for name in ("n8k_l3", "n16k_l4"):
        o = deep_oracles[name]
        h = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0)
        h.loadCryptoParams()
        assert h.load_status == 0, name
        cap = STATIC_BYTES // per_ct_bytes(h)
        assert cap >= 2, name
        two_calls_in_a_row(shelfi, h, (cap * h.batchSize, h.batchSize + 3, 2 * h.batchSize))


def test_static_decrypt_refuses_what_does_not_fit(shelfi, gpu_ckks):
    h = gpu_ckks["code"]
    L = shelfi.lib()
    cap = STATIC_BYTES // per_ct_bytes(h)
    n_ct = cap + 1
    n = n_ct * h.batchSize
    dct = torch.zeros(n_ct * h.ciphertext_words, dtype=torch.int64, device="cuda")
    out = torch.zeros(n, dtype=torch.float64, device="cuda")
    rc = L.shelfi_decrypt_static_device(h.handle, ptr(dct), n_ct, h.scaling_factor, ptr(out), n, stream())
    assert rc == -1   # SHELFI_ERR_ARG
    msg = h._err()
    assert "static array holds 8388608" in msg and f"at most {cap} ciphertexts" in msg, msg
    torch.cuda.synchronize()
    assert not out.any()                       # refused before any launch
    # the same request is fine through the allocating entry point, and one ciphertext fewer through the static one
    check(h, L.shelfi_decrypt_device(h.handle, ptr(dct), n_ct, h.scaling_factor, ptr(out), n, stream()))
    check(h, L.shelfi_decrypt_static_device(h.handle, ptr(dct), cap, h.scaling_factor, ptr(out), cap * h.batchSize, stream()))
    torch.cuda.synchronize()
    # a handle with nothing loaded is refused too (no lazy set-up in this entry point)
    empty = shelfi.CKKS("ckks", 4096, 52, "/nonexistent/", device=0)
    rc = L.shelfi_decrypt_static_device(empty.handle, ptr(dct), 1, 1.0, ptr(out), 1, stream())
    assert rc == -2   # SHELFI_ERR_NOT_LOADED


def test_static_decrypt_allocates_nothing(shelfi, gpu_ckks):
    """Device memory in use (driver view) does not move across calls of the new entry point."""
    h = gpu_ckks["code"]
    x = np.random.default_rng(5).normal(0, 1, 8 * h.batchSize)
    dct, n_ct = encrypt(shelfi, h, x, 77)
    out = torch.empty(len(x), dtype=torch.float64, device="cuda")
    L = shelfi.lib()
    check(h, L.shelfi_decrypt_static_device(h.handle, ptr(dct), n_ct, h.scaling_factor, ptr(out), len(x), stream()))
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(20):
        check(h, L.shelfi_decrypt_static_device(h.handle, ptr(dct), n_ct, h.scaling_factor, ptr(out), len(x), stream()))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0
    assert np.abs(out.cpu().numpy() - x).max() < 1e-6
