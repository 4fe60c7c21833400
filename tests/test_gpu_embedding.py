"""The canonical-embedding definition checks of tests/test_embedding_cpu.py run against the CUDA path through the C
ABI: the device encode is evaluated at zeta^(5^j) with plain complex arithmetic, and the device decrypt is fed
polynomials moved by ring automorphisms (as ciphertexts with c1 = 0, serialized by the oracle's blob writer — test
infrastructure, the wire grammar is pinned by the reference's fixtures)."""
from __future__ import annotations

import numpy as np
import pytest
import torch

from gpu_util import check, ptr, stream
from test_embedding_cpu import automorphism, evaluate_at_slots, mul_x_pow, trivial_ciphertext

pytestmark = pytest.mark.gpu


def gpu_encode(shelfi, h, x):
    n_ct = -(-x.size // h.batchSize)
    dm = torch.empty((n_ct, h.ring_dim), dtype=torch.int64, device="cuda")
    check(h, shelfi.lib().shelfi_encode_device(h.handle, ptr(torch.from_numpy(x).cuda()), x.size, ptr(dm), stream()))
    return dm.cpu().numpy()


def gpu_decode(h, o, coeff, n):
    """decrypt of the trivial ciphertext (NTT(m), 0) through the reference-facing call"""
    blob = o.blob_build(trivial_ciphertext(o, coeff)[None])
    return h.decrypt(blob, n)


def test_device_encode_is_evaluation_at_the_rotation_group(shelfi, gpu_ckks, fixture_name):
    h = gpu_ckks[fixture_name]
    N, slots, delta = h.ring_dim, h.slots, h.scaling_factor
    x = np.random.default_rng(11).normal(0, 1.0, slots)
    m = gpu_encode(shelfi, h, x)[0]
    z = evaluate_at_slots(m, N, slots) / delta
    assert np.abs(z.real - x).max() < 1e-9 and np.abs(z.imag).max() < 1e-9
    u = np.zeros(slots)
    u[3] = 1.0
    zu = evaluate_at_slots(gpu_encode(shelfi, h, u)[0], N, slots) / delta
    assert abs(zu[3] - 1.0) < 1e-9 and np.abs(np.delete(zu, 3)).max() < 1e-9


def test_device_decode_follows_the_automorphisms(shelfi, gpu_ckks, oracles):
    h, o = gpu_ckks["pybind"], oracles["pybind"]
    N, slots = h.ring_dim, h.slots
    rng = np.random.default_rng(12)
    x = rng.normal(0, 1.0, slots)
    m = [int(v) for v in gpu_encode(shelfi, h, x)[0]]
    assert np.abs(gpu_decode(h, o, m, slots) - x).max() < 1e-9
    for g, shift in ((5, 1), (pow(5, 9, 2 * N), 9), (pow(5, -1, 2 * N), -1)):
        out = gpu_decode(h, o, automorphism(m, N, g), slots)
        assert np.abs(out - np.roll(x, -shift)).max() < 1e-9, (g, shift)
    a, b = rng.normal(0, 1.0, slots), rng.normal(0, 1.0, slots)
    ma, mb = [int(v) for v in gpu_encode(shelfi, h, a)[0]], [int(v) for v in gpu_encode(shelfi, h, b)[0]]
    mz = [u + v for u, v in zip(ma, mul_x_pow(mb, N, N // 2))]               # a + i b
    assert np.abs(gpu_decode(h, o, mul_x_pow(mz, N, N // 2), slots) + b).max() < 1e-9       # Re(i (a + i b)) = -b
    conj = automorphism(mz, N, 2 * N - 1)
    assert np.abs(gpu_decode(h, o, mul_x_pow(conj, N, N // 2), slots) - b).max() < 1e-9     # Re(i (a - i b)) = +b


def test_device_decrypt_is_evaluation_of_the_decrypted_polynomial(shelfi, gpu_ckks, oracles):
    """decrypt(encrypt(x)) on the device against the direct evaluation of the device's own integer tap."""
    h, o = gpu_ckks["code"], oracles["code"]
    N, slots, sf = h.ring_dim, h.slots, h.scaling_factor
    x = np.random.default_rng(13).normal(0, 0.5, slots)
    dct = torch.empty(h.ciphertext_words, dtype=torch.int64, device="cuda")
    L = shelfi.lib()
    check(h, L.shelfi_encrypt_device(h.handle, ptr(torch.from_numpy(x).cuda()), slots, 0, ptr(dct), stream()))
    hi = torch.empty(N, dtype=torch.int64, device="cuda")
    lo = torch.empty(N, dtype=torch.int64, device="cuda")
    out = torch.empty(slots, dtype=torch.float64, device="cuda")
    check(h, L.shelfi_decrypt_integer_device(h.handle, ptr(dct), 1, ptr(hi), ptr(lo), stream()))
    check(h, L.shelfi_decrypt_device(h.handle, ptr(dct), 1, sf, ptr(out), slots, stream()))
    torch.cuda.synchronize()
    mt = [(int(a) << 64) | (int(b) & (2**64 - 1)) for a, b in zip(hi.cpu().numpy(), lo.cpu().numpy())]
    z = evaluate_at_slots(mt, N, slots) / sf
    got = out.cpu().numpy()
    assert np.abs(z.real - got).max() < 1e-9 and np.abs(got - x).max() < 1e-9
