"""Oracle-INDEPENDENT checks of the canonical embedding.

The engine and the oracle both restate PALISADE's FFTSpecial / FFTSpecialInv; they agree bit for bit, which proves
nothing about whether both follow the HEAAN / PALISADE convention.  These tests pin the convention from its DEFINITION,
with plain complex arithmetic and no FFT code at all:

    slot j of an encoded polynomial m(X) is  m(zeta^(5^j)) / Delta,   zeta = exp(i pi / N)  (a primitive 2N-th root),

so encode is checked by evaluating the integer polynomial at those points, decode by evaluating the decrypted integer
polynomial, and the slot ORDER by the ring automorphisms: X -> X^5 rotates the slots by one, X -> X^(2N-1) conjugates
them.  A wrong generator, a conjugated root, a bit-reversal left in or a swapped real/imaginary half each break one.
The GPU versions of the same checks are in tests/test_gpu_embedding.py.
"""
from __future__ import annotations

import numpy as np
import pytest


def slot_points(N: int, slots: int):
    """exponents e_j = 5^j mod 2N, j < slots"""
    M = 2 * N
    e = np.empty(slots, np.int64)
    v = 1
    for j in range(slots):
        e[j] = v
        v = (v * 5) % M
    return e


def evaluate_at_slots(coeff, N: int, slots: int, chunk: int = 256):
    """m(zeta^(5^j)) for j < slots by direct summation; coeff: length-N sequence of (big) integers or floats"""
    M = 2 * N
    c = np.array([float(v) for v in coeff], dtype=np.float64)     # exact enough: |m| < 2^63, compared at 1e-9 relative
    table = np.exp(1j * np.pi * np.arange(M) / N)
    e = slot_points(N, slots)
    k = np.arange(N, dtype=np.int64)
    out = np.empty(slots, np.complex128)
    for s in range(0, slots, chunk):
        idx = (e[s:s + chunk, None] * k[None, :]) % M
        out[s:s + chunk] = table[idx] @ c
    return out


def automorphism(coeff, N: int, g: int):
    """m(X) -> m(X^g) in Z[X]/(X^N + 1)"""
    out = [0] * N
    for k, v in enumerate(coeff):
        t = (k * g) % (2 * N)
        if t < N:
            out[t] += v
        else:
            out[t - N] -= v
    return out


def mul_x_pow(coeff, N: int, s: int):
    """m(X) * X^s, negacyclic"""
    out = [0] * N
    for k, v in enumerate(coeff):
        t = (k + s) % (2 * N)
        if t < N:
            out[t] += v
        else:
            out[t - N] -= v
    return out


def trivial_ciphertext(o, coeff):
    """(c0, c1) = (NTT(m), 0) in every limb: decrypts to m itself whatever the secret key (c1 = 0)"""
    L, N = o.num_limbs, o.ring_dim
    ct = np.zeros((2, L, N), np.uint64)
    for l in range(L):
        q = o.modulus(l)
        ct[0, l] = o.ntt_forward(l, np.array([int(v) % q for v in coeff], dtype=np.uint64))
    return ct


@pytest.fixture(scope="module")
def o(oracles):
    return oracles["pybind"]


def test_encode_is_evaluation_at_the_rotation_group(o):
    N, slots, delta = o.ring_dim, o.slots, o.scaling_factor
    x = np.random.default_rng(1).normal(0, 1.0, slots)
    m = o.encode_coeffs(x)
    z = evaluate_at_slots(m, N, slots) / delta
    assert np.abs(z.real - x).max() < 1e-9          # slot j <-> zeta^(5^j), no permutation, no conjugation
    assert np.abs(z.imag).max() < 1e-9              # a real vector encodes to a conjugate-symmetric polynomial
    # each single slot on its own (a unit vector) lands on its own evaluation point
    for j in (0, 1, 2, slots // 2, slots - 1):
        u = np.zeros(slots)
        u[j] = 1.0
        zu = evaluate_at_slots(o.encode_coeffs(u), N, slots) / delta
        assert abs(zu[j] - 1.0) < 1e-9 and np.abs(np.delete(zu, j)).max() < 1e-9


def test_decode_is_evaluation_of_the_decrypted_polynomial(o):
    N, slots, sf = o.ring_dim, o.slots, o.scaling_factor
    x = np.random.default_rng(2).normal(0, 0.5, slots)
    ct = o.encrypt_ct(x, 17, 0)
    out, hi, lo = o.decrypt_ct(ct, sf, want_coeffs=True)
    mt = [(int(h) << 64) | int(l) for h, l in zip(hi, lo)]      # centred CRT integers (two's complement hi, lo)
    z = evaluate_at_slots(mt, N, slots) / sf
    assert np.abs(z.real - out).max() < 1e-9
    assert np.abs(out - x).max() < 1e-9


def test_x_to_x5_rotates_the_slots_by_one(o):
    N, slots, sf = o.ring_dim, o.slots, o.scaling_factor
    x = np.random.default_rng(3).normal(0, 1.0, slots)
    m = [int(v) for v in o.encode_coeffs(x)]
    for g, shift in ((5, 1), (25, 2), (pow(5, 7, 2 * N), 7)):
        out = o.decrypt_ct(trivial_ciphertext(o, automorphism(m, N, g)), sf)
        assert np.abs(out - np.roll(x, -shift)).max() < 1e-9, (g, shift)
    # the inverse generator rotates the other way
    out = o.decrypt_ct(trivial_ciphertext(o, automorphism(m, N, pow(5, -1, 2 * N))), sf)
    assert np.abs(out - np.roll(x, 1)).max() < 1e-9


def test_x_to_x_inverse_conjugates_the_slots(o):
    """a + i b is encoded as Encode(a) + X^(N/2) Encode(b) (X^(N/2) evaluates to i at every zeta^(5^j), as 5^j = 1 mod
    4).  Multiplying by X^(N/2) once more turns the real part into -b; after conjugation it is +b."""
    N, slots, sf = o.ring_dim, o.slots, o.scaling_factor
    rng = np.random.default_rng(4)
    a, b = rng.normal(0, 1.0, slots), rng.normal(0, 1.0, slots)
    ma, mb = [int(v) for v in o.encode_coeffs(a)], [int(v) for v in o.encode_coeffs(b)]
    m = [u + v for u, v in zip(ma, mul_x_pow(mb, N, N // 2))]
    assert np.abs(o.decrypt_ct(trivial_ciphertext(o, m), sf) - a).max() < 1e-9                       # Re(a + ib)
    assert np.abs(o.decrypt_ct(trivial_ciphertext(o, mul_x_pow(m, N, N // 2)), sf) + b).max() < 1e-9  # Re(i(a + ib)) = -b
    conj = automorphism(m, N, 2 * N - 1)
    assert np.abs(o.decrypt_ct(trivial_ciphertext(o, conj), sf) - a).max() < 1e-9
    assert np.abs(o.decrypt_ct(trivial_ciphertext(o, mul_x_pow(conj, N, N // 2)), sf) - b).max() < 1e-9  # Re(i(a - ib)) = +b


def test_slotwise_product_is_the_negacyclic_product(o):
    """Evaluation is a ring homomorphism: the exact negacyclic product of two integer polynomials, decoded at scale
    Delta^2, is the slot-wise product of what the factors decode to.  (One factor is an encoding cut to its first 320
    coefficients, so that the exact big-integer product stays cheap; it no longer decodes to a real vector, which the
    identity does not need.)"""
    N, slots, delta = o.ring_dim, o.slots, o.scaling_factor
    rng = np.random.default_rng(5)
    mx = [int(v) for v in o.encode_coeffs(rng.normal(0, 1.0, slots))]
    my = [int(v) for v in o.encode_coeffs(rng.normal(0, 1.0, slots))[:320]] + [0] * (N - 320)
    A = np.array(mx, dtype=object)
    prod = np.zeros(N, dtype=object)
    for kk in range(320):                      # prod += my[kk] * X^kk * mx  (X^N = -1)
        b = my[kk]
        prod[kk:] += A[:N - kk] * b
        prod[:kk] -= A[N - kk:] * b
    zx = evaluate_at_slots(mx, N, slots) / delta
    zy = evaluate_at_slots(my, N, slots) / delta
    zp = evaluate_at_slots(list(prod), N, slots) / (delta * delta)
    assert np.abs(zp - zx * zy).max() < 1e-6 * max(1.0, np.abs(zx * zy).max())


@pytest.mark.parametrize("name", ["n16k_sparse"])
def test_sparse_packing_follows_the_same_definition(deep_oracles, name):
    """batch < N/2: coefficients sit gap apart and slot j is the evaluation at zeta^(gap 5^j)."""
    o = deep_oracles[name]
    N, slots, delta = o.ring_dim, o.slots, o.scaling_factor
    gap = (N // 2) // slots
    x = np.random.default_rng(6).normal(0, 1.0, slots)
    m = o.encode_coeffs(x)
    assert all(int(m[k]) == 0 for k in range(N) if k % gap)
    sub = [int(m[k]) for k in range(0, N, gap)]                 # p(Y), Y = X^gap, a ring of dimension N / gap
    z = evaluate_at_slots(sub, N // gap, slots) / delta
    assert np.abs(z.real - x).max() < 1e-9 and np.abs(z.imag).max() < 1e-9
