"""The PALISADE semantics no fixture pins, as switches of the oracle (the engine has the same ones under the same ids:
tests/test_gpu_options.py): EvalMult(ct, double)'s rounding, and the decode noise of PALISADE >= 1.10.6."""
from __future__ import annotations

import numpy as np
import pytest


@pytest.fixture()
def o(oracle_mod):
    from conftest import FIXTURE_DIRS
    orc = oracle_mod.Oracle(cryptodir=str(FIXTURE_DIRS["pybind"]) + "/")
    assert orc.loadCryptoParams() == 0
    return orc


def test_evalmult_rounding_switch_only_moves_negative_half_cases(o):
    delta = o.scaling_factor
    for w in (0.5, 0.2, 0.3, 1.0 / 3, 1e-3, 0.999):
        a = o.weight_constant(w)
        o.set_option(o.OPT_EVALMULT_ROUND, 1)
        b = o.weight_constant(w)
        o.set_option(o.OPT_EVALMULT_ROUND, 0)
        assert a == b == int(np.floor(float(np.float32(w)) * delta + 0.5))
    for w in (-0.5, -0.2, -1.0 / 3):          # truncation toward zero of (c Delta + 0.5) vs round half away from zero
        a = o.weight_constant(w)
        o.set_option(o.OPT_EVALMULT_ROUND, 1)
        b = o.weight_constant(w)
        o.set_option(o.OPT_EVALMULT_ROUND, 0)
        x = float(np.float32(w)) * delta
        assert a == int(x + 0.5) and b == -int(np.floor(-x + 0.5)) and a - b in (0, 1)   # llround: half away from zero
    with pytest.raises(ValueError):
        o.set_option(o.OPT_EVALMULT_ROUND, 2)


def test_decode_noise_is_floored_seeded_and_far_below_the_tolerance(o):
    n = 3 * 4096 + 100
    rng = np.random.default_rng(1)
    xs = [rng.normal(0, 0.05, n) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    avg = o.computeWeightedAverage([o.encrypt(x, seed=i) for i, x in enumerate(xs)], w)
    plain = o.decrypt(avg, n)
    o.set_option(o.OPT_DECODE_NOISE, 1)
    o.set_option(o.OPT_NOISE_SEED, 11)
    a = o.decrypt(avg, n)
    b = o.decrypt(avg, n)
    o.set_option(o.OPT_NOISE_SEED, 12)
    c = o.decrypt(avg, n)
    o.set_option(o.OPT_DECODE_NOISE, 0)
    assert np.array_equal(a, b) and not np.array_equal(a, c) and not np.array_equal(a, plain)
    # floor: sd >= sqrt(2) sqrt(N) / 8 at scale 2^52; FFTSpecial sums 2 * slots of them
    d = a - plain
    expect_sd = np.sqrt(2) * 0.125 * np.sqrt(8192) * 2.0**-52 * np.sqrt(4096)
    assert 0.5 * expect_sd < d.std() < 50 * expect_sd
    # different ciphertexts of one call draw different noise
    assert not np.array_equal(d[:4096], d[4096:8192])
    ref = sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, xs))
    assert np.abs(a - ref).max() / np.abs(ref).max() < 1e-6          # the 1e-6 tier holds with the noise on


def test_decode_noise_estimates_the_error_of_a_noisy_vector(o):
    """On a conjugate-symmetric vector plus asymmetric noise of known size the estimate tracks that size."""
    slots = 4096
    rng = np.random.default_rng(2)
    key = np.arange(8, dtype=np.uint32)
    for sigma in (1e3, 1e6):
        # a real message decodes to a vector z with Conjugate(z) = z; the decryption error breaks that symmetry
        z = np.zeros(slots, np.complex128)
        z += rng.normal(0, sigma, slots) + 1j * rng.normal(0, sigma, slots)
        noisy, sd = o.decode_noise(z, key, 0)
        assert 0.5 * sigma < sd / np.sqrt(2) < 2.0 * sigma
        added = noisy - z
        assert 0.8 * sd < added.real.std() < 1.2 * sd and 0.8 * sd < added.imag.std() < 1.2 * sd
    _, sd0 = o.decode_noise(np.zeros(slots, np.complex128), key, 0)
    assert sd0 == pytest.approx(np.sqrt(2) * 0.125 * np.sqrt(8192))
