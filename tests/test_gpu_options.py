"""The engine's switches for the unpinned PALISADE semantics (shelfi_set_option), against the oracle's under the same
ids: EvalMult rounding is bit-exact either way; the decode noise agrees with the oracle's to the accuracy of the device's
log / sin / cos (it is a random quantity by design: there is no bit-exact tier) and keeps the 1e-6 tier."""
from __future__ import annotations

import numpy as np
import pytest

from conftest import FIXTURE_DIRS

pytestmark = pytest.mark.gpu


@pytest.fixture()
def pair(shelfi, oracle_mod):
    d = str(FIXTURE_DIRS["code"]) + "/"
    h = shelfi.CKKS("ckks", 4096, 52, d, device=0)
    h.loadCryptoParams()
    o = oracle_mod.Oracle(cryptodir=d)
    assert h.load_status == 0 and o.loadCryptoParams() == 0
    return h, o


def test_evalmult_rounding_switch(pair):
    h, o = pair
    rng = np.random.default_rng(1)
    xs = [rng.normal(0, 0.05, 6000) for _ in range(3)]
    blobs = [o.encrypt(x, seed=i) for i, x in enumerate(xs)]
    w = [-0.5, 0.2, -1.0 / 3]
    assert h.get_option(h.OPT_EVALMULT_ROUND) == 0
    base = h.computeWeightedAverage(blobs, w)
    assert base == o.computeWeightedAverage(blobs, w)
    h.set_option(h.OPT_EVALMULT_ROUND, 1)
    o.set_option(o.OPT_EVALMULT_ROUND, 1)
    alt = h.computeWeightedAverage(blobs, w)
    assert alt == o.computeWeightedAverage(blobs, w) and alt != base
    assert [h.weight_constant(v) for v in w] == [o.weight_constant(v) for v in w]
    with pytest.raises(Exception):
        h.set_option(h.OPT_EVALMULT_ROUND, 5)


@pytest.mark.parametrize("resident", [False, True])
def test_decode_noise_matches_the_oracle_and_keeps_the_tolerance(pair, shelfi, resident):
    h, o = pair
    h.resident = resident
    n = 5 * 4096 + 17
    rng = np.random.default_rng(2)
    xs = [rng.normal(0, 0.05, n) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    blobs = [o.encrypt(x, seed=i) for i, x in enumerate(xs)]
    avg = o.computeWeightedAverage(blobs, w)
    src = h.to_resident(avg) if resident else avg
    plain = h.decrypt(src, n)
    assert np.array_equal(plain, o.decrypt(avg, n))
    h.set_option(h.OPT_DECODE_NOISE, 1)
    o.set_option(o.OPT_DECODE_NOISE, 1)
    h.set_sampler_seed(21)
    o.set_option(o.OPT_NOISE_SEED, 21)
    try:
        a, b = h.decrypt(src, n), h.decrypt(src, n)
    finally:
        h.set_sampler_seed(None)
    ref = o.decrypt(avg, n)
    noise = ref - plain
    assert np.array_equal(a, b) and not np.array_equal(a, plain)
    assert np.abs(a - ref).max() < 1e-3 * np.abs(noise).max()          # same noise up to libm differences
    c = h.decrypt(src, n)                                               # unseeded: a fresh key per call
    d = h.decrypt(src, n)
    assert not np.array_equal(c, d) and not np.array_equal(c, a)
    assert 0.3 < (c - plain).std() / noise.std() < 3.0
    expect = sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, xs))
    for got in (a, c, d):
        assert np.abs(got - expect).max() / np.abs(expect).max() < 1e-6
    h.set_option(h.OPT_DECODE_NOISE, 0)
    assert np.array_equal(h.decrypt(src, n), plain)


def test_decode_noise_on_a_deep_context(shelfi, deep_oracles):
    o = deep_oracles["n16k_l4"]
    h = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0)
    h.loadCryptoParams()
    n = 2 * 8192 + 5
    x = np.random.default_rng(3).normal(0, 0.05, n)
    blob = o.encrypt(x, seed=4)
    plain = h.decrypt(blob, n)
    h.set_option(h.OPT_DECODE_NOISE, 1)
    o.set_option(o.OPT_DECODE_NOISE, 1)
    o.set_option(o.OPT_NOISE_SEED, 5)
    h.set_sampler_seed(5)
    got = h.decrypt(blob, n)
    h.set_sampler_seed(None)
    ref = o.decrypt(blob, n)
    o.set_option(o.OPT_DECODE_NOISE, 0)
    assert not np.array_equal(got, plain)
    assert np.abs(got - ref).max() < 1e-3 * np.abs(ref - plain).max()
    assert np.abs(got - x).max() < 1e-6


def test_wire_variant_option_for_generated_contexts(shelfi, tmp_path):
    sizes = {}
    for v in (8, 9):
        d = tmp_path / f"v{v}"
        d.mkdir()
        g = shelfi.CKKS("ckks", 4096, 52, str(d) + "/", device=0)
        g.set_option(g.OPT_WIRE_VARIANT, v)
        assert g.genCryptoContextAndKeyGen() == 1
        sizes[v] = (d / "cryptocontext.txt").stat().st_size
        g.loadCryptoParams()
        assert g.load_status == 0 and g.get_option(g.OPT_WIRE_VARIANT) == v
    assert sizes == {8: 2647, 9: 2651}            # the two fixture sizes of the reference
