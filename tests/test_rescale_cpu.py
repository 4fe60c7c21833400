"""Rescale-before-transport on the CPU: the oracle's restatement of PALISADE's ModReduce under
EXACTRESCALE (DCRTPoly::DropLastElementAndScale), the wire form of a rescaled ciphertext through both the oracle's
and the product's parser, and the device phases replayed thread by thread (tests/emu) against the oracle.

The reference never rescales (ckks.cpp:287-303 leaves the average at depth 2) and holds no fixture of a rescaled
ciphertext, so this tier is "parity unpinned": it pins the product to the oracle and the oracle to the algebra."""
import ctypes as C

import numpy as np
import pytest

from test_wire_cpu import wire_map


def vp(a):
    return C.c_void_p(a.ctypes.data)


def _average(o, n, k=3, seed=0, amp=1.0):
    rng = np.random.default_rng(seed)
    xs = [rng.normal(0, amp, n) for _ in range(k)]
    w = rng.dirichlet(np.ones(k)).astype(np.float32)
    blobs = [o.encrypt(x, seed=11 + i) for i, x in enumerate(xs)]
    ref = sum(np.float64(wi) * x for wi, x in zip(w, xs))
    return o.computeWeightedAverage(blobs, [float(v) for v in w]), ref


def test_constants_are_what_the_algebra_says(oracles, deep_oracles):
    """PALISADE's QlQlInvModqlDivqlModq[i] = floor(P * (P^-1 mod q_l) / q_l) mod q_i equals -q_l^-1 mod q_i, which is
    what lets the device form (c_i - NTT_i(t)) * q_l^-1 instead of two products."""
    for o in list(oracles.values()) + list(deep_oracles.values()):
        L = o.num_limbs
        q = [o.modulus(i) for i in range(L)]
        for lin in range(2, L + 1):
            t, inv = o.rescale_constants(lin)
            for i in range(lin - 1):
                assert inv[i] * (q[lin - 1] % q[i]) % q[i] == 1
                assert t[i] == q[i] - inv[i]
            # and the definition itself, in Python integers
            P = 1
            for j in range(lin - 1):
                P *= q[j]
            want = (P * pow(P, -1, q[lin - 1])) // q[lin - 1]
            assert t == [want % q[i] for i in range(lin - 1)]


def test_rescaled_ciphertext_is_the_rounded_quotient(oracles):
    """CRT integer of the rescaled ciphertext's decryption == (c - [c]_{q_l}) / q_l of the original's, coefficient by
    coefficient: both polynomials are divided exactly, so m' = c0' + c1' s = (m - ([c0] + [c1] s)) / q_l up to the
    small term the centred remainders contribute — checked here on the decoded values instead."""
    o = oracles["pybind"]
    avg, ref = _average(o, 5000, seed=1)
    _, payload = __import__("oracle").blob_parse(avg)
    ct = payload[1]
    r = o.rescale_ct(ct)
    assert r.shape == (2, 1, o.ring_dim) and (r < o.modulus(0)).all()
    sf2 = o.scaling_factor * o.scaling_factor
    d0 = o.decrypt_ct(ct, sf2)
    d1, _ = o.decrypt_ct_level(r, sf2 / o.modulus(1))
    assert np.abs(d1 - d0).max() < 1e-9


@pytest.mark.parametrize("name", ["pybind", "code"])
def test_blob_rescale_round_trip(oracles, oracle_mod, name):
    o = oracles[name]
    n = 3 * 4096 + 17
    avg, ref = _average(o, n, seed=2)
    r = o.rescale(avg)
    info, payload = oracle_mod.blob_parse(r)
    assert (info.n_ct, info.n_limbs, info.depth, info.level) == (4, 1, 1, 1)
    assert info.scaling_factor == o.scaling_factor * o.scaling_factor / o.modulus(1)
    assert oracle_mod.blob_reemit(r) == r
    assert len(r) < 0.53 * len(avg)                     # one tower of two, plus the in-line parameter objects
    out = o.decrypt(r, n)
    assert np.abs(out - ref).max() < 1e-8
    assert np.abs(out - o.decrypt(avg, n)).max() < 1e-9


def test_deep_chain_rescale(deep_oracles, oracle_mod):
    """4 and 3 towers: the dropped tower is switched into every remaining modulus (primes above and below it)."""
    for name in ("n8k_l3", "n16k_l4"):
        o = deep_oracles[name]
        n = o.batchSize + 5
        avg, ref = _average(o, n, k=2, seed=3)
        r = o.rescale(avg)
        info, _ = oracle_mod.blob_parse(r)
        L = o.num_limbs
        assert (info.n_limbs, info.depth, info.level) == (L - 1, 1, 1)
        assert oracle_mod.blob_reemit(r) == r
        assert np.abs(o.decrypt(r, n) - ref).max() < 1e-6
        # a second rescale is refused at depth 1 by the product; the oracle's arithmetic still drops a tower correctly
        _, payload = oracle_mod.blob_parse(r)
        r2 = o.rescale_ct(payload[0])
        assert r2.shape == (2, L - 2, o.ring_dim)


@pytest.mark.parametrize("name", ["pybind", "code"])
def test_product_parser_maps_and_rebuilds_rescaled_blobs(emu, oracles, name):
    """wire.cpp reads the in-line reduced parameter object and re-creates the blob byte for byte."""
    o = oracles[name]
    avg, _ = _average(o, 2 * 4096 + 100, seed=4)
    r = o.rescale(avg)
    m = wire_map(emu, r)
    assert (m["n_ct"], m["limbs"], m["depth"], m["level"]) == (3, 1, 1, 1)
    assert m["sf"] == o.scaling_factor * o.scaling_factor / o.modulus(1)
    n = C.c_uint64()
    err = C.create_string_buffer(256)
    p = emu.L.emu_wire_rebuild(r, len(r), C.byref(n), err, 256)
    assert p, err.value.decode()
    assert emu.take(p, n.value) == r
    ctx = (__import__("conftest").FIXTURE_DIRS[name] / "cryptocontext.txt").read_bytes()
    for n_ct in (1, 2, 3, 50):
        assert emu.L.emu_wire_blob_size_level(ctx, len(ctx), n_ct, 2) == emu.L.emu_wire_blob_size(ctx, len(ctx), n_ct)
    assert emu.L.emu_wire_blob_size_level(ctx, len(ctx), 3, 1) == len(r)


def test_product_parser_rejects_inconsistent_rescaled_blobs(emu, oracles):
    o = oracles["pybind"]
    avg, _ = _average(o, 4096 + 1, seed=5)
    r = bytearray(o.rescale(avg))
    m = wire_map(emu, bytes(r))
    err = C.create_string_buffer(256)
    args = [C.c_uint64(), C.c_uint32(), np.zeros(16, np.uint64), np.zeros(16, np.uint64), C.c_uint64(), C.c_uint64(),
            C.c_uint64(), C.c_double(), C.create_string_buffer(33)]

    def maps(data):
        return emu.L.emu_wire_map(bytes(data), len(data), 0, C.byref(args[0]), C.byref(args[1]), args[2].ctypes.data,
                                  args[3].ctypes.data, C.byref(args[4]), C.byref(args[5]), C.byref(args[6]),
                                  C.byref(args[7]), args[8], err, 256) == 0

    assert maps(r)
    # the reduced modulus written in line must be q_0: flip one of its words (it follows c0's tower by 4+4+4+4+1 bytes
    # of format / marker / id / ring fields)
    after_c0 = int(m["run0"][0]) + 8 * o.ring_dim + 8 + 4 + 4 + 4      # q, format, SAME, REF(native)
    bad = bytearray(r)
    bad[after_c0 + 4 + 4 + 4 + 4 + 4 + 1 + 440] ^= 1                    # format SAME NEW M N pow2, last big-int byte
    assert not maps(bad)
    assert not maps(r[:-1]) and not maps(r + b"\0")


@pytest.mark.parametrize("name", ["pybind", "code"])
def test_replayed_rescale_kernels_match_oracle(emu, emu_ctx, oracles, oracle_mod, name):
    o, h = oracles[name], emu_ctx[name]
    avg, _ = _average(o, 4096 * 2, seed=6, amp=20.0)
    _, payload = oracle_mod.blob_parse(avg)
    rng = np.random.default_rng(9)
    q = [o.modulus(i) for i in range(2)]
    worst = np.stack([np.stack([rng.integers(0, q[l], o.ring_dim, dtype=np.uint64) for l in range(2)]) for _ in range(2)])
    worst[0, 1, :4] = [0, q[1] - 1, q[1] >> 1, (q[1] >> 1) + 1]        # the centring boundary of SwitchModulus
    for ct in (payload[0], payload[1], worst):
        ct = np.ascontiguousarray(ct)
        out = np.empty((2, 1, o.ring_dim), np.uint64)
        emu.L.emu_rescale(h, vp(ct), vp(out))
        assert np.array_equal(out, o.rescale_ct(ct))


@pytest.mark.parametrize("name", ["n8k_l3", "n16k_l4", "n16k_sparse", "n32k_l8"])
def test_replayed_rescale_kernels_match_oracle_on_deep_contexts(emu, deep_emu_ctx, deep_oracles, name):
    """Rings above one tile (head loader + tail pass) and several remaining towers, FP64- and integer-pipe limbs."""
    o, h = deep_oracles[name], deep_emu_ctx[name]
    L, N = o.num_limbs, o.ring_dim
    rng = np.random.default_rng(10)
    ct = np.stack([np.stack([rng.integers(0, o.modulus(l), N, dtype=np.uint64) for l in range(L)]) for _ in range(2)])
    out = np.empty((2, L - 1, N), np.uint64)
    emu.L.emu_rescale(h, vp(ct), vp(out))
    assert np.array_equal(out, o.rescale_ct(ct))
    emu.L.emu_set_fp64(h, 0)
    try:
        out2 = np.empty_like(out)
        emu.L.emu_rescale(h, vp(ct), vp(out2))
    finally:
        emu.L.emu_set_fp64(h, 1)
    assert np.array_equal(out2, out)
