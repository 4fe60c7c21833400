"""The CPU oracle against the reference's binary fixtures (KATs K1-K4).

These are the only hard pins the reference offers (it has no asserting tests): serialized context /
key pairs from palisade_pybind/SHELFI_FHE/resources/cryptoparams and code/resources/cryptoparams, and
two ciphertexts from code/mkhe/build.  An oracle that passes them has the right modulus chain, roots,
NTT ordering, key relation and wire grammar.
"""
import numpy as np
import pytest

from conftest import FIXTURE_DIRS, GOLDEN

Q0 = 0x0FFFFFFFFFFFC001
Q1 = 0x0010000000060001
PSI0 = 0x0000179C0F8FADCC
PSI1 = 0x0000010D0EF11890
TAGS = {"pybind": "83cfbab56f90dc0cfcc65516d614c9ce", "code": "a0e242e9f5e2686f01c283cd25a4cc22"}
E_HEAD = {"pybind": [-4, 2, 0, -1, -2, 3, -6, 0], "code": [-1, -1, -4, 0, 4, -2, 0, -3]}
S_HEAD = {"pybind": [1, 1, 0, 1, -1, 1, 0, 0, -1, 1, 1, -1], "code": [-1, -1, 0, 0, 1, 0, 1, 1, -1, 1, -1, 0]}


def centre(a, q):
    a = a.astype(object)
    return np.array([int(v) - q if int(v) > q // 2 else int(v) for v in a], dtype=np.int64)


def test_context_fields(oracles, fixture_name):
    o = oracles[fixture_name]
    assert (o.ring_dim, o.num_limbs, o.slots) == (8192, 2, 4096)
    assert (o.modulus(0), o.modulus(1)) == (Q0, Q1)
    assert (o.psi(0), o.psi(1)) == (PSI0, PSI1)
    assert o.scaling_factor == float(Q1)
    assert o.wire_variant == (8 if fixture_name == "pybind" else 9)
    assert o.key_tag() == TAGS[fixture_name]


def test_k1_files_reemit_byte_identical(oracle_mod, fixture_name):
    d = FIXTURE_DIRS[fixture_name]
    for kind, name in enumerate(["cryptocontext.txt", "key-public.txt", "key-private.txt"]):
        data = (d / name).read_bytes()
        assert oracle_mod.file_reemit(data, kind) == data


def test_k2_key_relation(oracles, fixture_name):
    """INTT(b + a*s) is one small error polynomial, identical in both limbs; INTT(s) is ternary."""
    o = oracles[fixture_name]
    b, a = o.public_key()
    s = o.secret_key()
    errs = []
    for l in range(2):
        q = o.modulus(l)
        t = np.array([(int(b[l, j]) + int(a[l, j]) * int(s[l, j])) % q for j in range(8192)], dtype=np.uint64)
        e = centre(o.ntt_inverse(l, t), q)
        assert np.abs(e).max() <= 12
        errs.append(e)
        sc = centre(o.ntt_inverse(l, s[l]), q)
        assert set(np.unique(sc)) <= {-1, 0, 1}
        assert list(sc[:12]) == S_HEAD[fixture_name]
    assert np.array_equal(errs[0], errs[1])
    assert list(errs[0][:8]) == E_HEAD[fixture_name]


def test_k3_parameter_rules(oracle_mod):
    M = 16384
    assert oracle_mod.first_prime(52, M) == Q1
    assert oracle_mod.previous_prime(oracle_mod.first_prime(60, M), M) == Q0
    assert oracle_mod.min_root_of_unity(M, Q0) == PSI0
    assert oracle_mod.min_root_of_unity(M, Q1) == PSI1
    q50 = oracle_mod.first_prime(50, M)
    assert q50 == 0x4000000024001
    assert oracle_mod.min_root_of_unity(M, q50) == 0xDE4616B52
    assert oracle_mod.select_ring_dim(52, 4096) == 8192


@pytest.mark.parametrize("name", ["CT1.txt", "CT2.txt"])
def test_k4_ciphertext_fixture(oracle_mod, name):
    data = (GOLDEN / "mkhe" / name).read_bytes()
    info, payload = oracle_mod.blob_parse(data, single=True)
    assert (info.n_ct, info.n_polys, info.n_limbs, info.ring_dim) == (1, 2, 2, 8192)
    assert (info.depth, info.level, info.encoding_type) == (1, 0, 4)
    assert info.scaling_factor == 1125899906990081.0 == float(0x4000000024001)
    assert info.key_tag.decode().startswith("750b9975")
    assert (payload[0, :, 0] < Q0).all() and (payload[0, :, 1] < 0x4000000024001).all()
    assert oracle_mod.blob_reemit(data, single=True) == data


def test_ntt_roundtrip_and_convolution(oracles):
    """forward/inverse are inverses, and pointwise products are negacyclic convolutions."""
    o = oracles["pybind"]
    rng = np.random.default_rng(5)
    for l in range(2):
        q = o.modulus(l)
        a = rng.integers(0, q, 8192, dtype=np.uint64)
        assert np.array_equal(o.ntt_inverse(l, o.ntt_forward(l, a)), a)
        # x * x^(N-1) = -1 in Z[x]/(x^N + 1)
        x1 = np.zeros(8192, np.uint64); x1[1] = 1
        xn = np.zeros(8192, np.uint64); xn[8191] = 1
        f1, fn = o.ntt_forward(l, x1), o.ntt_forward(l, xn)
        prod = np.array([int(u) * int(v) % q for u, v in zip(f1, fn)], dtype=np.uint64)
        back = o.ntt_inverse(l, prod)
        assert back[0] == q - 1 and not back[1:].any()


def test_oracle_end_to_end_matches_plaintext(oracles, fixture_name):
    """pythonApi/ckks_example.py turned into an asserting test (3 learners, weights .5/.2/.3)."""
    o = oracles[fixture_name]
    rng = np.random.default_rng(1)
    n = 10000
    xs = [rng.random(n) for _ in range(3)]
    w = [0.5, 0.2, 0.3]
    blobs = [o.encrypt(x, seed=11 + i) for i, x in enumerate(xs)]
    out = o.decrypt(o.computeWeightedAverage(blobs, w), n)
    expect = sum(float(np.float32(wi)) * x for wi, x in zip(w, xs))
    assert np.abs(out - expect).max() / np.abs(expect).max() < 1e-6
    assert o.computeWeightedAverage(blobs, w[:2]) == b""
