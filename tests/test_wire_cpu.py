"""The product's wire-format code (fhe-fed_b200/csrc/wire.cpp) against the reference fixtures and
against blobs produced by the oracle — on the CPU, through the replay harness build of the same source."""
import ctypes as C

import numpy as np
import pytest

from conftest import FIXTURE_DIRS, GOLDEN


def reemit(emu, kind, data):
    n = C.c_uint64()
    err = C.create_string_buffer(256)
    p = emu.L.emu_wire_reemit(kind, data, len(data), C.byref(n), err, 256)
    assert p, err.value.decode()
    return emu.take(p, n.value)


def wire_map(emu, data, single=False):
    n_ct, limbs, stride, depth, level = C.c_uint64(), C.c_uint32(), C.c_uint64(), C.c_uint64(), C.c_uint64()
    sf = C.c_double()
    run0, run1 = np.zeros(16, np.uint64), np.zeros(16, np.uint64)
    tag, err = C.create_string_buffer(33), C.create_string_buffer(256)
    rc = emu.L.emu_wire_map(data, len(data), int(single), C.byref(n_ct), C.byref(limbs), run0.ctypes.data,
                            run1.ctypes.data, C.byref(stride), C.byref(depth), C.byref(level), C.byref(sf), tag, err, 256)
    assert rc == 0, err.value.decode()
    return dict(n_ct=n_ct.value, limbs=limbs.value, run0=run0, run1=run1, stride=stride.value, depth=depth.value,
                level=level.value, sf=sf.value, tag=tag.value.decode())


def test_fixture_files_roundtrip(emu, fixture_name):
    d = FIXTURE_DIRS[fixture_name]
    for kind, name in enumerate(["cryptocontext.txt", "key-public.txt", "key-private.txt"]):
        data = (d / name).read_bytes()
        assert reemit(emu, kind, data) == data


def test_truncated_and_corrupt_inputs_are_rejected(emu):
    data = (FIXTURE_DIRS["pybind"] / "key-public.txt").read_bytes()
    err = C.create_string_buffer(256)
    n = C.c_uint64()
    assert not emu.L.emu_wire_reemit(1, data[:-5], len(data) - 5, C.byref(n), err, 256)
    assert not emu.L.emu_wire_reemit(1, data + b"\0", len(data) + 1, C.byref(n), err, 256)
    bad = bytearray(data); bad[0] = 0
    assert not emu.L.emu_wire_reemit(1, bytes(bad), len(bad), C.byref(n), err, 256)
    assert not emu.L.emu_wire_reemit(2, data, len(data), C.byref(n), err, 256)  # public key is not a private key


@pytest.mark.parametrize("name", ["CT1.txt", "CT2.txt"])
def test_single_ciphertext_fixture_maps(emu, oracle_mod, name):
    data = (GOLDEN / "mkhe" / name).read_bytes()
    m = wire_map(emu, data, single=True)
    info, _ = oracle_mod.blob_parse(data, single=True, want_payload=False)
    assert (m["n_ct"], m["limbs"], m["depth"], m["level"]) == (1, 2, 1, 0)
    assert m["sf"] == 1125899906990081.0
    assert m["run0"][0] == info.first_data_offset
    assert m["tag"] == info.key_tag.decode()


@pytest.mark.parametrize("n", [1, 4096, 4097, 10000, 3 * 4096])
def test_oracle_blobs_map_and_rebuild(emu, oracles, oracle_mod, fixture_name, n):
    """map -> unpack (replayed kernel) -> plan + pack (replayed kernel) reproduces the oracle's bytes."""
    o = oracles[fixture_name]
    x = np.random.default_rng(n).random(n)
    blob = o.encrypt(x, seed=3)
    m = wire_map(emu, blob)
    info, _ = oracle_mod.blob_parse(blob, want_payload=False)
    assert m["n_ct"] == info.n_ct == -(-n // 4096)
    assert m["run0"][0] == info.first_data_offset
    if info.n_ct > 1:
        assert m["stride"] == info.ct_stride
    size = emu.L.emu_wire_blob_size((FIXTURE_DIRS[fixture_name] / "cryptocontext.txt").read_bytes(),
                                    len((FIXTURE_DIRS[fixture_name] / "cryptocontext.txt").read_bytes()), info.n_ct)
    assert size == len(blob) == o.blob_size(info.n_ct)
    out_n = C.c_uint64()
    err = C.create_string_buffer(256)
    p = emu.L.emu_wire_rebuild(blob, len(blob), C.byref(out_n), err, 256)
    assert p, err.value.decode()
    assert emu.take(p, out_n.value) == blob


def test_depth2_blob_rebuild(emu, oracles):
    o = oracles["code"]
    rng = np.random.default_rng(2)
    blobs = [o.encrypt(rng.random(9000), seed=i) for i in range(2)]
    avg = o.computeWeightedAverage(blobs, [0.25, 0.75])
    m = wire_map(emu, avg)
    assert (m["depth"], m["level"]) == (2, 0)
    assert m["sf"] == o.scaling_factor * o.scaling_factor
    out_n = C.c_uint64()
    err = C.create_string_buffer(256)
    p = emu.L.emu_wire_rebuild(avg, len(avg), C.byref(out_n), err, 256)
    assert p, err.value.decode()
    assert emu.take(p, out_n.value) == avg


def test_parameter_rules(emu):
    M = 16384
    assert emu.L.emu_first_prime(52, M) == 0x0010000000060001
    assert emu.L.emu_previous_prime(emu.L.emu_first_prime(60, M), M) == 0x0FFFFFFFFFFFC001
    assert emu.L.emu_min_primitive_root(M, 0x0FFFFFFFFFFFC001) == 0x0000179C0F8FADCC
    assert emu.L.emu_min_primitive_root(M, 0x0010000000060001) == 0x0000010D0EF11890
    assert emu.L.emu_first_prime(50, M) == 0x4000000024001
    for bits in (14, 20, 33, 40, 52):  # code/benchmark_crypto.py:123-124 sweep
        assert emu.L.emu_select_ring_dim(bits) == 8192


def test_mutated_blobs_never_map_out_of_bounds(emu, oracles):
    """Ciphertext blobs arrive from other parties: a corrupted or truncated one must be rejected or map to runs that
    lie inside the buffer (the device copies whole runs from those offsets) — and must never crash the parser."""
    o = oracles["pybind"]
    blob = o.encrypt(np.random.default_rng(0).random(3 * 4096 + 5), seed=1)
    rng = np.random.default_rng(1)
    hdr, accepted = 3000, 0
    for it in range(1200):
        b = bytearray(blob)
        mode = it % 4
        if mode == 0:      # one header byte
            b[int(rng.integers(0, hdr))] = int(rng.integers(0, 256))
        elif mode == 1:    # a 64-bit field (counts, lengths, moduli)
            pos = int(rng.integers(0, hdr - 8))
            b[pos:pos + 8] = int(rng.integers(0, 2**63)).to_bytes(8, "little")
        elif mode == 2:    # truncation
            b = b[:int(rng.integers(0, len(b)))]
        else:              # later ciphertexts' headers / payload
            b[int(rng.integers(len(b) - 700000, len(b)))] ^= 0xFF
        data = bytes(b)
        n_ct, limbs, stride, depth, level = C.c_uint64(), C.c_uint32(), C.c_uint64(), C.c_uint64(), C.c_uint64()
        sf = C.c_double()
        run0, run1 = np.zeros(16, np.uint64), np.zeros(16, np.uint64)
        tag, err = C.create_string_buffer(33), C.create_string_buffer(256)
        rc = emu.L.emu_wire_map(data, len(data), 0, C.byref(n_ct), C.byref(limbs), run0.ctypes.data, run1.ctypes.data,
                                C.byref(stride), C.byref(depth), C.byref(level), C.byref(sf), tag, err, 256)
        if rc != 0:
            assert err.value
            continue
        accepted += 1
        runs = 2 * limbs.value
        last = int(run0[runs - 1]) if n_ct.value == 1 else int(run1[runs - 1]) + (n_ct.value - 2) * stride.value
        assert last + 8192 * 8 <= len(data), (it, mode)
    assert accepted > 100      # payload flips are legitimate blobs; header damage is what gets rejected


def test_mutated_key_files_are_rejected_or_reemitted(emu):
    rng = np.random.default_rng(2)
    for kind, name in enumerate(["cryptocontext.txt", "key-public.txt", "key-private.txt"]):
        data = (FIXTURE_DIRS["code"] / name).read_bytes()
        for it in range(300):
            b = bytearray(data)
            if it % 3 == 0:
                b = b[:int(rng.integers(0, len(b)))]
            elif it % 3 == 1:
                b[int(rng.integers(0, min(len(b), 2700)))] = int(rng.integers(0, 256))
            else:
                pos = int(rng.integers(0, min(len(b), 2700) - 8))
                b[pos:pos + 8] = int(rng.integers(0, 2**63)).to_bytes(8, "little")
            n, err = C.c_uint64(), C.create_string_buffer(256)
            p = emu.L.emu_wire_reemit(kind, bytes(b), len(b), C.byref(n), err, 256)
            if p:
                emu.take(p, n.value)
            else:
                assert err.value
