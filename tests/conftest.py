"""Shared fixtures.  `-m "not gpu"` runs everywhere; `-m gpu` needs a B200 and the built library."""
from __future__ import annotations

import ctypes as C
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
GOLDEN = ROOT / "tests" / "golden"
for p in (ROOT, ROOT / "fhe-fed_b200"):
    if str(p) not in sys.path:
        sys.path.insert(0, str(p))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200) and libshelfi_b200.so")


FIXTURE_DIRS = {"pybind": GOLDEN / "pybind_cryptoparams", "code": GOLDEN / "code_cryptoparams"}


@pytest.fixture(scope="session")
def oracle_mod():
    import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session", params=["pybind", "code"])
def fixture_name(request):
    return request.param


@pytest.fixture(scope="session")
def oracles(oracle_mod):
    out = {}
    for name, d in FIXTURE_DIRS.items():
        o = oracle_mod.Oracle(cryptodir=str(d) + "/")
        assert o.loadCryptoParams() == 0
        out[name] = o
    return out


class Emu:
    """ctypes front end of tests/emu (CPU replay of the device phases; test infrastructure)."""

    def __init__(self):
        d = ROOT / "tests" / "emu"
        subprocess.run(["make", "-s", "-C", str(d)], check=True, stdout=subprocess.DEVNULL)
        L = C.CDLL(str(d / "_build" / "libshelfi_emu.so"))
        vp, u32, u64, f64 = C.c_void_p, C.c_uint32, C.c_uint64, C.c_double
        sig = {
            "emu_create": (vp, [vp, vp, u32, u32, u32, f64]),
            "emu_destroy": (None, [vp]),
            "emu_error": (C.c_char_p, [vp]),
            "emu_fwd_mask": (u32, [vp, u32]),
            "emu_rmode": (u32, [vp, u32]),
            "emu_ntt_forward": (None, [vp, u32, vp]),
            "emu_ntt_inverse": (None, [vp, u32, vp]),
            "emu_encode": (None, [vp, vp, u64, vp]),
            "emu_set_selection": (None, [vp]),
            "emu_encode_sample": (None, [vp, vp, u64, u64, u64, vp, vp]),
            "emu_encrypt": (None, [vp, vp, vp, vp, vp, vp]),
            "emu_chacha20_block": (None, [vp, u32, vp, vp]),
            "emu_keygen_draws": (None, [vp, u64, vp, vp, vp]),
            "emu_gauss_lazy": (C.c_int, [vp, u32, u32, C.POINTER(C.c_int)]),
            "emu_ternary16": (C.c_int, [u32, C.c_int, C.POINTER(C.c_int)]),
            "emu_retry_ternary": (C.c_int, [vp, u64, u32, u32]),
            "emu_gauss_table": (None, [vp, vp]),
            "emu_decrypt": (None, [vp, vp, vp, f64, f64, vp, u64, vp, vp, vp]),
            "emu_create_ex": (vp, [vp, vp, u32, u32, u32, u32, f64]),
            "emu_weighted_sum": (None, [vp, u32, C.POINTER(vp), vp, u32, vp, u32]),
            "emu_unpack": (None, [vp, vp, u64, u32, u64, vp, vp]),
            "emu_pack": (None, [vp, vp, u64, u32, u64, vp, vp]),
            "emu_wire_reemit": (vp, [C.c_int, C.c_char_p, u64, C.POINTER(u64), C.c_char_p, C.c_size_t]),
            "emu_wire_map": (C.c_int, [C.c_char_p, u64, C.c_int, C.POINTER(u64), C.POINTER(u32), vp, vp, C.POINTER(u64),
                                       C.POINTER(u64), C.POINTER(u64), C.POINTER(f64), C.c_char_p, C.c_char_p,
                                       C.c_size_t]),
            "emu_wire_rebuild": (vp, [C.c_char_p, u64, C.POINTER(u64), C.c_char_p, C.c_size_t]),
            "emu_wire_blob_size": (u64, [C.c_char_p, u64, u64]),
            "emu_wire_blob_size_level": (u64, [C.c_char_p, u64, u64, u32]),
            "emu_rescale": (None, [vp, vp, vp]),
            "emu_free": (None, [vp]),
            "emu_first_prime": (u64, [u32, u32]),
            "emu_previous_prime": (u64, [u64, u32]),
            "emu_min_primitive_root": (u64, [u32, u64]),
            "emu_limb_fp64": (C.c_int, [vp, u32]),
            "emu_set_fp64": (None, [vp, C.c_int]),
            "emu_select_ring_dim": (u32, [u32]),
            "emu_select_ring_dim_ex": (u32, [u32, u32]),
            "emu_modulus_chain": (C.c_int, [u32, u32, u32, vp]),
        }
        for name, (res, args) in sig.items():
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        self.L = L

    def context(self, q, psi, slots=4096, batch=4096, sigma=float(np.float32(3.19)), logn=13):
        q = np.ascontiguousarray(q, np.uint64)
        psi = np.ascontiguousarray(psi, np.uint64)
        h = self.L.emu_create_ex(q.ctypes.data, psi.ctypes.data, len(q), logn, slots, batch, sigma)
        err = self.L.emu_error(h).decode()
        if err:
            self.L.emu_destroy(h)
            raise ValueError(err)
        return h

    def take(self, ptr, n):
        try:
            return C.string_at(ptr, n)
        finally:
            self.L.emu_free(ptr)


@pytest.fixture(scope="session")
def emu():
    return Emu()


@pytest.fixture(scope="session")
def emu_ctx(emu, oracles):
    """Replay contexts built from the two fixture contexts."""
    out = {}
    for name, o in oracles.items():
        L = o.num_limbs
        out[name] = emu.context([o.modulus(l) for l in range(L)], [o.psi(l) for l in range(L)], o.slots, 4096)
    return out


# Contexts beyond what the reference's hard-coded multDepth = 1 produces (BASELINE configs 3 and 4, plus a deep chain
# at ring dimension 8192 and sparse packing in a large ring): name -> (batch/slots, scale bits, multDepth, ring_dim)
DEEP_CONTEXTS = {
    "n16k_l4": (8192, 52, 3, 0),        # config 3: N = 16384, 4 limbs, full packing
    "n32k_l8": (16384, 52, 7, 0),       # config 4: N = 32768, 8 limbs, full packing
    "n8k_l3": (4096, 40, 2, 8192),      # 3 limbs in one tile: the general CRT without the tile split
    "n16k_sparse": (1024, 45, 3, 16384),  # slots < 4096 in a two-tile ring: gap 8
}


@pytest.fixture(scope="session")
def deep_oracles(oracle_mod, tmp_path_factory):
    """Oracle-generated key material (fixed seeds) for DEEP_CONTEXTS, loaded."""
    out = {}
    for i, (name, (batch, bits, depth, ring)) in enumerate(DEEP_CONTEXTS.items()):
        d = str(tmp_path_factory.mktemp(name)) + "/"
        o = oracle_mod.Oracle("ckks", batch, bits, d)
        assert o.genCryptoContextAndKeyGen(seed=100 + i, nine_int=True, mult_depth=depth, ring_dim=ring) == 1
        assert o.loadCryptoParams() == 0
        out[name] = o
    return out


@pytest.fixture(scope="session")
def deep_emu_ctx(emu, deep_oracles):
    out = {}
    for name, o in deep_oracles.items():
        L = o.num_limbs
        out[name] = emu.context([o.modulus(l) for l in range(L)], [o.psi(l) for l in range(L)], o.slots, o.batchSize,
                                logn=int(o.ring_dim).bit_length() - 1)
    return out


def has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


@pytest.fixture(scope="session")
def shelfi():
    import SHELFI_FHE
    return SHELFI_FHE


@pytest.fixture(scope="session")
def gpu_ckks(shelfi):
    """CKKS handles on cuda:0 for the two fixture key sets."""
    if not has_gpu():
        pytest.skip("no CUDA device")
    out = {}
    for name, d in FIXTURE_DIRS.items():
        h = shelfi.CKKS("ckks", 4096, 52, str(d) + "/", device=0)
        h.loadCryptoParams()
        assert h.load_status == 0
        out[name] = h
    return out
