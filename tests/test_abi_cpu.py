"""The C-ABI library: loads, exports every symbol include/shelfi_b200.h declares, and refuses to compute
without a GPU (no CPU fallback).  No compute calls here."""
import ctypes as C
import re
import subprocess

import pytest

from conftest import FIXTURE_DIRS, ROOT, has_gpu

HEADER = ROOT / "include" / "shelfi_b200.h"
LIB = ROOT / "fhe-fed_b200" / "lib" / "libshelfi_b200.so"


def declared_symbols():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(shelfi_[a-z0-9_]+)\s*\(", text)))


def test_library_is_built():
    assert LIB.exists(), "run __graft_entry__.build() or make -C fhe-fed_b200"


def test_every_declared_symbol_is_exported():
    lib = C.CDLL(str(LIB))
    names = declared_symbols()
    assert len(names) >= 35
    missing = [n for n in names if not hasattr(lib, n)]
    assert not missing, missing


def test_library_has_sm100a_code_and_no_torch_dependency():
    out = subprocess.run(["cuobjdump", "-lelf", str(LIB)], capture_output=True, text=True).stdout
    assert "sm_100a" in out
    needed = subprocess.run(["readelf", "-d", str(LIB)], capture_output=True, text=True).stdout
    assert "libcudart" in needed and "torch" not in needed and "oracle" not in needed


def test_shim_surface_matches_reference_binding(shelfi):
    # binding.cpp:16-31,45-49
    assert issubclass(shelfi.CKKS, shelfi.Scheme)
    for name in ["loadCryptoParams", "genCryptoContextAndKeyGen", "encrypt", "encrypt_cpp", "decrypt", "decrypt_cpp",
                 "computeWeightedAverage", "computeWeightedAverage_cpp"]:
        assert callable(getattr(shelfi.CKKS, name))
    assert shelfi.CKKS.encrypt_cpp is shelfi.CKKS.encrypt
    assert isinstance(shelfi.__version__, str)
    h = shelfi.CKKS()
    assert (h.scheme, h.batchSize, h.scaleFactorBits, h.cryptodir) == ("ckks", 4096, 52, "../resources/cryptoparams/")
    h = shelfi.CKKS(scheme="ckks", batchSize=1024, scaleFactorBits=40, cryptodir="/tmp/x/")
    assert (h.batchSize, h.scaleFactorBits, h.cryptodir) == (1024, 40, "/tmp/x/")
    # size mismatch: prints and returns b"" like ckks.cpp:265-268, before touching the device
    assert h.computeWeightedAverage([b"a", b"b"], [1.0]) == b""


@pytest.mark.skipif(has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback(shelfi):
    h = shelfi.CKKS(cryptodir=str(FIXTURE_DIRS["pybind"]) + "/")
    with pytest.raises(shelfi.ShelfiError) as e:
        h.loadCryptoParams()
    assert e.value.code == -5 and "no CPU path" in str(e.value)
    with pytest.raises(shelfi.ShelfiError):
        h.encrypt([1.0, 2.0])
    with pytest.raises(shelfi.ShelfiError):
        h.decrypt(b"\x01", 1)
    assert h.genCryptoContextAndKeyGen() == 0
    # the peer-buffer / decrypt-gather entry points too: no device, no work
    import ctypes as C
    lib = shelfi.lib()
    p, handle = C.c_void_p(), C.create_string_buffer(64)
    assert lib.shelfi_peer_buffer_create(h.handle, 4096, C.byref(p), handle) == -5 and not p.value
    assert lib.shelfi_peer_buffer_open(h.handle, handle, C.byref(p)) == -5
    outs = (C.c_void_p * 2)(8, 16)
    assert lib.shelfi_decrypt_gather_device(h.handle, C.c_void_p(8), 1, 1.0, outs, 2, 0, 1, None) == -5


def test_copy_threads_selftest(shelfi):
    """The staging copy threads (spin-then-sleep hand-off) on host memory only: many back-to-back copies, odd sizes,
    1 to 8 workers, and copies far enough apart that the workers have gone back to sleep."""
    import time
    lib = shelfi.lib()
    for threads in (1, 2, 3, 8):
        assert lib.shelfi_selftest_parallel_copy((5 << 20) + 12345, threads, 40) == 0
        assert lib.shelfi_selftest_parallel_copy((1 << 20) + 1, threads, 5) == 0      # below the split threshold
    for _ in range(3):
        time.sleep(0.01)
        assert lib.shelfi_selftest_parallel_copy(16 << 20, 8, 3) == 0
    assert lib.shelfi_selftest_parallel_copy(0, 8, 1) != 0


def test_product_never_imports_the_oracle():
    for path in list((ROOT / "fhe-fed_b200").rglob("*.py")) + list((ROOT / "fhe-fed_b200" / "csrc").iterdir()):
        if path.is_file():
            text = path.read_text(errors="ignore")
            assert "shelfi_oracle" not in text and "import oracle" not in text and "tests/emu" not in text.replace(
                "tests/emu/ (g++", "").replace("harness (tests/emu)", ""), path
