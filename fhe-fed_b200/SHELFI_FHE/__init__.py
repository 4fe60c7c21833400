"""SHELFI_FHE — drop-in for the reference's pybind module of the same name, backed by the B200 engine.

Mirrors /root/reference/palisade_pybind/SHELFI_FHE/src/binding.cpp:14-50: classes ``Scheme`` and
``CKKS(scheme="ckks", batchSize=4096, scaleFactorBits=52, cryptodir="../resources/cryptoparams/")``
with ``loadCryptoParams``, ``genCryptoContextAndKeyGen``, ``encrypt``, ``decrypt``,
``computeWeightedAverage`` and the three ``*_cpp`` aliases (binding.cpp:27,29,31 bind those names to
the same methods), plus ``__version__``.

All arithmetic happens in ``libshelfi_b200.so`` (hand-written sm_100a CUDA behind the C ABI of
``include/shelfi_b200.h``), loaded with ctypes.  There is no CPU path: without the library or
without a CUDA device the calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import sys
from pathlib import Path

import numpy as np

__version__ = "0.1.0-b200"

_PKG = Path(__file__).resolve().parent.parent
_LIB_PATH = Path(os.environ.get("SHELFI_B200_LIB", _PKG / "lib" / "libshelfi_b200.so"))

ERR_NAMES = {0: "OK", -1: "ARG", -2: "NOT_LOADED", -3: "FORMAT", -4: "PARAMS", -5: "CUDA", -6: "IO", -7: "UNSUPPORTED",
             -8: "MEMORY"}


class ShelfiError(RuntimeError):
    def __init__(self, code: int, message: str):
        super().__init__(f"[{ERR_NAMES.get(code, code)}] {message}")
        self.code = code


_lib = None


def lib() -> C.CDLL:
    """Load libshelfi_b200.so (built by fhe-fed_b200/Makefile or __graft_entry__.build())."""
    global _lib
    if _lib is not None:
        return _lib
    if not _LIB_PATH.exists():
        raise ImportError(f"{_LIB_PATH} not found: build it with `make -C {_PKG}` (nvcc, sm_100a); "
                          "SHELFI_FHE has no CPU fallback")
    L = C.CDLL(str(_LIB_PATH))
    vp, u8p, u32, u64, i64, f64 = C.c_void_p, C.POINTER(C.c_uint8), C.c_uint32, C.c_uint64, C.c_int64, C.c_double
    u64p = C.POINTER(C.c_uint64)
    sig = {
        "shelfi_ckks_create": (C.c_int, [C.c_char_p, u32, u32, C.c_char_p, C.c_int, C.POINTER(vp)]),
        "shelfi_ckks_destroy": (None, [vp]),
        "shelfi_last_error": (C.c_char_p, [vp]),
        "shelfi_load_crypto_params": (C.c_int, [vp]),
        "shelfi_load_crypto_params_from_memory": (C.c_int, [vp, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t,
                                                             C.c_char_p, C.c_size_t]),
        "shelfi_gen_crypto_context_and_keygen": (C.c_int, [vp]),
        "shelfi_gen_crypto_context_and_keygen_ex": (C.c_int, [vp, u32, u32]),
        "shelfi_set_sampler_seed": (C.c_int, [vp, u64]),
        "shelfi_clear_sampler_seed": (C.c_int, [vp]),
        "shelfi_chacha20_block_device": (C.c_int, [vp, vp, u32, vp, vp]),
        "shelfi_set_option": (C.c_int, [vp, C.c_int, i64]),
        "shelfi_get_option": (i64, [vp, C.c_int]),
        "shelfi_ring_dim": (u32, [vp]),
        "shelfi_num_limbs": (u32, [vp]),
        "shelfi_slots": (u32, [vp]),
        "shelfi_batch_size": (u32, [vp]),
        "shelfi_modulus": (u64, [vp, u32]),
        "shelfi_root_of_unity": (u64, [vp, u32]),
        "shelfi_scaling_factor": (f64, [vp]),
        "shelfi_key_tag": (C.c_int, [vp, C.c_char_p]),
        "shelfi_num_ciphertexts": (u64, [vp, u64]),
        "shelfi_ciphertext_words": (u64, [vp]),
        "shelfi_blob_size": (u64, [vp, u64]),
        "shelfi_weight_constant": (i64, [vp, C.c_float]),
        "shelfi_encrypt": (C.c_int, [vp, vp, u64, C.POINTER(vp), u64p]),
        "shelfi_encrypt_into": (C.c_int, [vp, vp, u64, vp, u64, u64p]),
        "shelfi_compute_weighted_average": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float), u32,
                                                       C.POINTER(vp), u64p]),
        "shelfi_compute_weighted_average_into": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float), u32, vp, u64,
                                                            u64p]),
        "shelfi_compute_weighted_average_rescaled": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float), u32,
                                                                C.POINTER(vp), u64p]),
        "shelfi_compute_weighted_average_rescaled_into": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float), u32,
                                                                     vp, u64, u64p]),
        "shelfi_rescale": (C.c_int, [vp, vp, u64, C.POINTER(vp), u64p]),
        "shelfi_rescale_into": (C.c_int, [vp, vp, u64, vp, u64, u64p]),
        "shelfi_blob_size_level": (u64, [vp, u64, u32]),
        "shelfi_rescale_device": (C.c_int, [vp, vp, u64, u32, vp, vp]),
        "shelfi_decrypt_level_device": (C.c_int, [vp, vp, u64, u32, f64, vp, u64, vp]),
        "shelfi_decrypt_words_level_device": (C.c_int, [vp, vp, u64, u32, vp, vp]),
        "shelfi_decrypt": (C.c_int, [vp, vp, u64, u64, vp]),
        "shelfi_encrypt_layers": (C.c_int, [vp, C.POINTER(vp), u64p, u32, C.POINTER(vp), u64p, u64p, u64p]),
        "shelfi_compute_weighted_average_layers": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float), u32, u32,
                                                             C.POINTER(vp), u64p, u64p, u64p]),
        "shelfi_decrypt_layers": (C.c_int, [vp, C.POINTER(vp), u64p, u64p, u32, vp]),
        "shelfi_encrypt_layers_into": (C.c_int, [vp, C.POINTER(vp), u64p, u32, vp, u64, u64p, u64p, u64p]),
        "shelfi_compute_weighted_average_layers_into": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float), u32, u32,
                                                                  vp, u64, u64p, u64p, u64p]),
        "shelfi_compute_weighted_average_layers_rescaled": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float), u32,
                                                                      u32, C.POINTER(vp), u64p, u64p, u64p]),
        "shelfi_compute_weighted_average_layers_rescaled_into": (C.c_int, [vp, C.POINTER(vp), u64p, C.POINTER(C.c_float),
                                                                           u32, u32, vp, u64, u64p, u64p, u64p]),
        "shelfi_ct_encrypt": (C.c_int, [vp, vp, u64, C.c_int, C.POINTER(vp)]),
        "shelfi_ct_from_blob": (C.c_int, [vp, vp, u64, C.POINTER(vp)]),
        "shelfi_ct_weighted_average": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_float), u32, C.c_int, C.POINTER(vp)]),
        "shelfi_ct_decrypt": (C.c_int, [vp, vp, u64, vp]),
        "shelfi_ct_blob_size": (u64, [vp]),
        "shelfi_ct_to_blob": (C.c_int, [vp, vp, vp, u64, u64p]),
        "shelfi_ct_info": (C.c_int, [vp, u64p, C.POINTER(u32), u64p, u64p, C.POINTER(f64)]),
        "shelfi_ct_device_ptr": (vp, [vp]),
        "shelfi_ct_free": (None, [vp, vp]),
        "shelfi_resident_stats": (C.c_int, [vp, u64p, u64p, u64p]),
        "shelfi_resident_set_limit": (C.c_int, [vp, u64]),
        "shelfi_synchronize": (C.c_int, [vp]),
        "shelfi_free": (None, [vp]),
        "shelfi_host_alloc": (vp, [C.c_size_t]),
        "shelfi_host_free": (None, [vp]),
        "shelfi_encrypt_device": (C.c_int, [vp, vp, u64, u64, vp, vp]),
        "shelfi_encrypt_many_device": (C.c_int, [vp, C.POINTER(vp), u32, u64, u64, C.POINTER(vp), vp]),
        "shelfi_weighted_sum_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_float), u32, u64, vp, vp]),
        "shelfi_weighted_sum_accumulate_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_float), u32, u64, vp, vp]),
        "shelfi_decrypt_device": (C.c_int, [vp, vp, u64, f64, vp, u64, vp]),
        "shelfi_decrypt_static_device": (C.c_int, [vp, vp, u64, f64, vp, u64, vp]),
        "shelfi_decrypt_gather_device": (C.c_int, [vp, vp, u64, f64, C.POINTER(vp), u32, u64, u64, vp]),
        "shelfi_peer_buffer_create": (C.c_int, [vp, u64, C.POINTER(vp), C.c_char_p]),
        "shelfi_peer_buffer_open": (C.c_int, [vp, C.c_char_p, C.POINTER(vp)]),
        "shelfi_peer_buffer_close": (C.c_int, [vp, vp]),
        "shelfi_peer_buffer_destroy": (C.c_int, [vp, vp]),
        "shelfi_peer_barrier_device": (C.c_int, [vp, C.POINTER(vp), u32, u32, u64, vp]),
        "shelfi_peer_barrier_guard_device": (C.c_int, [vp, C.POINTER(vp), u32, u32, u64, vp, u64, vp]),
        "shelfi_weighted_sum_vector_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(vp), u32, u64, vp, vp]),
        "shelfi_encrypt_selected_many_device": (C.c_int, [vp, C.POINTER(vp), u32, vp, u64, u64, C.POINTER(vp), vp]),
        "shelfi_decrypt_scatter_device": (C.c_int, [vp, vp, u64, f64, vp, vp, u64, vp]),
        "shelfi_plain_weighted_sum_device": (C.c_int, [vp, C.POINTER(vp), C.POINTER(C.c_float), u32, u64, vp, vp]),
        "shelfi_pack_device": (C.c_int, [vp, vp, u64, vp, vp]),
        "shelfi_unpack_device": (C.c_int, [vp, vp, u64, vp, vp]),
        "shelfi_ntt_forward_device": (C.c_int, [vp, vp, u32, vp]),
        "shelfi_ntt_inverse_device": (C.c_int, [vp, vp, u32, vp]),
        "shelfi_ntt_forward_shfl_device": (C.c_int, [vp, vp, u32, vp]),
        "shelfi_encode_device": (C.c_int, [vp, vp, u64, vp, vp]),
        "shelfi_decrypt_integer_device": (C.c_int, [vp, vp, u64, vp, vp, vp]),
        "shelfi_decrypt_words_device": (C.c_int, [vp, vp, u64, vp, vp]),
        "shelfi_profile_enable": (C.c_int, [vp, C.c_int]),
        "shelfi_profile_reset": (C.c_int, [vp]),
        "shelfi_profile_get": (C.c_int, [vp, C.c_int, C.POINTER(C.c_double), u64p]),
        "shelfi_selftest_parallel_copy": (C.c_int, [u64, u32, u32]),
        "shelfi_version": (C.c_char_p, []),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _as_bytes(blob):
    # pybind11 2.6.1's permissive std::string caster takes bytes or str (ckks.cpp:170,276)
    if isinstance(blob, (PinnedBlob, ResidentBlob)):
        return blob
    if isinstance(blob, (bytes, bytearray, memoryview)):
        return bytes(blob) if not isinstance(blob, bytes) else blob
    if isinstance(blob, str):
        return blob.encode("latin-1")
    raise TypeError("ciphertext blob must be bytes or str")


def _new_bytes(n: int) -> bytes:
    """An uninitialised bytes object of n bytes for the library to fill in place (what pybind's py::bytes(ptr, n)
    does in one step): saves the malloc + copy + free of going through a library-owned buffer."""
    f = C.pythonapi.PyBytes_FromStringAndSize
    f.restype, f.argtypes = C.py_object, [C.c_char_p, C.c_ssize_t]
    b = f(None, n)
    if n >= (8 << 20) and _HUGE:
        # a fresh 100 MB buffer is ~26,000 page faults before a byte lands in it; ask for 2 MB pages (no-op when the
        # kernel has transparent huge pages off)
        a = _addr(b)
        lo, hi = (a + _HP - 1) & ~(_HP - 1), (a + n) & ~(_HP - 1)
        if hi > lo:
            _madvise(C.c_void_p(lo), C.c_size_t(hi - lo), 14)      # MADV_HUGEPAGE
    return b


_HP = 2 << 20
_HUGE = os.environ.get("SHELFI_HUGEPAGES", "1") != "0"
try:
    _madvise = C.CDLL(None, use_errno=True).madvise
    _madvise.restype, _madvise.argtypes = C.c_int, [C.c_void_p, C.c_size_t, C.c_int]
except Exception:   # pragma: no cover
    _HUGE = False


def _addr(buf) -> int:
    """Address of a bytes / PinnedBlob / buffer-protocol object's memory (borrowed for the duration of the call)."""
    if isinstance(buf, PinnedBlob):
        return buf.address
    if isinstance(buf, ResidentBlob):
        buf = bytes(buf)     # cached on the blob, which the caller keeps alive for the call
    if isinstance(buf, bytes):
        return C.cast(C.c_char_p(buf), C.c_void_p).value
    return np.frombuffer(buf, dtype=np.uint8).ctypes.data


class _PinnedPool:
    """Page-locked host buffers are expensive to create (cudaHostAlloc pins at a few GB/s) and cheap to reuse:
    freed blobs go back to a size-keyed free list, up to SHELFI_PINNED_POOL_MB (default 2048) of parked memory;
    beyond that they are returned to the driver."""

    def __init__(self):
        self.free = {}
        self.parked = 0
        self.cap_bytes = int(os.environ.get("SHELFI_PINNED_POOL_MB", "2048")) << 20

    def take(self, n: int):
        cap = max(4096, 1 << (int(n) - 1).bit_length()) if n > (1 << 20) else max(4096, int(n))
        lst = self.free.get(cap)
        if lst:
            self.parked -= cap
            return lst.pop(), cap
        p = lib().shelfi_host_alloc(cap)
        if not p:
            self.trim()
            p = lib().shelfi_host_alloc(cap)
            if not p:
                raise MemoryError(f"cudaHostAlloc({cap}) failed")
        return p, cap

    def give(self, p, cap):
        if self.parked + cap > self.cap_bytes:
            lib().shelfi_host_free(p)
            return
        self.parked += cap
        self.free.setdefault(cap, []).append(p)

    def trim(self):
        for lst in self.free.values():
            for p in lst:
                lib().shelfi_host_free(p)
        self.free.clear()
        self.parked = 0


_pool = _PinnedPool()


class _KeepArray(np.ndarray):
    """A numpy view of memory owned by another object (`_keep`): views and slices keep this array as their base, so
    the owner — and with it the page-locked buffer — stays alive as long as anything refers to the memory."""

    _keep = None

    def __array_finalize__(self, obj):
        if obj is not None and getattr(self, "_keep", None) is None:
            self._keep = getattr(obj, "_keep", None)

    def __reduce__(self):          # pickle / deepcopy hand back a plain array
        return np.asarray(self).copy().__reduce__()


class _PinnedBuffer:
    """One page-locked allocation from the pool; returns to it when the last array / blob over it is gone."""

    __slots__ = ("p", "cap")

    def __init__(self, n: int):
        self.p, self.cap = _pool.take(n)

    def __del__(self):
        if getattr(self, "p", None) and _lib is not None:
            _pool.give(self.p, self.cap)
            self.p = None

    def array(self, offset: int, n: int, dtype=np.uint8) -> np.ndarray:
        count = int(n) // np.dtype(dtype).itemsize
        a = np.ctypeslib.as_array((C.c_uint8 * int(n)).from_address(self.p + offset)).view(dtype)[:count].view(_KeepArray)
        a._keep = self
        return a


def pinned_empty(shape, dtype=np.float64) -> np.ndarray:
    """An uninitialised numpy array in page-locked host memory (pooled): inputs placed here are read by the GPU's copy
    engine straight from the caller's pages, at the PCIe rate, instead of through the pageable staging path."""
    shape = (int(shape),) if np.isscalar(shape) else tuple(int(v) for v in shape)
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    return _PinnedBuffer(max(n, 1)).array(0, n, dtype).reshape(shape)


def pinned_array(a, dtype=None) -> np.ndarray:
    """A page-locked copy of `a` (see pinned_empty)."""
    a = np.asarray(a, dtype=dtype)
    out = pinned_empty(a.shape, a.dtype)
    np.copyto(out, a)
    return out


class PinnedBlob:
    """A serialized ciphertext vector held in page-locked host memory (opt-in: ``CKKS(..., pinned=True)``).

    Bytes-like: ``len``, ``bytes(b)``, ``memoryview(b)`` (buffer protocol), ``b.array`` (numpy uint8 view that keeps
    the buffer alive), equality and hash like the ``bytes`` it stands for.  Passing it back to
    ``computeWeightedAverage`` / ``decrypt`` moves it over PCIe at the link rate instead of through the driver's
    pageable staging path; ``bytes(b)`` gives the reference's exact type when the payload has to leave the process."""

    def __init__(self, n: int, _parent=None, _offset: int = 0):
        if _parent is None:
            self._buf, self._off = _PinnedBuffer(n), 0
        else:   # a slice of another blob's buffer (the per-layer results of the *_layers calls)
            self._buf, self._off = _parent._buf, _parent._off + _offset
        self._n = int(n)

    def view(self, offset: int, n: int) -> "PinnedBlob":
        return PinnedBlob(n, self, offset)

    @property
    def array(self) -> np.ndarray:
        return self._buf.array(self._off, self._n)

    @property
    def address(self) -> int:
        return self._buf.p + self._off

    def __len__(self):
        return self._n

    def __bytes__(self):
        return C.string_at(self.address, self._n)

    tobytes = __bytes__

    def __reduce__(self):          # pickle / copy.deepcopy (benchmark_selection.py:171) hand back plain bytes
        return (bytes, (bytes(self),))

    def __buffer__(self, flags):   # PEP 688: memoryview(b), sock.sendall(b), f.write(b) read the pinned pages directly
        return memoryview(self.array)

    def __eq__(self, other):
        if isinstance(other, PinnedBlob):
            return self._n == other._n and bool(np.array_equal(self.array, other.array))
        if isinstance(other, (bytes, bytearray, memoryview, ResidentBlob)):
            return self._n == len(other) and bytes(self) == bytes(other)
        return NotImplemented

    def __hash__(self):            # equal to the bytes it compares equal to
        return hash(bytes(self))


class ResidentBlob:
    """A serialized ciphertext vector that lives in GPU memory (opt-in: ``CKKS(..., resident=True)``).

    What ``encrypt`` and ``computeWeightedAverage`` return in resident mode.  The reference's callers never look inside
    these objects: they collect them in lists and dicts and pass them to the next call (benchmark.py:490-528), which is
    all a handle has to support to cost nothing.  Whenever the bytes themselves are needed — ``bytes(b)``,
    ``memoryview(b)``, pickling, ``==``, writing to a socket or file — the exact wire bytes of the bytes-mode call are
    produced on demand (and cached); ``len(b)`` is known without producing them."""

    __slots__ = ("_ck", "_p", "_len", "_cache", "__weakref__")

    def __init__(self, ckks: "CKKS", ptr):
        self._ck, self._p, self._cache = ckks, ptr, None
        self._len = int(lib().shelfi_ct_blob_size(ptr))

    def __del__(self):
        p, ck = getattr(self, "_p", None), getattr(self, "_ck", None)
        if p and ck is not None and _lib is not None and getattr(ck, "_h", None):
            _lib.shelfi_ct_free(ck._h, p)
        self._p = None

    def __len__(self):
        return self._len

    def __bytes__(self):
        if self._cache is None:
            dst = _new_bytes(self._len)
            n = C.c_uint64()
            self._ck._check(lib().shelfi_ct_to_blob(self._ck._h, self._p, _addr(dst), self._len, C.byref(n)))
            assert n.value == self._len
            self._cache = dst
        return self._cache

    tobytes = __bytes__

    def to_pinned(self) -> PinnedBlob:
        dst = PinnedBlob(self._len)
        n = C.c_uint64()
        self._ck._check(lib().shelfi_ct_to_blob(self._ck._h, self._p, dst.address, self._len, C.byref(n)))
        return dst

    def __reduce__(self):
        return (bytes, (bytes(self),))

    def __buffer__(self, flags):
        return memoryview(bytes(self))

    def __eq__(self, other):
        if isinstance(other, (bytes, bytearray, memoryview, PinnedBlob, ResidentBlob)):
            return self._len == len(other) and bytes(self) == bytes(other)
        return NotImplemented

    def __hash__(self):
        return hash(bytes(self))

    def _info(self):
        n_ct, limbs, depth, level, sf = C.c_uint64(), C.c_uint32(), C.c_uint64(), C.c_uint64(), C.c_double()
        lib().shelfi_ct_info(self._p, C.byref(n_ct), C.byref(limbs), C.byref(depth), C.byref(level), C.byref(sf))
        return n_ct.value, limbs.value, depth.value, level.value, sf.value

    n_ct = property(lambda self: self._info()[0])
    limbs = property(lambda self: self._info()[1])
    depth = property(lambda self: self._info()[2])
    level = property(lambda self: self._info()[3])
    scaling_factor = property(lambda self: self._info()[4])
    device_ptr = property(lambda self: lib().shelfi_ct_device_ptr(self._p))


_PinnedArray = _KeepArray   # decrypt results in pinned / resident mode


class Scheme:
    """Opaque base class, as bound at binding.cpp:16 (include/scheme.h:15-32)."""

    def __init__(self, scheme: str = "ckks"):
        self.scheme = scheme


class CKKS(Scheme):
    """CKKS helper with the reference's constructor and methods (binding.cpp:18-31)."""

    def __init__(self, scheme: str = "ckks", batchSize: int = 4096, scaleFactorBits: int = 52,
                 cryptodir: str = "../resources/cryptoparams/", device: int | None = None, pinned: bool = False,
                 rescale: bool = False, resident: bool = False):
        super().__init__(scheme)
        # resident mode (extension, off by default): encrypt / computeWeightedAverage return ResidentBlob objects whose
        # payload stays in GPU memory; decrypt and computeWeightedAverage accept them (and bytes, and PinnedBlob) in any
        # mix.  The wire bytes are produced only when something asks for them.
        self.resident = bool(resident) or os.environ.get("SHELFI_RESIDENT", "0") == "1"
        # rescale-before-transport (extension, off by default like the reference): computeWeightedAverage returns the
        # average one tower shorter (PALISADE ModReduce), i.e. half the bytes at the shipped two-tower context
        self.rescale = bool(rescale) or os.environ.get("SHELFI_RESCALE", "0") == "1"
        self.pinned = bool(pinned) or os.environ.get("SHELFI_PINNED", "0") == "1"   # return PinnedBlob objects instead of bytes (extension; see PinnedBlob)
        self.batchSize, self.scaleFactorBits, self.cryptodir = int(batchSize), int(scaleFactorBits), str(cryptodir)
        if device is None:
            device = int(os.environ.get("SHELFI_DEVICE", os.environ.get("LOCAL_RANK", "0")))
        self.device = device
        h = C.c_void_p()
        rc = lib().shelfi_ckks_create(scheme.encode(), self.batchSize, self.scaleFactorBits, self.cryptodir.encode(),
                                      device, C.byref(h))
        if rc:
            raise ShelfiError(rc, "could not create CKKS handle")
        self._h = h
        self.load_status = None

    def __del__(self):
        h = getattr(self, "_h", None)
        if h and _lib is not None:
            _lib.shelfi_ckks_destroy(h)
            self._h = None

    # ---- helpers ----
    def _err(self) -> str:
        return lib().shelfi_last_error(self._h).decode(errors="replace")

    def _check(self, rc: int):
        if rc < 0:
            raise ShelfiError(rc, self._err())

    # ---- reference API ----
    def loadCryptoParams(self) -> None:
        """ckks.cpp:11-23: read cryptocontext.txt / key-public.txt / key-private.txt from cryptodir.

        Like the reference, a missing file is reported on stderr-ish output and does not raise; the
        status mask (1 context, 2 public key, 4 private key) is kept in ``load_status``.  A missing CUDA
        device does raise: nothing can run without it.
        """
        rc = lib().shelfi_load_crypto_params(self._h)
        self._check(rc)
        self.load_status = rc
        if rc & 1:  # same words as ckks.cpp:13,17,21
            print("Could not read serialization from " + self.cryptodir + "cryptocontext.txt")
        if rc & 2:
            print("Could not read public key")
        if rc & 4:
            print("Could not read secret key", file=sys.stderr)

    def genCryptoContextAndKeyGen(self, mult_depth: int = 1, ring_dim: int = 0) -> int:
        """ckks.cpp:25-59: write a fresh context and key pair into cryptodir; returns 1 or 0; does not load.

        The reference takes no arguments (multDepth = 1 is hard-coded at ckks.cpp:26, the ring dimension is
        PALISADE's choice); ``mult_depth`` / ``ring_dim`` are extensions for deeper modulus chains and larger
        rings (BASELINE configs 3-4: (3, 0) -> N = 16384, 4 limbs; (7, 0) -> N = 32768, 8 limbs at 52 bits).
        """
        ok = lib().shelfi_gen_crypto_context_and_keygen_ex(self._h, int(mult_depth), int(ring_dim))
        if not ok:
            print("Error generating crypto context and keys: " + self._err())
        return int(ok)

    def encrypt(self, data_array) -> bytes:
        """ckks.cpp:61-104: any array-like convertible to float64 (flattened) -> serialized ciphertext vector."""
        if self.resident:
            return self._encrypt_resident(data_array)
        x = np.ascontiguousarray(data_array, dtype=np.float64).ravel()
        cap = self.blob_size(self.num_ciphertexts(x.size)) if x.size else 0
        if cap == 0:   # not loaded / empty input: let the library report it
            out, n = C.c_void_p(), C.c_uint64()
            self._check(lib().shelfi_encrypt(self._h, x.ctypes.data, x.size, C.byref(out), C.byref(n)))
            raise ShelfiError(-1, "encrypt produced no ciphertext")
        dst = PinnedBlob(cap) if self.pinned else _new_bytes(cap)
        n = C.c_uint64()
        self._check(lib().shelfi_encrypt_into(self._h, x.ctypes.data, x.size, _addr(dst), cap, C.byref(n)))
        assert n.value == cap
        return dst

    def _encrypt_resident(self, data_array) -> ResidentBlob:
        """encrypt with the ciphertexts left in GPU memory.  float32 arrays (what the reference's callers pass: the
        state_dict entries as numpy, benchmark.py:490-493) are uploaded as they are, 4 bytes per parameter, and widened
        on the device — the values pybind's array_t<double> conversion would produce; everything else goes as float64.
        Objects with ``__cuda_array_interface__`` (a torch.cuda tensor, a cupy array; float32 or float64, contiguous,
        already computed: synchronise their stream first) are read where they are."""
        cai = getattr(data_array, "__cuda_array_interface__", None)
        out = C.c_void_p()
        if cai is not None:
            if cai.get("strides") is not None or cai["typestr"] not in ("<f8", "<f4"):
                raise TypeError("device input must be contiguous float64 or float32")
            n = int(np.prod(cai["shape"])) if cai["shape"] else 1
            self._check(lib().shelfi_ct_encrypt(self._h, cai["data"][0], n, int(cai["typestr"] == "<f4"), C.byref(out)))
            return ResidentBlob(self, out)
        a = data_array if isinstance(data_array, np.ndarray) else np.asarray(data_array)
        f32 = a.dtype == np.float32
        x = np.ascontiguousarray(a, dtype=np.float32 if f32 else np.float64).ravel()
        self._check(lib().shelfi_ct_encrypt(self._h, x.ctypes.data, x.size, int(f32), C.byref(out)))
        return ResidentBlob(self, out)

    def _as_resident(self, blob):
        """(handle pointer, owner to keep alive) for any accepted blob type; bytes are uploaded and unpacked."""
        if isinstance(blob, ResidentBlob) and blob._ck is self:
            return blob._p, blob
        raw = _as_bytes(blob)
        out = C.c_void_p()
        self._check(lib().shelfi_ct_from_blob(self._h, _addr(raw), len(raw), C.byref(out)))
        return out, ResidentBlob(self, out)

    def to_resident(self, blob) -> ResidentBlob:
        """bytes / PinnedBlob (from this engine or a PALISADE peer) -> ResidentBlob on this handle's GPU."""
        return self._as_resident(blob)[1]

    def _wavg_resident(self, learner_data, scaling_factors, rescale: bool):
        k = len(learner_data)
        if k == 0:
            raise ShelfiError(-1, "no learners")
        held = [self._as_resident(b) for b in learner_data]
        ptrs = (C.c_void_p * k)(*[p if isinstance(p, int) else p.value for p, _ in held])
        w = (C.c_float * k)(*[float(v) for v in scaling_factors])
        out = C.c_void_p()
        self._check(lib().shelfi_ct_weighted_average(self._h, ptrs, w, k, int(rescale), C.byref(out)))
        return ResidentBlob(self, out)

    def computeWeightedAverage(self, learner_data, scaling_factors, rescale: bool | None = None) -> bytes:
        """ckks.cpp:264-320: sum_i float32(w_i) * ct_i over learners; size mismatch prints and returns b"".

        ``rescale`` (extension; default: the handle's setting, normally off) appends PALISADE's ModReduce to every
        ciphertext of the result."""
        rescale = self.rescale if rescale is None else bool(rescale)
        learner_data, scaling_factors = list(learner_data), list(scaling_factors)
        if len(learner_data) != len(scaling_factors):
            print("Error: learner_data and scaling_factors size mismatch")  # ckks.cpp:266
            return b""
        if self.resident or any(isinstance(b, ResidentBlob) for b in learner_data):
            res = self._wavg_resident(learner_data, scaling_factors, rescale)
            return res if self.resident else (res.to_pinned() if self.pinned else bytes(res))
        blobs = [_as_bytes(b) for b in learner_data]
        k = len(blobs)
        ptrs = (C.c_void_p * k)(*[_addr(b) for b in blobs])
        lens = (C.c_uint64 * k)(*[len(b) for b in blobs])
        w = (C.c_float * k)(*[float(v) for v in scaling_factors])
        cap = len(blobs[0]) if k else 0   # the result has learner 0's layout (ckks.cpp:291-292)
        into = lib().shelfi_compute_weighted_average_into
        alloc = lib().shelfi_compute_weighted_average
        if rescale:
            into, alloc = lib().shelfi_compute_weighted_average_rescaled_into, lib().shelfi_compute_weighted_average_rescaled
            n_ct = (cap - self.blob_size(1)) // (self.blob_size(2) - self.blob_size(1)) + 1 if cap else 0
            cap = lib().shelfi_blob_size_level(self._h, n_ct, self.num_limbs - 1) if n_ct > 0 and self.num_limbs > 1 else 0
        if cap:
            dst = PinnedBlob(cap) if self.pinned else _new_bytes(cap)
            n = C.c_uint64()
            rc = into(self._h, ptrs, lens, w, k, _addr(dst), cap, C.byref(n))
            if rc == 0 and n.value == cap:
                return dst
        out, n = C.c_void_p(), C.c_uint64()   # errors, or a result of another size: the allocating call reports / returns it
        self._check(alloc(self._h, ptrs, lens, w, k, C.byref(out), C.byref(n)))
        try:
            return C.string_at(out.value, n.value)
        finally:
            lib().shelfi_free(out)

    def rescaleCiphertexts(self, learner_data) -> bytes:
        """Extension: PALISADE ModReduce on every ciphertext of an aggregated (depth-2) blob — one tower fewer."""
        blob = _as_bytes(learner_data)
        out, n = C.c_void_p(), C.c_uint64()
        self._check(lib().shelfi_rescale(self._h, _addr(blob), len(blob), C.byref(out), C.byref(n)))
        try:
            return C.string_at(out.value, n.value)
        finally:
            lib().shelfi_free(out)

    def decrypt(self, learner_data, data_dimensions: int) -> np.ndarray:
        """ckks.cpp:170-213: serialized ciphertext vector -> float64[data_dimensions] (level-0 or rescaled)."""
        blob = _as_bytes(learner_data)
        n = int(data_dimensions)
        if (self.pinned or self.resident) and n * 8 >= (1 << 20):
            # pinned / resident mode: the result lives in pooled page-locked memory (an ndarray that keeps its buffer
            # alive), so the values cross PCIe at the link rate and the engine's pipelined decrypt applies
            out = pinned_empty(n, np.float64)
        else:
            out = np.empty(n, dtype=np.float64)
        if isinstance(blob, ResidentBlob) and blob._ck is self:
            self._check(lib().shelfi_ct_decrypt(self._h, blob._p, n, out.ctypes.data))
        else:
            self._check(lib().shelfi_decrypt(self._h, _addr(blob), len(blob), n, out.ctypes.data))
        return out

    # ---- layer-wise callers (benchmark.py:490-493, benchmark_selection.py:152-193) in one call each ----
    _PIPELINE_MIN_CTS = 128    # engine.cpp kPipelineMinCts: from here a single vector is pipelined against its own copies
    def encrypt_layers(self, arrays) -> list:
        """[encrypt(a) for a in arrays] with one host->device copy, one stream of launches, one copy back."""
        if self.resident or not self.pinned:   # resident: nothing crosses PCIe but the inputs; bytes objects cannot
            return [self.encrypt(a) for a in arrays]   # share one buffer: filling each in place beats slicing copies
        flat = [np.ascontiguousarray(a, dtype=np.float64).ravel() for a in arrays]
        # a layer of >= 128 ciphertexts goes through the single-vector call, whose copy back overlaps its own
        # compute group by group; the small layers of the model share one call (one launch sequence, one copy back)
        big = [i for i, a in enumerate(flat) if self.num_ciphertexts(a.size) >= self._PIPELINE_MIN_CTS]
        if big and len(flat) > 1:
            out = [None] * len(flat)
            for i in big:
                out[i] = self.encrypt(flat[i])
            rest = [i for i in range(len(flat)) if out[i] is None]
            for i, b in zip(rest, self.encrypt_layers([flat[i] for i in rest]) if rest else []):
                out[i] = b
            return out
        n = len(flat)
        xp = (C.c_void_p * n)(*[a.ctypes.data for a in flat])
        ll = (C.c_uint64 * n)(*[a.size for a in flat])
        offs, lens, total = (C.c_uint64 * n)(), (C.c_uint64 * n)(), C.c_uint64()
        cap = sum((self.blob_size(self.num_ciphertexts(a.size)) + 7) // 8 * 8 for a in flat if a.size)
        dst = PinnedBlob(max(cap, 8))
        self._check(lib().shelfi_encrypt_layers_into(self._h, xp, ll, n, _addr(dst), cap, offs, lens,
                                                     C.byref(total)))
        return self._split(dst, offs, lens, n)

    def _split(self, dst, offs, lens, n) -> list:
        return [dst.view(offs[i], lens[i]) for i in range(n)]

    def computeWeightedAverage_layers(self, learner_layers, scaling_factors, rescale: bool | None = None) -> list:
        """learner_layers[i][s] = blob of learner i, layer s -> [computeWeightedAverage over learners for each layer];
        ``rescale`` as in computeWeightedAverage."""
        rescale = self.rescale if rescale is None else bool(rescale)
        learner_layers, scaling_factors = [list(l) for l in learner_layers], list(scaling_factors)
        if len(learner_layers) != len(scaling_factors):
            print("Error: learner_data and scaling_factors size mismatch")  # ckks.cpp:266
            return []
        k, n = len(learner_layers), len(learner_layers[0]) if learner_layers else 0
        if any(len(l) != n for l in learner_layers):
            raise ValueError("every learner must supply the same number of layers")
        if self.resident or not self.pinned or any(isinstance(b, ResidentBlob) for l in learner_layers for b in l):
            return [self.computeWeightedAverage([l[s] for l in learner_layers], scaling_factors, rescale) for s in range(n)]
        big_len = self.blob_size(self._PIPELINE_MIN_CTS)
        big = [s for s in range(n) if k and len(learner_layers[0][s]) >= big_len]
        if big and n > 1:    # as in encrypt_layers: large layers take the pipelined single-vector call
            out = [None] * n
            for s in big:
                out[s] = self.computeWeightedAverage([l[s] for l in learner_layers], scaling_factors, rescale)
            rest = [s for s in range(n) if out[s] is None]
            if rest:
                for s, b in zip(rest, self.computeWeightedAverage_layers([[l[s] for s in rest] for l in learner_layers],
                                                                       scaling_factors, rescale)):
                    out[s] = b
            return out
        blobs = [_as_bytes(b) for l in learner_layers for b in l]
        ptrs = (C.c_void_p * (k * n))(*[_addr(b) for b in blobs])
        blens = (C.c_uint64 * (k * n))(*[len(b) for b in blobs])
        w = (C.c_float * k)(*[float(v) for v in scaling_factors])
        offs, lens, total = (C.c_uint64 * n)(), (C.c_uint64 * n)(), C.c_uint64()
        cap = sum((len(b) + 7) // 8 * 8 for b in blobs[:n])   # results have learner 0's layout (ckks.cpp:291-292)
        dst = PinnedBlob(max(cap, 8))   # a rescaled result is shorter than learner 0's blobs, never longer
        into = lib().shelfi_compute_weighted_average_layers_rescaled_into if rescale else lib().shelfi_compute_weighted_average_layers_into
        self._check(into(self._h, ptrs, blens, w, k, n, _addr(dst), cap, offs, lens, C.byref(total)))
        return self._split(dst, offs, lens, n)

    def decrypt_layers(self, blobs, data_dimensions) -> list:
        """[decrypt(b, n) for b, n in zip(blobs, data_dimensions)] in one pass; returns one float64 array per layer."""
        blobs = [_as_bytes(b) for b in blobs]
        dims = [int(d) for d in data_dimensions]
        n = len(blobs)
        if n != len(dims):
            raise ValueError("one data_dimensions entry per blob")
        if any(isinstance(b, ResidentBlob) for b in blobs):
            return [self.decrypt(b, d) for b, d in zip(blobs, dims)]
        ptrs = (C.c_void_p * n)(*[_addr(b) for b in blobs])
        blens = (C.c_uint64 * n)(*[len(b) for b in blobs])
        dd = (C.c_uint64 * n)(*dims)
        out = np.empty(sum(dims), dtype=np.float64)
        self._check(lib().shelfi_decrypt_layers(self._h, ptrs, blens, dd, n, out.ctypes.data))
        return np.split(out, np.cumsum(dims)[:-1]) if n else []

    encrypt_cpp = encrypt
    decrypt_cpp = decrypt
    computeWeightedAverage_cpp = computeWeightedAverage

    # ---- extras beyond the reference (tests, benchmarks, sharded callers) ----
    def set_sampler_seed(self, seed: int | None) -> None:
        if seed is None:
            lib().shelfi_clear_sampler_seed(self._h)
        else:
            lib().shelfi_set_sampler_seed(self._h, int(seed) & (2**64 - 1))

    OPT_EVALMULT_ROUND, OPT_DECODE_NOISE, OPT_WIRE_VARIANT = 1, 2, 3

    def set_option(self, option: int, value: int) -> None:
        """The PALISADE semantics no fixture pins (include/shelfi_b200.h: SHELFI_OPT_*): EvalMult rounding, decode
        noise, wire variant of generated contexts."""
        self._check(lib().shelfi_set_option(self._h, int(option), int(value)))

    def get_option(self, option: int) -> int:
        return int(lib().shelfi_get_option(self._h, int(option)))

    def synchronize(self) -> None:
        """Wait for everything this handle has queued on the GPU (resident calls return before their kernels end)."""
        self._check(lib().shelfi_synchronize(self._h))

    def resident_stats(self) -> dict:
        """Bytes of GPU memory held by live ResidentBlobs, parked on the pool's free list, and the pool's limit."""
        a, b, c = C.c_uint64(), C.c_uint64(), C.c_uint64()
        self._check(lib().shelfi_resident_stats(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return {"in_use": a.value, "pooled": b.value, "limit": c.value}

    def set_resident_limit(self, nbytes: int) -> None:
        self._check(lib().shelfi_resident_set_limit(self._h, int(nbytes)))

    def load_from_memory(self, cc: bytes | None, pk: bytes | None, sk: bytes | None) -> int:
        rc = lib().shelfi_load_crypto_params_from_memory(self._h, cc, len(cc or b""), pk, len(pk or b""), sk,
                                                         len(sk or b""))
        self._check(rc)
        self.load_status = rc
        return rc

    @property
    def handle(self) -> C.c_void_p:
        return self._h

    @property
    def ring_dim(self) -> int: return lib().shelfi_ring_dim(self._h)
    @property
    def num_limbs(self) -> int: return lib().shelfi_num_limbs(self._h)
    @property
    def slots(self) -> int: return lib().shelfi_slots(self._h)
    @property
    def scaling_factor(self) -> float: return lib().shelfi_scaling_factor(self._h)
    @property
    def ciphertext_words(self) -> int: return lib().shelfi_ciphertext_words(self._h)

    def modulus(self, limb: int) -> int: return lib().shelfi_modulus(self._h, limb)
    def root_of_unity(self, limb: int) -> int: return lib().shelfi_root_of_unity(self._h, limb)
    def num_ciphertexts(self, n: int) -> int: return lib().shelfi_num_ciphertexts(self._h, n)
    def blob_size(self, n_ct: int) -> int: return lib().shelfi_blob_size(self._h, n_ct)
    def weight_constant(self, w: float) -> int: return lib().shelfi_weight_constant(self._h, w)

    def key_tag(self) -> str:
        buf = C.create_string_buffer(33)
        lib().shelfi_key_tag(self._h, buf)
        return buf.value.decode()
