"""CPU oracle for the SHELFI_FHE CKKS path — TEST INFRASTRUCTURE ONLY.

ctypes front end of oracle/shelfi_oracle.c (a plain-C restatement of
/root/reference/palisade_pybind/SHELFI_FHE/src/ckks.cpp and the PALISADE semantics it calls).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this package; the product under fhe-fed_b200/ never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess
from pathlib import Path

import numpy as np

_HERE = Path(__file__).resolve().parent
_LIB_PATH = _HERE / "_build" / "libshelfi_oracle.so"


def _cpu_stamp() -> str:
    """The oracle is compiled -march=native (it is the timed CPU baseline): a library built on another machine
    (it travels with the repo snapshot) may hold instructions this CPU lacks, so it is rebuilt when the flags differ."""
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("flags"):
                import hashlib
                return hashlib.sha256(line.encode()).hexdigest()
    except OSError:
        pass
    return "unknown"


def build(force: bool = False) -> Path:
    """Compile the oracle with gcc (oracle/Makefile)."""
    src = [_HERE / "shelfi_oracle.c", _HERE / "shelfi_oracle.h", _HERE / "Makefile"]
    stamp = _HERE / "_build" / "cpu.stamp"
    stale = (not _LIB_PATH.exists()) or any(s.stat().st_mtime > _LIB_PATH.stat().st_mtime for s in src)
    foreign = not stamp.exists() or stamp.read_text() != _cpu_stamp()
    if force or stale or foreign:
        env = dict(os.environ)
        env.pop("CC", None)
        subprocess.run(["make", "-C", str(_HERE), "-B" if (force or foreign) else "-s"], check=True, env=env,
                       stdout=subprocess.DEVNULL)
        stamp.write_text(_cpu_stamp())
    return _LIB_PATH


class BlobInfo(C.Structure):
    _fields_ = [("n_ct", C.c_uint64), ("n_polys", C.c_uint32), ("n_limbs", C.c_uint32), ("ring_dim", C.c_uint32),
                ("depth", C.c_uint64), ("level", C.c_uint64), ("scaling_factor", C.c_double),
                ("encoding_type", C.c_uint32), ("key_tag", C.c_char * 33), ("first_data_offset", C.c_uint64),
                ("ct_stride", C.c_uint64)]


_lib = None


def lib():
    global _lib
    if _lib is not None:
        return _lib
    L = C.CDLL(str(build()))
    u8p, u64p, f64p, i64p = (C.POINTER(C.c_uint8), C.POINTER(C.c_uint64), C.POINTER(C.c_double),
                             C.POINTER(C.c_int64))
    sig = {
        "orc_create": (C.c_void_p, [C.c_uint32, C.c_uint32, C.c_char_p]),
        "orc_destroy": (None, [C.c_void_p]),
        "orc_load_crypto_params": (C.c_int, [C.c_void_p]),
        "orc_load_from_memory": (C.c_int, [C.c_void_p, C.c_char_p, C.c_size_t, C.c_char_p, C.c_size_t, C.c_char_p,
                                           C.c_size_t]),
        "orc_gen_context_and_keys": (C.c_int, [C.c_uint32, C.c_uint32, C.c_char_p, C.c_uint64, C.c_int]),
        "orc_gen_context_and_keys_ex": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_uint32, C.c_char_p,
                                                  C.c_uint64, C.c_int]),
        "orc_select_ring_dim_ex": (C.c_uint32, [C.c_uint32, C.c_uint32]),
        "orc_num_large_digits": (C.c_uint32, [C.c_uint32]),
        "orc_modulus_chain": (C.c_int, [C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p]),
        "orc_decrypt_ct_words": (C.c_int, [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p]),
        "orc_ct_mul_plain_accumulate": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]),
        "orc_rescale_constants": (None, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
        "orc_rescale_ct": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
        "orc_decrypt_ct_level": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_double, C.c_void_p, C.c_void_p]),
        "orc_rescale_blob": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.POINTER(C.c_void_p), u64p]),
        "orc_last_error": (C.c_char_p, [C.c_void_p]),
        "orc_ring_dim": (C.c_uint32, [C.c_void_p]),
        "orc_num_limbs": (C.c_uint32, [C.c_void_p]),
        "orc_slots": (C.c_uint32, [C.c_void_p]),
        "orc_modulus": (C.c_uint64, [C.c_void_p, C.c_int]),
        "orc_psi": (C.c_uint64, [C.c_void_p, C.c_int]),
        "orc_scaling_factor": (C.c_double, [C.c_void_p]),
        "orc_wire_variant": (C.c_int, [C.c_void_p]),
        "orc_get_public_key": (C.c_int, [C.c_void_p, C.c_void_p, C.c_void_p]),
        "orc_get_secret_key": (C.c_int, [C.c_void_p, C.c_void_p]),
        "orc_get_key_tag": (None, [C.c_void_p, C.c_char_p]),
        "orc_first_prime": (C.c_uint64, [C.c_uint32, C.c_uint32]),
        "orc_previous_prime": (C.c_uint64, [C.c_uint64, C.c_uint32]),
        "orc_min_root_of_unity": (C.c_uint64, [C.c_uint32, C.c_uint64]),
        "orc_select_ring_dim": (C.c_uint32, [C.c_uint32, C.c_uint32]),
        "orc_ntt_forward": (None, [C.c_void_p, C.c_int, C.c_void_p]),
        "orc_ntt_inverse": (None, [C.c_void_p, C.c_int, C.c_void_p]),
        "orc_fft_special": (None, [C.c_void_p, C.c_void_p, C.c_uint32]),
        "orc_fft_special_inv": (None, [C.c_void_p, C.c_void_p, C.c_uint32]),
        "orc_chacha20_block": (None, [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]),
        "orc_seed_key": (None, [C.c_uint64, C.c_uint32, C.c_void_p]),
        "orc_set_option": (C.c_int, [C.c_void_p, C.c_int, C.c_int64]),
        "orc_decode_noise": (C.c_double, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_uint64]),
        "orc_gauss_from_words": (C.c_int64, [C.c_void_p, C.c_uint32, C.c_uint32]),
        "orc_ternary_from_half": (C.c_int64, [C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32]),
        "orc_sample_encrypt_noise": (None, [C.c_void_p, C.c_uint64, C.c_uint64, C.c_void_p, C.c_void_p, C.c_void_p]),
        "orc_encode_coeffs": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p]),
        "orc_encrypt_ct": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint64, C.c_void_p]),
        "orc_decrypt_ct": (C.c_int, [C.c_void_p, C.c_void_p, C.c_double, C.c_void_p, C.c_void_p, C.c_void_p]),
        "orc_weight_constant": (C.c_int64, [C.c_void_p, C.c_float]),
        "orc_num_ciphertexts": (C.c_uint64, [C.c_void_p, C.c_uint64]),
        "orc_blob_size": (C.c_uint64, [C.c_void_p, C.c_uint64]),
        "orc_encrypt": (C.c_int, [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint64, C.POINTER(C.c_void_p), u64p]),
        "orc_weighted_average": (C.c_int, [C.c_void_p, C.POINTER(C.c_char_p), u64p, C.POINTER(C.c_float), C.c_uint32,
                                           C.POINTER(C.c_void_p), u64p]),
        "orc_decrypt": (C.c_int, [C.c_void_p, C.c_char_p, C.c_uint64, C.c_uint64, C.c_void_p]),
        "orc_free": (None, [C.c_void_p]),
        "orc_blob_parse": (C.c_int, [C.c_char_p, C.c_uint64, C.c_int, C.POINTER(BlobInfo), C.c_void_p, C.c_uint64]),
        "orc_blob_reemit": (C.c_int, [C.c_char_p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p), u64p]),
        "orc_file_reemit": (C.c_int, [C.c_char_p, C.c_uint64, C.c_int, C.POINTER(C.c_void_p), u64p]),
        "orc_blob_build": (C.c_int, [C.c_void_p, C.c_uint64, C.c_void_p, C.c_uint64, C.c_uint64, C.c_double,
                                     C.POINTER(C.c_void_p), u64p]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype, fn.argtypes = res, args
    _lib = L
    return L


def _take(ptr: C.c_void_p, n: int) -> bytes:
    try:
        return C.string_at(ptr.value, n)
    finally:
        lib().orc_free(ptr)


def _u64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.uint64)


class Oracle:
    """Mirror of the reference's CKKS class (binding.cpp:18-31) running on the CPU oracle."""

    def __init__(self, scheme: str = "ckks", batchSize: int = 4096, scaleFactorBits: int = 52,
                 cryptodir: str = "../resources/cryptoparams/"):
        self.batchSize, self.scaleFactorBits, self.cryptodir = batchSize, scaleFactorBits, cryptodir
        self._h = lib().orc_create(batchSize, scaleFactorBits, cryptodir.encode())

    def __del__(self):
        if getattr(self, "_h", None):
            lib().orc_destroy(self._h)
            self._h = None

    # --- reference API ---
    def loadCryptoParams(self) -> int:
        return lib().orc_load_crypto_params(self._h)

    def genCryptoContextAndKeyGen(self, seed: int = 1, nine_int: bool = False, mult_depth: int = 1,
                                  ring_dim: int = 0) -> int:
        """mult_depth / ring_dim are extensions (ckks.cpp:26 hard-codes multDepth = 1, ring_dim automatic)."""
        return lib().orc_gen_context_and_keys_ex(self.batchSize, self.scaleFactorBits, mult_depth, ring_dim,
                                                 self.cryptodir.encode(), seed, int(nine_int))

    def encrypt(self, data, seed: int = 0) -> bytes:
        x = np.ascontiguousarray(data, dtype=np.float64).ravel()
        out, n = C.c_void_p(), C.c_uint64()
        rc = lib().orc_encrypt(self._h, x.ctypes.data, x.size, seed, C.byref(out), C.byref(n))
        if rc:
            raise RuntimeError(f"oracle encrypt failed ({rc}): {self.last_error()}")
        return _take(out, n.value)

    def computeWeightedAverage(self, blobs, weights) -> bytes:
        if len(blobs) != len(weights):  # ckks.cpp:265-268
            print("Error: learner_data and scaling_factors size mismatch")
            return b""
        k = len(blobs)
        arr = (C.c_char_p * k)(*blobs)
        lens = (C.c_uint64 * k)(*[len(b) for b in blobs])
        w = (C.c_float * k)(*[float(v) for v in weights])
        out, n = C.c_void_p(), C.c_uint64()
        rc = lib().orc_weighted_average(self._h, arr, lens, w, k, C.byref(out), C.byref(n))
        if rc:
            raise RuntimeError(f"oracle weighted average failed ({rc}): {self.last_error()}")
        return _take(out, n.value)

    def decrypt(self, blob: bytes, n: int) -> np.ndarray:
        out = np.empty(n, dtype=np.float64)
        rc = lib().orc_decrypt(self._h, blob, len(blob), n, out.ctypes.data)
        if rc:
            raise RuntimeError(f"oracle decrypt failed ({rc}): {self.last_error()}")
        return out

    # --- introspection / primitives for tests ---
    def last_error(self) -> str:
        return lib().orc_last_error(self._h).decode()

    @property
    def ring_dim(self): return lib().orc_ring_dim(self._h)
    @property
    def num_limbs(self): return lib().orc_num_limbs(self._h)
    @property
    def slots(self): return lib().orc_slots(self._h)
    @property
    def scaling_factor(self): return lib().orc_scaling_factor(self._h)
    @property
    def wire_variant(self): return lib().orc_wire_variant(self._h)
    def modulus(self, l): return lib().orc_modulus(self._h, l)
    def psi(self, l): return lib().orc_psi(self._h, l)

    def key_tag(self) -> str:
        buf = C.create_string_buffer(33)
        lib().orc_get_key_tag(self._h, buf)
        return buf.value.decode()

    def public_key(self):
        shape = (self.num_limbs, self.ring_dim)
        b, a = np.empty(shape, np.uint64), np.empty(shape, np.uint64)
        if not lib().orc_get_public_key(self._h, b.ctypes.data, a.ctypes.data):
            raise RuntimeError("public key not loaded")
        return b, a

    def secret_key(self):
        s = np.empty((self.num_limbs, self.ring_dim), np.uint64)
        if not lib().orc_get_secret_key(self._h, s.ctypes.data):
            raise RuntimeError("secret key not loaded")
        return s

    def ntt_forward(self, limb, a):
        a = _u64(a).copy()
        lib().orc_ntt_forward(self._h, limb, a.ctypes.data)
        return a

    def ntt_inverse(self, limb, a):
        a = _u64(a).copy()
        lib().orc_ntt_inverse(self._h, limb, a.ctypes.data)
        return a

    def fft_special(self, z):
        v = np.ascontiguousarray(z, dtype=np.complex128).copy()
        lib().orc_fft_special(self._h, v.ctypes.data, v.size)
        return v

    def fft_special_inv(self, z):
        v = np.ascontiguousarray(z, dtype=np.complex128).copy()
        lib().orc_fft_special_inv(self._h, v.ctypes.data, v.size)
        return v

    OPT_EVALMULT_ROUND, OPT_DECODE_NOISE, OPT_NOISE_SEED = 1, 2, 4

    def set_option(self, option: int, value: int) -> None:
        if lib().orc_set_option(self._h, option, value):
            raise ValueError(f"bad option {option} = {value}")

    def decode_noise(self, z, key, ct_index: int):
        """PALISADE-style decode noise on a complex vector at scale 2^p: (noisy copy, standard deviation added)."""
        v = np.ascontiguousarray(z, dtype=np.complex128).copy()
        k = np.ascontiguousarray(key, dtype=np.uint32)
        sd = lib().orc_decode_noise(self._h, v.ctypes.data, v.size, k.ctypes.data, ct_index)
        return v, sd

    def gauss_from_words(self, h: int, l: int) -> int:
        return lib().orc_gauss_from_words(self._h, h, l)

    def sample_noise(self, seed, ct_index):
        n = self.ring_dim
        v, e0, e1 = (np.empty(n, np.int64) for _ in range(3))
        lib().orc_sample_encrypt_noise(self._h, seed, ct_index, v.ctypes.data, e0.ctypes.data, e1.ctypes.data)
        return v, e0, e1

    def encode_coeffs(self, x):
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty(self.ring_dim, np.int64)
        rc = lib().orc_encode_coeffs(self._h, x.ctypes.data, x.size, out.ctypes.data)
        if rc:
            raise RuntimeError("encode failed")
        return out

    def encrypt_ct(self, x, seed, ct_index):
        x = np.ascontiguousarray(x, dtype=np.float64)
        out = np.empty((2, self.num_limbs, self.ring_dim), np.uint64)
        rc = lib().orc_encrypt_ct(self._h, x.ctypes.data, x.size, seed, ct_index, out.ctypes.data)
        if rc:
            raise RuntimeError("encrypt_ct failed")
        return out

    def decrypt_ct(self, ct, sf, want_coeffs=False):
        ct = _u64(ct)
        out = np.empty(self.slots, np.float64)
        hi = np.empty(self.ring_dim, np.int64) if want_coeffs else None
        lo = np.empty(self.ring_dim, np.uint64) if want_coeffs else None
        rc = lib().orc_decrypt_ct(self._h, ct.ctypes.data, sf, out.ctypes.data,
                                  hi.ctypes.data if want_coeffs else None, lo.ctypes.data if want_coeffs else None)
        if rc:
            raise RuntimeError(f"decrypt_ct failed ({rc})")
        return (out, hi, lo) if want_coeffs else out

    def decrypt_ct_words(self, ct, sf):
        """(doubles, centred CRT integers as Python ints) for any limb count."""
        ct = _u64(ct)
        out = np.empty(self.slots, np.float64)
        words = np.empty((self.ring_dim, 8), np.uint64)
        rc = lib().orc_decrypt_ct_words(self._h, ct.ctypes.data, sf, out.ctypes.data, words.ctypes.data)
        if rc:
            raise RuntimeError(f"decrypt_ct_words failed ({rc})")
        return out, words

    def ct_mul_plain_accumulate(self, ct, w, acc=None):
        """acc (+)= ct (*) Encode(w): EvalMult(ct, plaintext vector) + EvalAdd on one ciphertext [2][L][N]."""
        ct = _u64(ct)
        w = np.ascontiguousarray(w, dtype=np.float64)
        first = acc is None
        if first:
            acc = np.empty((2, self.num_limbs, self.ring_dim), np.uint64)
        rc = lib().orc_ct_mul_plain_accumulate(self._h, ct.ctypes.data, w.ctypes.data, w.size, int(first), acc.ctypes.data)
        if rc:
            raise RuntimeError(f"ct_mul_plain_accumulate failed ({rc})")
        return acc

    # --- rescale-before-transport: PALISADE ModReduce under EXACTRESCALE ---
    def rescale_constants(self, limbs_in=None):
        """(QlQlInvModqlDivqlModq[i], qlInvModq[i]) for i < limbs_in - 1."""
        lin = self.num_limbs if limbs_in is None else limbs_in
        t, qi = np.zeros(8, np.uint64), np.zeros(8, np.uint64)
        lib().orc_rescale_constants(self._h, lin, t.ctypes.data, qi.ctypes.data)
        return [int(v) for v in t[:lin - 1]], [int(v) for v in qi[:lin - 1]]

    def rescale_ct(self, ct):
        """[2][Lin][N] -> [2][Lin-1][N]"""
        ct = _u64(ct)
        lin = ct.shape[-2]
        out = np.empty((2, lin - 1, self.ring_dim), np.uint64)
        rc = lib().orc_rescale_ct(self._h, ct.ctypes.data, lin, out.ctypes.data)
        if rc:
            raise RuntimeError(f"rescale_ct failed ({rc})")
        return out

    def decrypt_ct_level(self, ct, sf):
        """Decrypt+Decode of a ciphertext [2][limbs][N] with limbs <= L: (doubles, centred CRT words)."""
        ct = _u64(ct)
        out = np.empty(self.slots, np.float64)
        words = np.empty((self.ring_dim, 8), np.uint64)
        rc = lib().orc_decrypt_ct_level(self._h, ct.ctypes.data, ct.shape[-2], sf, out.ctypes.data, words.ctypes.data)
        if rc:
            raise RuntimeError(f"decrypt_ct_level failed ({rc})")
        return out, words

    def rescale(self, blob: bytes) -> bytes:
        out, n = C.c_void_p(), C.c_uint64()
        rc = lib().orc_rescale_blob(self._h, blob, len(blob), C.byref(out), C.byref(n))
        if rc:
            raise RuntimeError(f"oracle rescale failed ({rc}): {self.last_error()}")
        return _take(out, n.value)

    def weight_constant(self, w: float) -> int:
        return lib().orc_weight_constant(self._h, w)

    def blob_size(self, n_ct: int) -> int:
        return lib().orc_blob_size(self._h, n_ct)

    def blob_build(self, payload, depth=1, level=0, sf=None) -> bytes:
        payload = _u64(payload)
        words = 2 * (self.num_limbs - level) * self.ring_dim
        n_ct = payload.size // words
        out, n = C.c_void_p(), C.c_uint64()
        lib().orc_blob_build(self._h, n_ct, payload.ctypes.data, depth, level,
                             self.scaling_factor if sf is None else sf, C.byref(out), C.byref(n))
        return _take(out, n.value)


def blob_parse(blob: bytes, single: bool = False, want_payload: bool = True):
    info = BlobInfo()
    rc = lib().orc_blob_parse(blob, len(blob), int(single), C.byref(info), None, 0)
    if rc:
        raise ValueError(f"blob parse failed ({rc})")
    payload = None
    if want_payload:
        payload = np.empty((info.n_ct, 2, info.n_limbs, info.ring_dim), np.uint64)
        rc = lib().orc_blob_parse(blob, len(blob), int(single), C.byref(info), payload.ctypes.data, payload.size)
        if rc:
            raise ValueError(f"blob parse failed ({rc})")
    return info, payload


def blob_reemit(blob: bytes, single: bool = False) -> bytes:
    out, n = C.c_void_p(), C.c_uint64()
    rc = lib().orc_blob_reemit(blob, len(blob), int(single), C.byref(out), C.byref(n))
    if rc:
        raise ValueError(f"blob re-emit failed ({rc})")
    return _take(out, n.value)


def file_reemit(data: bytes, kind: int) -> bytes:
    """kind: 0 = cryptocontext, 1 = public key, 2 = private key."""
    out, n = C.c_void_p(), C.c_uint64()
    rc = lib().orc_file_reemit(data, len(data), kind, C.byref(out), C.byref(n))
    if rc:
        raise ValueError(f"file re-emit failed ({rc})")
    return _take(out, n.value)


def first_prime(bits, m): return lib().orc_first_prime(bits, m)
def previous_prime(q, m): return lib().orc_previous_prime(q, m)
def min_root_of_unity(m, q): return lib().orc_min_root_of_unity(m, q)
def select_ring_dim(bits, batch): return lib().orc_select_ring_dim(bits, batch)
def select_ring_dim_ex(bits, mult_depth): return lib().orc_select_ring_dim_ex(bits, mult_depth)


def modulus_chain(bits, mult_depth, ring_dim):
    q = np.zeros(8, np.uint64)
    n = lib().orc_modulus_chain(bits, mult_depth, ring_dim, q.ctypes.data)
    if n < 0:
        raise ValueError("unsupported chain")
    return [int(v) for v in q[:n]]


def words_to_int(words) -> list:
    """rows of 8 little-endian u64 (two's complement, 512 bit) -> Python ints"""
    out = []
    for row in np.asarray(words, dtype=np.uint64):
        v = sum(int(w) << (64 * i) for i, w in enumerate(row))
        out.append(v - (1 << 512) if v >> 511 else v)
    return out


def chacha20_block(key, counter: int, nonce) -> np.ndarray:
    """RFC 8439 2.3: key 8 words, block counter, nonce 3 words -> 16 words."""
    k = np.ascontiguousarray(key, dtype=np.uint32)
    n = np.ascontiguousarray(nonce, dtype=np.uint32)
    out = np.empty(16, np.uint32)
    lib().orc_chacha20_block(k.ctypes.data, counter, n.ctypes.data, out.ctypes.data)
    return out


def seed_key(seed: int, domain: int) -> np.ndarray:
    key = np.empty(8, np.uint32)
    lib().orc_seed_key(seed, domain, key.ctypes.data)
    return key
