"""Probe: encrypt_device time per ciphertext vs batch size (wave effects)."""
import ctypes as C, sys
from pathlib import Path
ROOT = Path(__file__).resolve().parent.parent
sys.path[:0] = [str(ROOT), str(ROOT / "fhe-fed_b200")]
import numpy as np, torch
import SHELFI_FHE
L = SHELFI_FHE.lib()
h = SHELFI_FHE.CKKS("ckks", 4096, 52, str(ROOT / "tests/golden/pybind_cryptoparams") + "/", device=0)
h.loadCryptoParams()
sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
for n_ct in (148, 296, 407, 444, 592, 814, 888, 1332, 1776, 3552):
    n = n_ct * 4096
    x = torch.randn(n, dtype=torch.float64, device="cuda") * 0.05
    ct = torch.empty(n_ct * h.ciphertext_words, dtype=torch.int64, device="cuda")
    for _ in range(3):
        L.shelfi_encrypt_device(h.handle, C.c_void_p(x.data_ptr()), n, 0, C.c_void_p(ct.data_ptr()), sp)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); a.record()
    for _ in range(10):
        L.shelfi_encrypt_device(h.handle, C.c_void_p(x.data_ptr()), n, 0, C.c_void_p(ct.data_ptr()), sp)
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / 10
    print(f"n_ct {n_ct:5d}  {ms:8.4f} ms  {1e6*ms/n_ct:8.2f} ns/ct  {n/ms/1e3:8.1f} M params/s")
