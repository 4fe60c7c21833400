"""Where the host-buffer (`e2e`) step spends its time: per-call wall clock of the three C-ABI calls on page-locked
buffers next to the raw PCIe copy rates of the same box (one B200).  python tools/e2e_phases.py"""
import ctypes as C
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "fhe-fed_b200"))
import SHELFI_FHE  # noqa: E402

P, K = 1663370, 10


def main():
    L = SHELFI_FHE.lib()
    h = SHELFI_FHE.CKKS("ckks", 4096, 52, str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/", device=0)
    h.loadCryptoParams()
    n_ct = h.num_ciphertexts(P)
    cap = h.blob_size(n_ct)
    rng = np.random.default_rng(2)
    xs = [torch.from_numpy(rng.normal(0, 0.05, P)).pin_memory() for _ in range(K)]
    blobs = [torch.empty(cap, dtype=torch.uint8).pin_memory() for _ in range(K)]
    avg = torch.empty(cap, dtype=torch.uint8).pin_memory()
    out = torch.empty(P, dtype=torch.float64).pin_memory()
    bp = (C.c_void_p * K)(*[t.data_ptr() for t in blobs])
    bl = (C.c_uint64 * K)(*[cap] * K)
    w = (C.c_float * K)(*[1.0 / K] * K)
    olen = C.c_uint64()
    ck = h._check

    def step(t):
        for k in range(K):
            t0 = time.perf_counter()
            ck(L.shelfi_encrypt_into(h.handle, C.c_void_p(xs[k].data_ptr()), P, C.c_void_p(blobs[k].data_ptr()), cap, C.byref(olen)))
            t["encrypt"].append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        ck(L.shelfi_compute_weighted_average_into(h.handle, bp, bl, w, K, C.c_void_p(avg.data_ptr()), cap, C.byref(olen)))
        t["wavg"].append(time.perf_counter() - t0)
        t0 = time.perf_counter()
        ck(L.shelfi_decrypt(h.handle, C.c_void_p(avg.data_ptr()), olen.value, P, C.c_void_p(out.data_ptr())))
        t["decrypt"].append(time.perf_counter() - t0)

    step({"encrypt": [], "wavg": [], "decrypt": []})
    t = {"encrypt": [], "wavg": [], "decrypt": []}
    t0 = time.perf_counter()
    reps = 5
    for _ in range(reps):
        step(t)
    total = (time.perf_counter() - t0) / reps
    # raw link rates: one 1 GiB copy each way, and both at once on two streams
    nbytes = 1 << 30
    hbuf, hbuf2 = torch.empty(nbytes, dtype=torch.uint8).pin_memory(), torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dbuf, dbuf2 = torch.empty(nbytes, dtype=torch.uint8, device="cuda"), torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()

    def rate(fn, moved):
        fn(); torch.cuda.synchronize()
        a = time.perf_counter()
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        return moved * 3 / (time.perf_counter() - a) / 1e9

    def h2d():
        with torch.cuda.stream(s1):
            dbuf.copy_(hbuf, non_blocking=True)

    def d2h():
        with torch.cuda.stream(s2):
            hbuf2.copy_(dbuf2, non_blocking=True)

    def both():
        h2d(); d2h()

    link = {"h2d_GBps": rate(h2d, nbytes), "d2h_GBps": rate(d2h, nbytes), "both_GBps_total": rate(both, 2 * nbytes)}
    ms = {k: 1e3 * float(np.mean(v)) for k, v in t.items()}
    moved = {"encrypt": (8 * P, cap), "wavg": (K * cap, cap), "decrypt": (cap, 8 * P)}
    ideal = {k: 1e3 * (moved[k][0] / link["h2d_GBps"] + moved[k][1] / link["d2h_GBps"]) / 1e9 for k in moved}
    print(json.dumps({"tool": "e2e_phases", "ms_per_step": 1e3 * total, "params_per_s": P / total,
                      "ms_per_call": ms, "calls_per_step": {"encrypt": K, "wavg": 1, "decrypt": 1},
                      "ms_if_only_the_link": ideal, "link": link,
                      "step_ms_if_only_the_link": K * ideal["encrypt"] + ideal["wavg"] + ideal["decrypt"]}))


if __name__ == "__main__":
    main()
