#!/usr/bin/env python
"""Where the drop-in's end-to-end step goes (BASELINE configs[1], resident mode, float32 page-locked inputs): wall
clock of the full step and of its prefixes (encrypt x K; + computeWeightedAverage; + decrypt), each ended by a
synchronize, plus the host time spent inside the calls.  Documentation numbers for profiles/, not the bench contract.
usage: e2e_probe.py [--steps N]"""
import argparse
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fhe-fed_b200"):
    sys.path.insert(0, str(p))
import SHELFI_FHE  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=20)
ap.add_argument("--pageable", action="store_true", help="ordinary numpy arrays instead of page-locked ones")
ap.add_argument("--f64", action="store_true", help="float64 inputs")
args = ap.parse_args()
P, K, BATCH = 1_663_370, 10, 4096
h = SHELFI_FHE.CKKS("ckks", BATCH, 52, str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/", device=0, resident=True)
h.loadCryptoParams()
rng = np.random.default_rng(2)
xs = [rng.normal(0, 0.05, P).astype(np.float64 if args.f64 else np.float32) for _ in range(K)]
if not args.pageable:
    xs = [SHELFI_FHE.pinned_array(x) for x in xs]
w = [1.0 / K] * K


def run(upto, steps):
    host = [0.0, 0.0, 0.0]
    h.synchronize()
    t0 = time.perf_counter()
    for _ in range(steps):
        a = time.perf_counter()
        enc = [h.encrypt(x) for x in xs]
        b = time.perf_counter()
        host[0] += b - a
        if upto >= 1:
            avg = h.computeWeightedAverage(enc, w)
            c = time.perf_counter()
            host[1] += c - b
            if upto >= 2:
                out = h.decrypt(avg, P)
                host[2] += time.perf_counter() - c
                del out
            del avg
        del enc
        h.synchronize()
    dt = time.perf_counter() - t0
    return 1e3 * dt / steps, [round(1e3 * t / steps, 4) for t in host]


res = {}
for name, upto in (("encrypt", 0), ("encrypt+wavg", 1), ("full", 2)):
    run(upto, 3)
    ms, host = run(upto, args.steps)
    res[name] = {"ms_per_step": round(ms, 4), "host_ms_in_calls[enc,wavg,dec]": host}
res["Mparams_per_s_full"] = round(P / res["full"]["ms_per_step"] / 1e3, 1)
print(json.dumps(res))
import os  # noqa: E402
if os.environ.get("SHELFI_TIMELINE"):   # one more step with the per-launch device timeline (stderr)
    import ctypes as C
    L = SHELFI_FHE.lib()
    L.shelfi_profile_enable(h.handle, 1)
    L.shelfi_profile_reset(h.handle)
    run(2, 1)
    ms, n = C.c_double(), C.c_uint64()
    L.shelfi_profile_get(h.handle, 1, C.byref(ms), C.byref(n))
