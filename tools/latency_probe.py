#!/usr/bin/env python
"""Per-call wall-clock latency of the Python drop-in on small inputs (BASELINE config 1 sizes)."""
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fhe-fed_b200"):
    sys.path.insert(0, str(p))
import SHELFI_FHE  # noqa: E402

d = str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/"
for pinned in (False, True):
    h = SHELFI_FHE.CKKS("ckks", 4096, 52, d, device=0, pinned=pinned)
    h.loadCryptoParams()
    for n in (100, 10_000, 100_000):
        xs = [np.random.default_rng(i).random(n) for i in range(3)]
        w = [0.5, 0.2, 0.3]
        best = {}
        for _ in range(7):
            t = [time.perf_counter()]
            blobs = [h.encrypt(x) for x in xs]; t.append(time.perf_counter())
            avg = h.computeWeightedAverage(blobs, w); t.append(time.perf_counter())
            out = h.decrypt(avg, n); t.append(time.perf_counter())
            for k, v in zip(("encrypt_x3", "wavg", "decrypt"), np.diff(t)):
                best[k] = min(best.get(k, 1e9), v)
        print(("pinned" if pinned else "bytes "), n, {k: round(v * 1e3, 3) for k, v in best.items()}, "ms", flush=True)
