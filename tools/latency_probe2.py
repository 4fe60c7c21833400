#!/usr/bin/env python
"""Per-call wall clock of BASELINE configs[0] (the pythonApi example: 3 learners, weights .5/.2/.3) through the drop-in,
in bytes mode (the reference's types) and in resident mode.  usage: latency_probe2.py [n ...]"""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fhe-fed_b200"):
    sys.path.insert(0, str(p))
import SHELFI_FHE  # noqa: E402

d = str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/"
sizes = [int(a) for a in sys.argv[1:]] or [10_000, 100_000]
for mode in ("bytes", "pinned", "resident"):
    h = SHELFI_FHE.CKKS("ckks", 4096, 52, d, device=0, pinned=mode == "pinned", resident=mode == "resident")
    h.loadCryptoParams()
    for n in sizes:
        data = [np.random.default_rng(1 + i).random(n) for i in range(3)]
        w = [0.5, 0.2, 0.3]
        best = None
        for it in range(30):
            t = [time.perf_counter()]
            blobs = []
            for x in data:
                blobs.append(h.encrypt(x)); t.append(time.perf_counter())
            avg = h.computeWeightedAverage(blobs, w); t.append(time.perf_counter())
            out = h.decrypt(avg, n); t.append(time.perf_counter())
            dt = np.diff(t)
            if it >= 5 and (best is None or dt.sum() < best.sum()):
                best = dt
        ref = sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, data))
        assert np.abs(out - ref).max() / np.abs(ref).max() < 1e-6
        print(json.dumps({"mode": mode, "n": n, "total_us": round(1e6 * best.sum(), 1),
                          "encrypt_us": [round(1e6 * v, 1) for v in best[:3]], "wavg_us": round(1e6 * best[3], 1),
                          "decrypt_us": round(1e6 * best[4], 1)}), flush=True)
