"""Wall clock of the three Python drop-in calls on the config-2 workload, `bytes` results and PinnedBlob results.
python tools/py_api_phases.py"""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "fhe-fed_b200"))
import SHELFI_FHE  # noqa: E402

P, K = 1663370, 10
h = SHELFI_FHE.CKKS("ckks", 4096, 52, str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/", device=0)
h.loadCryptoParams()
rng = np.random.default_rng(2)
xs = [rng.normal(0, 0.05, P) for _ in range(K)]
w = [1.0 / K] * K
expect = sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, xs))
for mode in ("bytes", "pinned"):
    h.pinned = mode == "pinned"
    best = None
    for rep in range(5):
        t0 = time.perf_counter(); blobs = [h.encrypt(x) for x in xs]; t1 = time.perf_counter()
        avg = h.computeWeightedAverage(blobs, w); t2 = time.perf_counter()
        out = h.decrypt(avg, P); t3 = time.perf_counter()
        del blobs, avg
        row = {"encrypt_x10": (t1 - t0) * 1e3, "wavg": (t2 - t1) * 1e3, "decrypt": (t3 - t2) * 1e3, "step": (t3 - t0) * 1e3}
        if rep and (best is None or row["step"] < best["step"]):
            best = row
    assert np.abs(out - expect).max() / np.abs(expect).max() < 1e-6
    print(json.dumps({"tool": "py_api_phases", "mode": mode, "ms": best, "params_per_s": P / (best["step"] / 1e3)}))
