#!/usr/bin/env python
"""Per-layer calls (the reference's loop, benchmark.py:490-493) vs the *_layers entry points, host buffers, wall clock.
Models: the config-2 CNN (8 state_dict entries) and a ResNet-18-shaped list of 62 tensors."""
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fhe-fed_b200"):
    sys.path.insert(0, str(p))
import SHELFI_FHE  # noqa: E402

CNN = [800, 32, 51200, 64, 1605632, 512, 5120, 10]
RESNET18 = [9408, 64, 64] + [36864, 64, 64] * 4 + [73728, 128, 128, 147456, 128, 128, 8192, 128, 128, 147456, 128, 128,
            147456, 128, 128] + [294912, 256, 256, 589824, 256, 256, 32768, 256, 256, 589824, 256, 256, 589824, 256, 256] + \
           [1179648, 512, 512, 2359296, 512, 512, 131072, 512, 512, 2359296, 512, 512, 2359296, 512, 512] + [512000, 1000]


def run(name, sizes, K, h, tag=""):
    rng = np.random.default_rng(1)
    models = [[rng.normal(0, 0.05, n) for n in sizes] for _ in range(K)]
    w = [1.0 / K] * K
    res = {}
    for mode in ("per_layer", "layers"):
        best = None
        enc = avg = dec = None
        for _ in range(3):
            del enc, avg, dec                      # hand the previous pass's buffers back before timing the next
            t0 = time.perf_counter()
            if mode == "per_layer":
                enc = [[h.encrypt(a) for a in m] for m in models]
                avg = [h.computeWeightedAverage([enc[i][s] for i in range(K)], w) for s in range(len(sizes))]
                dec = [h.decrypt(avg[s], n) for s, n in enumerate(sizes)]
            else:
                enc = [h.encrypt_layers(m) for m in models]
                avg = h.computeWeightedAverage_layers(enc, w)
                dec = h.decrypt_layers(avg, sizes)
            dt = time.perf_counter() - t0
            best = dt if best is None else min(best, dt)
        res[mode] = best
    P = sum(sizes)
    print(json.dumps({"model": name, "mode": tag, "tensors": len(sizes), "params": P, "learners": K,
                      "per_layer_calls_s": round(res["per_layer"], 4), "layers_api_s": round(res["layers"], 4),
                      "params_per_s_per_layer": P / res["per_layer"], "params_per_s_layers": P / res["layers"]}), flush=True)


if __name__ == "__main__":
    d = str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/"
    h = SHELFI_FHE.CKKS("ckks", 4096, 52, d, device=0)
    h.loadCryptoParams()
    run("cnn_fedavg", CNN, 10, h, "bytes")
    run("resnet18_shaped", RESNET18, 10, h, "bytes")
    hp = SHELFI_FHE.CKKS("ckks", 4096, 52, d, device=0, pinned=True)
    hp.loadCryptoParams()
    run("cnn_fedavg", CNN, 10, hp, "pinned blobs")
    run("resnet18_shaped", RESNET18, 10, hp, "pinned blobs")
