"""Writes tests/golden/vectors_{pybind,code}.npz from the CPU oracle on seeded inputs, so the GPU box can
check the CUDA path against committed outputs.  Run from the repo root: python tests/golden/make_vectors.py"""
import hashlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent.parent
sys.path.insert(0, str(ROOT))
import oracle  # noqa: E402

for name, sub in (("pybind", "pybind_cryptoparams"), ("code", "code_cryptoparams")):
    o = oracle.Oracle(cryptodir=str(ROOT / "tests" / "golden" / sub) + "/")
    assert o.loadCryptoParams() == 0
    rng = np.random.default_rng(20261104)
    n, seed = 10000, 1234567
    inputs = rng.normal(0, 0.5, (3, n))
    weights = np.array([0.5, 0.2, 0.3])
    blobs = [o.encrypt(x, seed=seed) for x in inputs]
    avg = o.computeWeightedAverage(blobs, list(weights))
    np.savez_compressed(ROOT / "tests" / "golden" / f"vectors_{name}.npz", inputs=inputs, weights=weights, seed=seed,
                        encrypt_sha256=np.array([hashlib.sha256(b).hexdigest() for b in blobs]),
                        wavg_sha256=hashlib.sha256(avg).hexdigest(), decrypted=o.decrypt(avg, n))
    print(name, "ok")
