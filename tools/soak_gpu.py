#!/usr/bin/env python
"""Randomised soak of the public API against the CPU oracle (bit-exact blobs, identical decoded doubles).
Contexts: the two fixture key sets plus generated ones (batch 1024/2048/4096, bits 20..59, depths 1..3, ring
dimensions 8192..32768).  Each case: K learners x n values -> encrypt (seeded) -> weighted average -> decrypt.
usage: python tools/soak_gpu.py [cases] [seed]      (the oracle is the checker here, as in tests/)"""
import sys
import tempfile
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fhe-fed_b200"):
    sys.path.insert(0, str(p))
import oracle  # noqa: E402
import SHELFI_FHE  # noqa: E402


def contexts(rng):
    g = ROOT / "tests" / "golden"
    yield "fixture-pybind", 4096, 52, str(g / "pybind_cryptoparams") + "/", None
    yield "fixture-code", 4096, 52, str(g / "code_cryptoparams") + "/", None
    for batch, bits, depth, ring in ((1024, 20, 1, 0), (2048, 40, 1, 0), (4096, 59, 1, 0), (4096, 33, 1, 0),
                                     (4096, 52, 2, 0), (8192, 52, 3, 0), (2048, 45, 3, 16384), (16384, 50, 3, 32768),
                                     (512, 52, 5, 32768), (4096, 14, 1, 0)):
        yield f"gen-b{batch}-s{bits}-d{depth}-r{ring}", batch, bits, tempfile.mkdtemp() + "/", (depth, ring)


def main():
    cases = int(sys.argv[1]) if len(sys.argv) > 1 else 60
    rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
    done = 0
    t0 = time.time()
    ctxs = list(contexts(rng))
    per_ctx = max(1, cases // len(ctxs))
    for name, batch, bits, d, gen in ctxs:
        o = oracle.Oracle("ckks", batch, bits, d)
        h = SHELFI_FHE.CKKS("ckks", batch, bits, d, device=0)
        if gen:
            assert o.genCryptoContextAndKeyGen(seed=int(rng.integers(1, 2**31)), mult_depth=gen[0], ring_dim=gen[1]) == 1
        assert o.loadCryptoParams() == 0
        h.loadCryptoParams()
        assert h.load_status == 0, name
        hr = SHELFI_FHE.CKKS("ckks", batch, bits, d, device=0, resident=True)   # the same calls on resident vectors
        hr.loadCryptoParams()
        assert (h.ring_dim, h.num_limbs) == (o.ring_dim, o.num_limbs)
        for _ in range(per_ctx):
            k = int(rng.integers(1, 7))
            n = int(rng.choice([1, batch - 1, batch, batch + 1, int(rng.integers(1, 6 * batch))]))
            # |sum w x| * Delta^2 must stay below Q/2: generated chains with few bits have little headroom
            amp = 10.0 ** rng.uniform(-3, 0 if bits >= 40 else -1)
            xs = [rng.normal(0, amp, n) for _ in range(k)]
            w = [float(v) for v in rng.uniform(-1, 1, k) / k]
            seed = int(rng.integers(0, 2**62))
            h.set_sampler_seed(seed)
            blobs = [h.encrypt(x) for x in xs]
            h.set_sampler_seed(None)
            for b, x in zip(blobs, xs):
                assert b == o.encrypt(x, seed=seed), (name, "encrypt", k, n)
            avg = h.computeWeightedAverage(blobs, w)
            assert avg == o.computeWeightedAverage(blobs, w), (name, "wavg", k, n)
            got = h.decrypt(avg, n)
            assert np.array_equal(got, o.decrypt(avg, n)), (name, "decrypt", k, n)
            # resident mode: handles from encrypt (float32 / float64 inputs), uploaded blobs and plain bytes mixed
            hr.set_sampler_seed(seed)
            enc_r = [hr.encrypt(x) for x in xs]
            hr.set_sampler_seed(None)
            assert all(len(a) == len(b) for a, b in zip(enc_r, blobs))
            mixed = [enc_r[i] if i % 3 == 0 else (hr.to_resident(blobs[i]) if i % 3 == 1 else blobs[i]) for i in range(k)]
            avg_r = hr.computeWeightedAverage(mixed, w)
            assert bytes(avg_r) == avg, (name, "resident wavg", k, n)
            assert bytes(enc_r[-1]) == blobs[-1], (name, "resident encrypt", k, n)
            assert np.array_equal(hr.decrypt(avg_r, n), got), (name, "resident decrypt", k, n)
            del enc_r, mixed, avg_r
            expect = sum(np.float64(np.float32(wi)) * x for wi, x in zip(w, xs))
            tol = max(1e-6 * np.abs(expect).max(), 2.0 ** (26 - bits))
            assert np.abs(got - expect).max() <= tol, (name, "plain", k, n, np.abs(got - expect).max(), tol)
            done += 1
        print(f"{name}: ok ({per_ctx} cases, N={h.ring_dim}, L={h.num_limbs})", flush=True)
    print(f"soak ok: {done} cases in {time.time() - t0:.1f}s")


if __name__ == "__main__":
    main()
