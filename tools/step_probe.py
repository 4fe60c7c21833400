#!/usr/bin/env python
"""One device-resident step of BASELINE configs[1] (K learners' encryptions in one launch, weighted sum, decrypt) plus
a standalone forward / inverse NTT batch, with the library's per-kernel CUDA-event timers.  The short command that
ncu wraps and that kernel variants are A/B-timed with; bench.py holds the contract numbers.
usage: step_probe.py [--steps N] [--k K] [--per-learner]"""
import argparse
import ctypes as C
import json
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fhe-fed_b200"):
    sys.path.insert(0, str(p))
import torch  # noqa: E402

import SHELFI_FHE  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--k", type=int, default=10)
ap.add_argument("--per-learner", action="store_true", help="one encrypt launch per learner (what the drop-in's calls do)")
args = ap.parse_args()
P, K, BATCH = 1_663_370, args.k, 4096
L = SHELFI_FHE.lib()
h = SHELFI_FHE.CKKS("ckks", BATCH, 52, str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/", device=0)
h.loadCryptoParams()
n_ct, words = h.num_ciphertexts(P), h.ciphertext_words
rng = np.random.default_rng(2)
xs = [torch.from_numpy(rng.normal(0, 0.05, P).astype(np.float32).astype(np.float64)).cuda() for _ in range(K)]
cts = [torch.empty(n_ct * words, dtype=torch.int64, device="cuda") for _ in range(K)]
avg = torch.empty(n_ct * words, dtype=torch.int64, device="cuda")
out = torch.empty(P, dtype=torch.float64, device="cuda")
xp = (C.c_void_p * K)(*[t.data_ptr() for t in xs])
cp = (C.c_void_p * K)(*[t.data_ptr() for t in cts])
w = (C.c_float * K)(*[1.0 / K] * K)
st = torch.cuda.current_stream()
sp = C.c_void_p(st.cuda_stream)
sf2 = h.scaling_factor ** 2


def ck(rc):
    if rc:
        raise RuntimeError(h._err())


def step():
    if args.per_learner:
        for i in range(K):
            ck(L.shelfi_encrypt_device(h.handle, C.c_void_p(xs[i].data_ptr()), P, 0, C.c_void_p(cts[i].data_ptr()), sp))
    else:
        ck(L.shelfi_encrypt_many_device(h.handle, xp, K, P, 0, cp, sp))
    ck(L.shelfi_weighted_sum_device(h.handle, cp, w, K, n_ct, C.c_void_p(avg.data_ptr()), sp))
    ck(L.shelfi_decrypt_device(h.handle, C.c_void_p(avg.data_ptr()), n_ct, sf2, C.c_void_p(out.data_ptr()), P, sp))


for _ in range(3):
    step()
torch.cuda.synchronize()
expect = sum(np.float64(np.float32(1.0 / K)) * x.cpu().numpy() for x in xs)
rel = float(np.abs(out.cpu().numpy() - expect).max() / np.abs(expect).max())
assert rel < 1e-6, rel
L.shelfi_profile_reset(h.handle)
L.shelfi_profile_enable(h.handle, 1)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(st)
for _ in range(args.steps):
    step()
e1.record(st)
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / args.steps
L.shelfi_profile_enable(h.handle, 0)
prof = {}
for i, name in enumerate(["encode_sample", "encrypt", "weighted_sum", "decrypt_intt", "decode"]):
    tot, cnt = C.c_double(), C.c_uint64()
    L.shelfi_profile_get(h.handle, i, C.byref(tot), C.byref(cnt))
    prof[name] = round(tot.value / args.steps, 4)
rows = K * n_ct * 2
d_rows = torch.randint(0, 2**50, (rows * 8192,), dtype=torch.int64, device="cuda")
res = {}
for name, fn in (("ntt_fwd", L.shelfi_ntt_forward_device), ("ntt_inv", L.shelfi_ntt_inverse_device)):
    for _ in range(2):
        ck(fn(h.handle, C.c_void_p(d_rows.data_ptr()), rows, sp))
    e0.record(st)
    for _ in range(5):
        ck(fn(h.handle, C.c_void_p(d_rows.data_ptr()), rows, sp))
    e1.record(st)
    torch.cuda.synchronize()
    t = e0.elapsed_time(e1) / 5
    res[name] = {"ms": round(t, 4), "ns_per_row": round(t * 1e6 / rows, 1), "GBps": round(16 * 8192 * rows / t / 1e6, 1)}
print(json.dumps({"ms_per_step": round(ms, 4), "Mparams_per_s": round(P / ms / 1e3, 1), "kernels_ms": prof, **res,
                  "rel_err": rel, "per_learner": args.per_learner, "k": K}))
