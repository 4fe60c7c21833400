#!/usr/bin/env python
"""Selective encryption on the config-2 workload (10 learners x 1,663,370 params, fixture keys), device-resident:
a fraction r of the parameters (random index set shared by the learners) is encrypted / aggregated / decrypted and
scattered over the plaintext weighted sum of all parameters.  One JSON line per ratio; r = 1.0 is the dense path
with the gather/scatter still in place, "dense" is bench.py's path for reference."""
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

P, K, BATCH = 1_663_370, 10, 4096


def main():
    L = SHELFI_FHE.lib()
    h = SHELFI_FHE.CKKS("ckks", BATCH, 52, str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/", device=0)
    h.loadCryptoParams()
    rng = np.random.default_rng(2)
    xs = [rng.normal(0, 0.05, P).astype(np.float32).astype(np.float64) for _ in range(K)]
    dx = [torch.from_numpy(x).cuda() for x in xs]
    xp = (C.c_void_p * K)(*[t.data_ptr() for t in dx])
    w = (C.c_float * K)(*[1.0 / K] * K)
    expect = sum(np.float64(np.float32(1.0 / K)) * x for x in xs)
    stream = torch.cuda.current_stream()
    sp = C.c_void_p(stream.cuda_stream)
    words = h.ciphertext_words
    sf2 = h.scaling_factor ** 2
    dout = torch.empty(P, dtype=torch.float64, device="cuda")

    def ck(rc):
        if rc:
            raise RuntimeError(h._err())

    for r in (0.01, 0.1, 0.3, 1.0, "dense"):
        n_sel = P if r == "dense" else max(1, int(round(P * r)))
        n_ct = -(-n_sel // BATCH)
        sel = np.sort(rng.permutation(P)[:n_sel]).astype(np.uint32)
        dsel = torch.from_numpy(sel.view(np.int32)).cuda()
        dct = [torch.empty(n_ct * words, dtype=torch.int64, device="cuda") for _ in range(K)]
        cp = (C.c_void_p * K)(*[t.data_ptr() for t in dct])
        davg = torch.empty(n_ct * words, dtype=torch.int64, device="cuda")

        def step():
            if r == "dense":
                ck(L.shelfi_encrypt_many_device(h.handle, xp, K, P, 0, cp, sp))
                ck(L.shelfi_weighted_sum_device(h.handle, cp, w, K, n_ct, C.c_void_p(davg.data_ptr()), sp))
                ck(L.shelfi_decrypt_device(h.handle, C.c_void_p(davg.data_ptr()), n_ct, sf2, C.c_void_p(dout.data_ptr()), P, sp))
                return
            ck(L.shelfi_encrypt_selected_many_device(h.handle, xp, K, C.c_void_p(dsel.data_ptr()), n_sel, 0, cp, sp))
            ck(L.shelfi_weighted_sum_device(h.handle, cp, w, K, n_ct, C.c_void_p(davg.data_ptr()), sp))
            ck(L.shelfi_plain_weighted_sum_device(h.handle, xp, w, K, P, C.c_void_p(dout.data_ptr()), sp))
            ck(L.shelfi_decrypt_scatter_device(h.handle, C.c_void_p(davg.data_ptr()), n_ct, sf2,
                                               C.c_void_p(dout.data_ptr()), C.c_void_p(dsel.data_ptr()), n_sel, sp))

        for _ in range(3):
            step()
        torch.cuda.synchronize()
        rel = float(np.abs(dout.cpu().numpy() - expect).max() / np.abs(expect).max())
        assert rel < 1e-6, rel
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        steps = 20
        e0.record(stream)
        for _ in range(steps):
            step()
        e1.record(stream)
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / steps
        print(json.dumps({"ratio": r, "encrypted_params": n_sel, "ciphertexts_per_learner": n_ct, "ms_per_step": round(ms, 4),
                          "model_params_per_s": P / (ms / 1e3), "max_rel_err": rel}), flush=True)


if __name__ == "__main__":
    main()
