#!/usr/bin/env python
"""Forward-NTT row rate on single-limb contexts: a 52-bit prime (FP64-pipe path unless SHELFI_FP64_NTT=0) and a
59-bit prime (integer path).  Key material comes from the oracle (test infrastructure; this is a measurement tool)."""
import ctypes as C
import json
import sys
import tempfile
from pathlib import Path

ROOT = Path(__file__).resolve().parent.parent
for p in (ROOT, ROOT / "fhe-fed_b200"):
    sys.path.insert(0, str(p))
import torch  # noqa: E402

import oracle  # noqa: E402
import SHELFI_FHE  # noqa: E402

L = SHELFI_FHE.lib()
for bits in (52, 59):
    d = tempfile.mkdtemp() + "/"
    o = oracle.Oracle("ckks", 4096, bits, d)
    assert o.genCryptoContextAndKeyGen(seed=1, nine_int=True, mult_depth=0, ring_dim=8192) == 1
    h = SHELFI_FHE.CKKS("ckks", 4096, bits, d, device=0)
    h.loadCryptoParams()
    assert h.load_status == 0 and h.num_limbs == 1
    rows = 8140
    src = torch.randint(0, 2**bits - 2**(bits - 2), (rows * 8192,), dtype=torch.int64, device="cuda")
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    out = {}
    ref = None
    for name, fn in (("smem_5_pass", L.shelfi_ntt_forward_device), ("lane_exchange_last_stage", L.shelfi_ntt_forward_shfl_device)):
        t = src.clone()
        assert fn(h.handle, C.c_void_p(t.data_ptr()), rows, sp) == 0
        torch.cuda.synchronize()
        if ref is None:
            ref = t.clone()
        assert torch.equal(t, ref), name            # bit-identical results
        for _ in range(3):
            fn(h.handle, C.c_void_p(t.data_ptr()), rows, sp)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(10):
            fn(h.handle, C.c_void_p(t.data_ptr()), rows, sp)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 10
        out[name] = {"ms": round(ms, 4), "ns_per_row": round(ms * 1e6 / rows, 1), "GBps": round(16 * 8192 * rows / ms / 1e6, 1)}
    print(json.dumps({"bits": bits, "rows": rows, "fp64_pipe": bool(bits <= 52), **out}), flush=True)
