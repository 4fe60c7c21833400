#!/usr/bin/env python
"""PCIe copy times on this box for the sizes of the drop-in's step (one learner's float32 upload, the float64 result):
whole and in four pieces, wall clock around a synchronised copy into / out of page-locked memory."""
import json
import time

import torch

res = {}
for name, nbytes in (("h2d_6.65MB", 1_663_370 * 4), ("d2h_13.3MB", 1_663_370 * 8)):
    host = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dev = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    for pieces in (1, 4):
        step = -(-nbytes // pieces)
        ts = []
        for it in range(30):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for p in range(pieces):
                a, b = p * step, min(nbytes, (p + 1) * step)
                if name.startswith("h2d"):
                    dev[a:b].copy_(host[a:b], non_blocking=True)
                else:
                    host[a:b].copy_(dev[a:b], non_blocking=True)
            torch.cuda.synchronize()
            ts.append(time.perf_counter() - t0)
        ts.sort()
        med = ts[len(ts) // 2]
        res[f"{name}_x{pieces}"] = {"ms": round(1e3 * med, 4), "GBps": round(nbytes / med / 1e9, 1)}
print(json.dumps(res))
