"""Decrypt-gather across the GPUs of one box, checked and timed against decrypt + NCCL all-gather.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29533 \
        tools/peer_gather_check.py [--params P] [--learners K] [--iters I]

Every rank encrypts and aggregates its ciphertext-index shard of a P-parameter model (BASELINE config 2 shape by
default), then reassembles the decoded model twice: (a) shelfi_decrypt_device + dist.all_gather_into_tensor (NCCL),
(b) shelfi_decrypt_gather_device into peer-mapped buffers between two stream-ordered flag barriers.  The two results must be identical on every
rank; rank 0 prints one JSON line with both times (CUDA events on the launching stream, max over ranks)."""
import argparse
import ctypes as C
import json
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT / "fhe-fed_b200"))
import SHELFI_FHE  # noqa: E402
from SHELFI_FHE.sharding import PeerGather, gather_decoded, shard_range  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--params", type=int, default=1663370)
    ap.add_argument("--learners", type=int, default=3)
    ap.add_argument("--iters", type=int, default=20)
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device(f"cuda:{local}"))
    L = SHELFI_FHE.lib()
    h = SHELFI_FHE.CKKS("ckks", 4096, 52, str(ROOT / "tests" / "golden" / "pybind_cryptoparams") + "/", device=local)
    h.loadCryptoParams()
    assert h.load_status == 0
    batch, P, K = 4096, a.params, a.learners
    sh = shard_range(P, batch, rank, world)
    words = L.shelfi_ciphertext_words(h.handle)
    sp = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ck = h._check
    # every rank draws the whole model from one seed and keeps its slice
    rng = np.random.default_rng(2)
    xs = [rng.normal(0, 0.05, P) for _ in range(K)]
    w = [1.0 / K] * K
    lo, hi = sh.param_first, sh.param_first + sh.param_count
    d_x = [torch.from_numpy(x[lo:hi].copy()).cuda() for x in xs]
    d_ct = [torch.empty(max(1, sh.ct_count) * words, dtype=torch.int64, device="cuda") for _ in range(K)]
    d_avg = torch.empty(max(1, sh.ct_count) * words, dtype=torch.int64, device="cuda")
    vp = lambda ts: (C.c_void_p * len(ts))(*[t.data_ptr() for t in ts])  # noqa: E731
    if sh.ct_count:
        ck(L.shelfi_encrypt_many_device(h.handle, vp(d_x), K, sh.param_count, sh.ct_first, vp(d_ct), sp))
        ck(L.shelfi_weighted_sum_device(h.handle, vp(d_ct), (C.c_float * K)(*w), K, sh.ct_count, C.c_void_p(d_avg.data_ptr()), sp))
    sf2 = L.shelfi_scaling_factor(h.handle) ** 2
    d_out = torch.empty(max(1, sh.param_count), dtype=torch.float64, device="cuda")

    def nccl_path():
        if sh.ct_count:
            ck(L.shelfi_decrypt_device(h.handle, C.c_void_p(d_avg.data_ptr()), sh.ct_count, sf2,
                                       C.c_void_p(d_out.data_ptr()), sh.param_count, sp))
        return gather_decoded(d_out[:sh.param_count], P, batch)

    pg = PeerGather(h, P, batch)

    def peer_path():          # barrier, decode with peer stores, barrier: three launches on the stream, nothing else
        pg.decrypt(d_avg.data_ptr(), sh.ct_count, sf2, torch.cuda.current_stream().cuda_stream)
        return pg.tensor()[:P]

    def peer_path_host_barrier():
        pg.decrypt(d_avg.data_ptr(), sh.ct_count, sf2, torch.cuda.current_stream().cuda_stream, device_barriers=False)
        pg.barrier()
        return pg.tensor()[:P]

    def timed(fn):
        for _ in range(3):
            fn()
        torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(a.iters):
            fn()
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / a.iters], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    ref = nccl_path().clone()
    torch.cuda.synchronize()
    pg.tensor().zero_()
    torch.cuda.synchronize(); dist.barrier()
    got = peer_path()
    torch.cuda.synchronize()
    same = bool(torch.equal(ref, got))
    expect = sum(np.float32(wi).astype(np.float64) * x for wi, x in zip(w, xs))
    rel = float(np.abs(got.cpu().numpy() - expect).max() / np.abs(expect).max())
    flag = torch.tensor([int(same and rel < 1e-6)], device="cuda")
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    ms_nccl, ms_peer, ms_peer_host = timed(nccl_path), timed(peer_path), timed(peer_path_host_barrier)
    again = peer_path()
    torch.cuda.synchronize()
    flag2 = torch.tensor([int(torch.equal(ref, again) and pg.timed_out() == 0)], device="cuda")
    dist.all_reduce(flag2, op=dist.ReduceOp.MIN)
    flag = torch.minimum(flag, flag2)
    # decode alone, for scale
    ms_dec = timed(lambda: sh.ct_count and ck(L.shelfi_decrypt_device(
        h.handle, C.c_void_p(d_avg.data_ptr()), sh.ct_count, sf2, C.c_void_p(d_out.data_ptr()), sh.param_count, sp)))
    if rank == 0:
        print(json.dumps({"tool": "peer_gather_check", "n_gpus": world, "params": P, "learners": K,
                          "identical_on_every_rank": bool(flag.item()), "max_rel_err_vs_plaintext": rel,
                          "ms_decrypt_plus_nccl_all_gather": ms_nccl, "ms_decrypt_gather_device_barriers": ms_peer,
                          "ms_decrypt_gather_plus_nccl_barrier": ms_peer_host,
                          "ms_decrypt_alone": ms_dec, "bytes_gathered_per_gpu": 8 * pg.values}))
    pg.close()
    dist.destroy_process_group()
    if not flag.item():
        sys.exit(1)


if __name__ == "__main__":
    main()
