"""Ciphertext-index sharding across the GPUs of one box (one process per GPU, torch.distributed plumbing).

The reference has no distributed code: its three calls loop over ciphertext indices independently
(ckks.cpp:71-83 chunks, :286-297 per-index scale/add, :187-205 per-ciphertext decrypt), and the learner sum is
*within* an index, so the path shards with no data-path collective.  Each rank owns a contiguous block of
ciphertext indices, i.e. a contiguous parameter range; the only collective is an optional all-gather of the
decoded float64 slices when the caller wants one contiguous buffer (NCCL over NVLink on GPUs, gloo in CPU tests).
"""
from __future__ import annotations

from dataclasses import dataclass


@dataclass(frozen=True)
class Shard:
    rank: int
    world: int
    ct_first: int      # first ciphertext index owned by this rank (the sampler's ct_index0)
    ct_count: int      # ciphertexts owned (0 for trailing ranks of a tiny model)
    param_first: int   # first parameter index
    param_count: int   # parameters owned (the last owner gets the short tail)
    ct_per_rank: int   # common block size: ceil(n_ct / world)


def shard_range(n_params: int, batch: int, rank: int, world: int) -> Shard:
    """Contiguous block partition of ceil(n_params / batch) ciphertexts over `world` ranks."""
    if n_params <= 0 or batch <= 0 or world <= 0 or not 0 <= rank < world:
        raise ValueError("bad shard arguments")
    n_ct = -(-n_params // batch)
    per = -(-n_ct // world)
    first = min(rank * per, n_ct)
    count = min(per, n_ct - first)
    p_first = first * batch
    p_count = max(0, min(n_params, (first + count) * batch) - p_first) if count else 0
    return Shard(rank, world, first, count, p_first, p_count, per)


def gather_decoded(local, n_params: int, batch: int, group=None):
    """All-gather the ranks' decoded slices into one contiguous float64 tensor of n_params values.

    `local` is this rank's decoded slice (torch tensor, on the device the process group works on). Slices are
    padded to the common block size so one all_gather_into_tensor suffices; the padding is dropped afterwards.
    """
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    sh = shard_range(n_params, batch, rank, world)
    if local.numel() != sh.param_count:
        raise ValueError(f"rank {rank}: slice has {local.numel()} values, shard owns {sh.param_count}")
    block = sh.ct_per_rank * batch
    padded = torch.zeros(block, dtype=local.dtype, device=local.device)
    padded[:sh.param_count] = local
    out = torch.empty(world * block, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded, group=group)
    return out[:n_params]


class PeerGather:
    """Decoded-model reassembly with no collective: every rank owns one device buffer of `world * block` float64 that
    its peers map (CUDA IPC through the C ABI), and `decrypt` makes the decode kernel store this rank's slice into all
    of them at the slice's offset (shelfi_decrypt_gather_device).  The stores ride NVLink while the kernel computes;
    what remains of the all-gather is a flag exchange between the GPUs, queued on the same stream before and after the
    decode (shelfi_peer_barrier_device) — no host synchronisation, no NCCL launch.  `torch.distributed` only carries
    the 64-byte handles at construction.

        pg = PeerGather(ckks, n_params, batch)              # collective: every rank of `group` calls it
        pg.decrypt(d_ct_ptr, n_ct, scaling_factor, stream)  # this rank's ciphertexts -> everybody's buffer
        full = pg.tensor()[:n_params]                       # valid for work queued on `stream` after decrypt()
    """

    FLAG_WORDS = 16    # 8 rank slots + status slot, padded to 128 bytes

    def __init__(self, ckks, n_params: int, batch: int, group=None):
        import ctypes as C

        import torch.distributed as dist

        from . import lib

        self._C, self._lib, self._ckks, self._dist, self._group = C, lib(), ckks, dist, group
        self.world, self.rank = dist.get_world_size(group), dist.get_rank(group)
        if self.world > 8:
            raise ValueError("PeerGather covers the GPUs of one box (at most 8 ranks)")
        self.shard = shard_range(n_params, batch, self.rank, self.world)
        self.n_params, self.block = n_params, self.shard.ct_per_rank * batch
        self.values = self.world * self.block
        self._flag_off = (self.values * 8 + 255) // 256 * 256
        self._epoch = 0
        # Every rank goes through the same collectives whatever fails locally, and all of them raise together: a rank
        # that dropped out alone would leave the others waiting in the next collective.
        own, handle = C.c_void_p(), C.create_string_buffer(64)
        rc = self._lib.shelfi_peer_buffer_create(ckks.handle, self._flag_off + 8 * self.FLAG_WORDS, C.byref(own), handle)
        self._own = own.value if rc == 0 else None
        handles = [None] * self.world
        dist.all_gather_object(handles, bytes(handle.raw) if rc == 0 else None, group=group)
        self._ptrs, failed = [], None
        if all(hb is not None for hb in handles):
            for r, hb in enumerate(handles):
                if r == self.rank:
                    self._ptrs.append(self._own)
                    continue
                p = C.c_void_p()
                if self._lib.shelfi_peer_buffer_open(ckks.handle, hb, C.byref(p)) != 0:
                    failed = f"rank {self.rank}: cannot map rank {r}'s buffer: {ckks._err()}"
                    break
                self._ptrs.append(p.value)
        else:
            failed = f"rank {self.rank}: peer buffer allocation failed on rank(s) {[r for r, hb in enumerate(handles) if hb is None]}"
        verdicts = [None] * self.world
        dist.all_gather_object(verdicts, failed, group=group)
        if any(v is not None for v in verdicts):
            for r, p in enumerate(self._ptrs):
                if r != self.rank:
                    self._lib.shelfi_peer_buffer_close(ckks.handle, C.c_void_p(p))
            dist.barrier(group=group)
            if self._own is not None:
                self._lib.shelfi_peer_buffer_destroy(ckks.handle, C.c_void_p(self._own))
            self._own = None
            raise RuntimeError("; ".join(v for v in verdicts if v is not None))
        # own copy first (shelfi_decrypt_gather_device's d_outs[0]), then the peers in rank order
        order = [self.rank] + [r for r in range(self.world) if r != self.rank]
        self._outs = (C.c_void_p * self.world)(*[self._ptrs[r] for r in order])
        self._flags = (C.c_void_p * self.world)(*[p + self._flag_off for p in self._ptrs])

    def device_barrier(self, stream: int = 0, guard: bool = False) -> None:
        """Queue a barrier between the ranks' streams (every rank calls it the same number of times).  guard=True (the
        barrier after a decrypt): should the wait for a rank time out, the first value of that rank's slice in this
        GPU's copy becomes NaN, so a partial gather cannot pass for a whole one (`timed_out()` tells which epoch)."""
        self._epoch += 1
        if guard:
            self._ckks._check(self._lib.shelfi_peer_barrier_guard_device(
                self._ckks.handle, self._flags, self.world, self.rank, self._epoch, self._C.c_void_p(self._own),
                self.block, self._C.c_void_p(stream)))
            return
        self._ckks._check(self._lib.shelfi_peer_barrier_device(self._ckks.handle, self._flags, self.world, self.rank,
                                                               self._epoch, self._C.c_void_p(stream)))

    def decrypt(self, d_ct_ptr: int, n_ct: int, scaling_factor: float, stream: int = 0, device_barriers: bool = True) -> None:
        """Queue on `stream`: barrier (the peers have consumed the previous result), decrypt + scatter of this rank's
        `n_ct` ciphertexts, barrier (every slice has landed here).  Returns at once; with device_barriers=False the
        caller orders the ranks itself (synchronise the stream, then `barrier()`)."""
        C, sh = self._C, self.shard
        if n_ct != sh.ct_count:
            raise ValueError(f"rank {self.rank}: {n_ct} ciphertexts, shard owns {sh.ct_count}")
        if device_barriers:
            self.device_barrier(stream)
        if n_ct:
            self._ckks._check(self._lib.shelfi_decrypt_gather_device(
                self._ckks.handle, C.c_void_p(d_ct_ptr), n_ct, scaling_factor, self._outs, self.world, sh.param_first,
                sh.param_count, C.c_void_p(stream)))
        if device_barriers:
            self.device_barrier(stream, guard=True)

    def timed_out(self) -> int:
        """Epoch of the last barrier this rank gave up on (0 = never); synchronises the device."""
        import torch

        torch.cuda.synchronize(self._ckks.device)
        flags = torch.as_tensor(_RawDevice(self._own + self._flag_off, (self.FLAG_WORDS,), "<i8"),
                                device=f"cuda:{self._ckks.device}")
        return int(flags[8].item())

    def barrier(self) -> None:
        """Host-side barrier of the process group (the caller has synchronised its own stream)."""
        self._dist.barrier(group=self._group)

    def tensor(self):
        """This rank's buffer as a torch tensor (no copy); [:n_params] is the decoded model."""
        import torch

        return torch.as_tensor(_RawDevice(self._own, (self.values,), "<f8"), device=f"cuda:{self._ckks.device}")

    def close(self) -> None:
        if self._own is None:
            return
        import torch

        torch.cuda.synchronize(self._ckks.device)
        for r, p in enumerate(self._ptrs):
            if r != self.rank:
                self._lib.shelfi_peer_buffer_close(self._ckks.handle, self._C.c_void_p(p))
        self._dist.barrier(group=self._group)      # nobody has the buffer mapped any more
        self._lib.shelfi_peer_buffer_destroy(self._ckks.handle, self._C.c_void_p(self._own))
        self._own = None


class _RawDevice:
    """__cuda_array_interface__ view of library-owned device memory (torch wraps it without copying)."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (ptr, False), "version": 3,
                                         "strides": None}
