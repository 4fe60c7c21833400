"""Host-side multi-GPU logic on the CPU: the block partition by ciphertext index and the single all-gather of
decoded slices, exercised with torch.distributed (gloo, world_size 2 and 3).  No compute calls."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from SHELFI_FHE.sharding import gather_decoded, shard_range


@pytest.mark.parametrize("n,world", [(1663370, 8), (1663370, 1), (10000, 2), (4096, 4), (1, 8), (100000, 3),
                                     (25557032, 8)])
def test_partition_covers_every_parameter_once(n, world):
    batch = 4096
    shards = [shard_range(n, batch, r, world) for r in range(world)]
    n_ct = -(-n // batch)
    assert sum(s.ct_count for s in shards) == n_ct
    assert sum(s.param_count for s in shards) == n
    pos_ct = pos_p = 0
    for s in shards:
        if s.ct_count:
            assert (s.ct_first, s.param_first) == (pos_ct, pos_p)
            assert s.param_first == s.ct_first * batch          # shards start on ciphertext boundaries
            assert -(-s.param_count // batch) == s.ct_count
        pos_ct += s.ct_count
        pos_p += s.param_count
    assert max(s.ct_count for s in shards) == shards[0].ct_per_rank


def test_bad_arguments():
    with pytest.raises(ValueError):
        shard_range(0, 4096, 0, 1)
    with pytest.raises(ValueError):
        shard_range(10, 4096, 2, 2)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, n, batch, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sh = shard_range(n, batch, rank, world)
        full = np.random.default_rng(7).normal(size=n)               # what the unsharded decrypt would return
        local = torch.from_numpy(full[sh.param_first:sh.param_first + sh.param_count].copy())
        got = gather_decoded(local, n, batch)
        # device-timing convention of bench.py: max over ranks
        t = torch.tensor([float(rank + 1)], dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ok = bool(np.array_equal(got.numpy(), full)) and t.item() == world
        with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
            f.write("ok" if ok else "mismatch")
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 100000), (3, 3 * 4096 + 5), (2, 100)])
def test_gather_decoded_gloo(world, n, tmp_path):
    mp.spawn(_worker, args=(world, _free_port(), n, 4096, str(tmp_path)), nprocs=world, join=True)
    assert [(tmp_path / f"rank{r}.txt").read_text() for r in range(world)] == ["ok"] * world


def _peer_worker(rank, world, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        import SHELFI_FHE
        from SHELFI_FHE.sharding import PeerGather
        h = SHELFI_FHE.CKKS(device=0)
        try:
            PeerGather(h, 100000, 4096)
            verdict = "constructed"
        except RuntimeError as e:            # no device here: the allocation fails on every rank, and all learn of it
            verdict = "raised:" + str(e)
        t = torch.tensor([1.0])
        dist.all_reduce(t)                   # the process group is still usable and nobody is stuck in a collective
        with open(os.path.join(out_dir, f"rank{rank}.txt"), "w") as f:
            f.write(verdict + f"|{int(t.item())}")
    finally:
        dist.destroy_process_group()


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-device failure mode")
def test_peer_gather_fails_on_every_rank_together(tmp_path):
    """PeerGather's constructor is collective: a local failure is shared before anyone raises, so no rank is left
    waiting in all_gather_object (here every rank fails: there is no device)."""
    world = 2
    mp.spawn(_peer_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    for r in range(world):
        text = (tmp_path / f"rank{r}.txt").read_text()
        assert text.startswith("raised:") and "allocation failed on rank(s) [0, 1]" in text and text.endswith("|2"), text
