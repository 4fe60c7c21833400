"""-m gpu: decrypt-gather (shelfi_decrypt_gather_device, SHELFI_FHE.sharding.PeerGather): the decode kernel stores every
value to several destinations.  On one GPU the destinations are local buffers and must equal shelfi_decrypt_device's
output at the slice offset; with two or more GPUs visible the torchrun check compares against decrypt + NCCL all-gather."""
import ctypes as C
import os
import subprocess
import sys
from pathlib import Path

import numpy as np
import pytest

torch = pytest.importorskip("torch")
from gpu_util import check, ptr, stream  # noqa: E402

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


@pytest.fixture(scope="module")
def handles(shelfi, gpu_ckks, oracles, deep_oracles):
    out = {"pybind": (gpu_ckks["pybind"], oracles["pybind"])}
    for name in ("n16k_l4", "n32k_l8", "n8k_l3"):     # decode tail of 2 and 4 blocks, general CRT in one tile
        o = deep_oracles[name]
        h = shelfi.CKKS("ckks", o.batchSize, o.scaleFactorBits, o.cryptodir, device=0)
        h.loadCryptoParams()
        assert h.load_status == 0
        out[name] = (h, o)
    return out


@pytest.mark.parametrize("which", ["pybind", "n16k_l4", "n32k_l8", "n8k_l3"])
def test_every_destination_gets_the_decoded_slice(shelfi, handles, which):
    h, o = handles[which]
    lib, batch = shelfi.lib(), o.batchSize
    words = lib.shelfi_ciphertext_words(h.handle)
    rng = np.random.default_rng(5)
    n_ct, n = 3, 2 * batch + 77                     # last ciphertext short
    x = rng.normal(0, 0.05, n)
    dx = torch.from_numpy(x).cuda()
    dct = torch.empty(n_ct * words, dtype=torch.int64, device="cuda")
    check(h, lib.shelfi_encrypt_device(h.handle, ptr(dx), n, 0, ptr(dct), stream()))
    sf = lib.shelfi_scaling_factor(h.handle)
    ref = torch.empty(n, dtype=torch.float64, device="cuda")
    check(h, lib.shelfi_decrypt_device(h.handle, ptr(dct), n_ct, sf, ptr(ref), n, stream()))
    offset, total = 1234, 1234 + n + 99
    outs = [torch.full((total,), -7.0, dtype=torch.float64, device="cuda") for _ in range(3)]
    arr = (C.c_void_p * 3)(*[t.data_ptr() for t in outs])
    check(h, lib.shelfi_decrypt_gather_device(h.handle, ptr(dct), n_ct, sf, arr, 3, offset, n, stream()))
    torch.cuda.synchronize()
    for t in outs:
        assert torch.equal(t[offset:offset + n], ref)
        assert bool((t[:offset] == -7.0).all()) and bool((t[offset + n:] == -7.0).all())   # nothing outside the slice
    assert np.abs(ref.cpu().numpy() - x).max() < 1e-6
    # one destination = plain decrypt at an offset
    one = torch.zeros(total, dtype=torch.float64, device="cuda")
    check(h, lib.shelfi_decrypt_gather_device(h.handle, ptr(dct), n_ct, sf, (C.c_void_p * 1)(one.data_ptr()), 1, offset, n, stream()))
    assert torch.equal(one[offset:offset + n], ref)


def test_argument_errors(shelfi, handles):
    h, o = handles["pybind"]
    lib = shelfi.lib()
    buf = torch.zeros(8192 * 4 + 4096, dtype=torch.int64, device="cuda")
    out = torch.zeros(4096, dtype=torch.float64, device="cuda")
    nine = (C.c_void_p * 9)(*[out.data_ptr()] * 9)
    assert lib.shelfi_decrypt_gather_device(h.handle, ptr(buf), 1, 1.0, nine, 9, 0, 4096, stream()) != 0
    assert lib.shelfi_decrypt_gather_device(h.handle, ptr(buf), 1, 1.0, nine, 0, 0, 4096, stream()) != 0
    hole = (C.c_void_p * 2)(out.data_ptr(), None)
    assert lib.shelfi_decrypt_gather_device(h.handle, ptr(buf), 1, 1.0, hole, 2, 0, 4096, stream()) != 0
    assert lib.shelfi_decrypt_gather_device(h.handle, ptr(buf), 1, 1.0, nine, 2, 0, 4097, stream()) != 0
    assert "destination" in h._err() or "inconsistent" in h._err()


def test_peer_buffer_lifecycle(shelfi, handles):
    h, _ = handles["pybind"]
    lib = shelfi.lib()
    p, handle = C.c_void_p(), C.create_string_buffer(64)
    check(h, lib.shelfi_peer_buffer_create(h.handle, 1 << 20, C.byref(p), handle))
    assert p.value and any(handle.raw)
    check(h, lib.shelfi_peer_buffer_destroy(h.handle, p))
    assert lib.shelfi_peer_buffer_create(h.handle, 0, C.byref(p), handle) != 0


def test_stream_barrier_between_two_ranks_on_one_gpu(shelfi, handles):
    """Two 'ranks' = two streams of one process: each barrier kernel must wait for the other's flag."""
    h, _ = handles["pybind"]
    lib = shelfi.lib()
    flags = [torch.zeros(16, dtype=torch.int64, device="cuda") for _ in range(2)]
    arr = (C.c_void_p * 2)(*[f.data_ptr() for f in flags])
    s = [torch.cuda.Stream() for _ in range(2)]
    marks = torch.zeros(2, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()
    for epoch in (1, 2, 3):
        with torch.cuda.stream(s[0]):
            torch.cuda._sleep(20_000_000)                      # rank 0 arrives ~10 ms late
            marks[0] = epoch                                   # ... having written this before its barrier
            check(h, lib.shelfi_peer_barrier_device(h.handle, arr, 2, 0, epoch, C.c_void_p(s[0].cuda_stream)))
        with torch.cuda.stream(s[1]):
            check(h, lib.shelfi_peer_barrier_device(h.handle, arr, 2, 1, epoch, C.c_void_p(s[1].cuda_stream)))
            seen = marks[0].clone()                            # after the barrier rank 1 must see rank 0's write
        torch.cuda.synchronize()
        assert int(seen.item()) == epoch
        assert flags[0][:2].tolist() == [epoch, epoch] and flags[1][:2].tolist() == [epoch, epoch]
        assert int(flags[0][8].item()) == 0 and int(flags[1][8].item()) == 0
    assert lib.shelfi_peer_barrier_device(h.handle, arr, 9, 0, 4, None) != 0
    assert lib.shelfi_peer_barrier_device(h.handle, arr, 2, 2, 4, None) != 0
    assert lib.shelfi_peer_barrier_device(h.handle, arr, 2, 0, 0, None) != 0


def test_barrier_gives_up_after_the_timeout(shelfi, handles, monkeypatch):
    h, _ = handles["pybind"]
    lib = shelfi.lib()
    monkeypatch.setenv("SHELFI_PEER_TIMEOUT_MS", "50")
    flags = [torch.zeros(16, dtype=torch.int64, device="cuda") for _ in range(2)]
    arr = (C.c_void_p * 2)(*[f.data_ptr() for f in flags])
    check(h, lib.shelfi_peer_barrier_device(h.handle, arr, 2, 0, 5, stream()))   # rank 1 never comes
    torch.cuda.synchronize()
    assert int(flags[0][8].item()) == 5 and int(flags[1][0].item()) == 5


def test_guarded_barrier_poisons_the_missing_slice(shelfi, handles, monkeypatch):
    """The barrier after a decrypt-gather: a peer that never arrives leaves a NaN at the start of ITS slice in this
    GPU's copy (and the status word), nothing else is touched; a barrier that completes leaves the values alone."""
    h, _ = handles["pybind"]
    lib = shelfi.lib()
    monkeypatch.setenv("SHELFI_PEER_TIMEOUT_MS", "50")
    flags = [torch.zeros(16, dtype=torch.int64, device="cuda") for _ in range(2)]
    arr = (C.c_void_p * 2)(*[f.data_ptr() for f in flags])
    block = 1000
    values = torch.ones(2 * block, dtype=torch.float64, device="cuda")
    check(h, lib.shelfi_peer_barrier_guard_device(h.handle, arr, 2, 0, 7, C.c_void_p(values.data_ptr()), block, stream()))
    torch.cuda.synchronize()
    got = values.cpu().numpy()
    assert int(flags[0][8].item()) == 7
    assert np.isnan(got[block]) and not np.isnan(got[0])          # rank 1 owed the second slice
    assert np.all(np.delete(got, block) == 1.0)
    # a one-rank barrier always completes: no poison
    solo = torch.zeros(16, dtype=torch.int64, device="cuda")
    arr1 = (C.c_void_p * 1)(solo.data_ptr())
    v1 = torch.ones(block, dtype=torch.float64, device="cuda")
    check(h, lib.shelfi_peer_barrier_guard_device(h.handle, arr1, 1, 0, 1, C.c_void_p(v1.data_ptr()), block, stream()))
    torch.cuda.synchronize()
    assert bool((v1 == 1.0).all()) and int(solo[8].item()) == 0
    assert lib.shelfi_peer_barrier_guard_device(h.handle, arr1, 1, 0, 2, None, block, stream()) != 0


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs on the box")
def test_two_ranks_equal_nccl_all_gather():
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29537", str(ROOT / "tools" / "peer_gather_check.py"),
                        "--params", "100000", "--iters", "3"], capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert '"identical_on_every_rank": true' in r.stdout
