"""Helpers for the -m gpu tests: torch is used only as the owner of device memory and streams."""
import ctypes as C

import numpy as np
import torch


def dev_u64(a: np.ndarray) -> torch.Tensor:
    return torch.from_numpy(np.ascontiguousarray(a, np.uint64).view(np.int64)).cuda()


def host_u64(t: torch.Tensor) -> np.ndarray:
    return t.cpu().numpy().view(np.uint64)


def ptr(t: torch.Tensor) -> C.c_void_p:
    return C.c_void_p(t.data_ptr())


def stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def check(h, rc):
    assert rc == 0, h._err()
