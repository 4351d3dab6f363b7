"""Helpers shared by the -m gpu parity tests (torch is only the device-memory allocator)."""
import numpy as np
import torch

from hpc_b200 import capi

_ctx = None


def ctx():
    global _ctx
    if _ctx is None:
        _ctx = capi.Context(0)
    return _ctx


def dev(x):
    """numpy -> device tensor with the same bytes (uint16 patterns go through int16)."""
    x = np.ascontiguousarray(x)
    if x.dtype == np.uint16:
        return torch.from_numpy(x.view(np.int16)).cuda()
    return torch.from_numpy(x).cuda()


def host(t, dtype=None):
    a = t.detach().cpu().numpy()
    if dtype is not None:
        a = a.view(dtype)
    return a


def bits32(x):
    return np.ascontiguousarray(x).view(np.uint32)


def sync():
    torch.cuda.synchronize()
