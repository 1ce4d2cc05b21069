"""Zero-copy views of context-owned device buffers as torch tensors (plumbing for torch.distributed
collectives and device-side checks; no computation)."""
import numpy as np
import torch


class _CudaView:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 3, "strides": None}


def as_torch(ptr, shape, dtype=torch.float64, device="cuda"):
    typestr = {torch.float64: "<f8", torch.int32: "<i4", torch.int64: "<i8", torch.uint8: "|u1"}[dtype]
    return torch.as_tensor(_CudaView(ptr, shape, typestr), device=device)


def buffer(ctx, which, dtype=torch.float64):
    ptr, nbytes = ctx.device_buffer(which)
    n = nbytes // torch.tensor([], dtype=dtype).element_size()
    return as_torch(ptr, (n,), dtype)
