"""Zero-copy views of exported device columns: the role of the reference's py::Tensor
(include/madrona/python.hpp:21-73, src/python/bindings.cpp:78-128): (pointer, dtype, dims, gpu id) -> a torch
tensor over the same memory. PyTorch is device-memory plumbing here, nothing is computed with it."""
import numpy as np


class DeviceArray:
    """Carries __cuda_array_interface__ for a raw device pointer (also consumable by cupy / numba)."""

    def __init__(self, ptr, shape, typestr, owner=None):
        self.owner = owner          # keeps the executor alive while views exist
        self.__cuda_array_interface__ = {"shape": tuple(int(x) for x in shape), "typestr": typestr,
                                         "data": (int(ptr), False), "version": 3, "strides": None}


def as_torch(ptr, shape, dtype, gpu_id, owner=None):
    import torch
    if ptr is None or int(ptr) == 0:
        raise ValueError("null device pointer")
    typestr = np.dtype(dtype).str
    if 0 in tuple(shape):
        return torch.empty(tuple(shape), dtype=getattr(torch, np.dtype(dtype).name), device=f"cuda:{gpu_id}")
    return torch.as_tensor(DeviceArray(ptr, shape, typestr, owner), device=f"cuda:{gpu_id}")
