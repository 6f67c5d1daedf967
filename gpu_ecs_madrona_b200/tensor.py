"""Exported columns as the reference hands them to a trainer (include/madrona/python.hpp:21-73,
src/python/bindings.cpp:78-128): `madrona_python.Tensor` = (pointer, element type, dims, gpu id) with
`.to_torch()` and the DLPack protocol, `madrona_python.CudaSync.wait(stream)` for the stream-ordered
hand-off. Zero copy: the torch tensor aliases the executor's HBM table."""
import numpy as np

_ELEMENT = None


def module():
    """The compiled `madrona_python` extension (gpu_ecs_madrona_b200/python/bindings.cpp)."""
    try:
        from . import madrona_python
    except ImportError as e:
        raise ImportError("gpu_ecs_madrona_b200/madrona_python*.so is missing: build it with "
                          "`python -c 'import __graft_entry__ as g; g.build()'`") from e
    return madrona_python


def element_type(dtype):
    mp = module()
    table = {np.dtype(np.uint8): mp.ElementType.UInt8, np.dtype(np.int8): mp.ElementType.Int8,
             np.dtype(np.int16): mp.ElementType.Int16, np.dtype(np.int32): mp.ElementType.Int32,
             np.dtype(np.int64): mp.ElementType.Int64, np.dtype(np.float16): mp.ElementType.Float16,
             np.dtype(np.float32): mp.ElementType.Float32}
    return table[np.dtype(dtype)]


def as_tensor(ptr, shape, dtype, gpu_id):
    """madrona_python.Tensor over device memory at `ptr` (gpu_id -1: host memory)."""
    return module().Tensor.from_pointer(int(ptr), element_type(dtype), [int(x) for x in shape], int(gpu_id))


def as_torch(ptr, shape, dtype, gpu_id, owner=None):
    """Zero-copy torch view [shape] of device memory; `owner` is kept alive by the tensor."""
    if ptr is None or int(ptr) == 0:
        raise ValueError("null device pointer")
    if 0 in tuple(shape):       # a staged export of an archetype without rows
        import torch
        return torch.empty(tuple(shape), dtype=getattr(torch, np.dtype(dtype).name), device=f"cuda:{gpu_id}")
    t = as_tensor(ptr, shape, dtype, gpu_id).to_torch()
    t._mdr_owner = owner
    return t


def cuda_sync(done_event):
    return module().CudaSync(int(done_event))
