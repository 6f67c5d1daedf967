"""the `madrona_python` module (Tensor, CudaSync) -- include/madrona/python.hpp:21-73,
src/python/bindings.cpp:78-128 -- over pybind11 (nanobind is not in this image)."""
import numpy as np
import pytest
import torch

from gpu_ecs_madrona_b200 import tensor as T


def test_module_surface():
    mp = T.module()
    assert {"Tensor", "CudaSync", "ElementType"} <= set(dir(mp))
    assert {"to_torch", "__dlpack__", "__dlpack_device__", "from_pointer"} <= set(dir(mp.Tensor))
    assert "wait" in dir(mp.CudaSync)
    assert [e for e in ("UInt8", "Int8", "Int16", "Int32", "Int64", "Float16", "Float32") if hasattr(mp.ElementType, e)] \
        == ["UInt8", "Int8", "Int16", "Int32", "Int64", "Float16", "Float32"]


@pytest.mark.parametrize("dtype", [np.uint8, np.int8, np.int16, np.int32, np.int64, np.float16, np.float32])
def test_host_tensor_is_zero_copy_both_ways(dtype):
    a = (np.arange(24) % 100).astype(dtype).reshape(2, 3, 4)
    t = T.as_tensor(a.ctypes.data, a.shape, dtype, -1)
    assert t.dims == [2, 3, 4] and not t.is_on_gpu and t.__dlpack_device__() == (1, 0)
    x = t.to_torch()
    assert tuple(x.shape) == (2, 3, 4) and x.numpy().dtype == np.dtype(dtype)
    assert np.array_equal(x.numpy(), a)
    x[1, 2, 3] = 7                      # a view, not a copy
    assert a[1, 2, 3] == 7
    y = torch.from_dlpack(t)            # the DLPack protocol itself
    assert y.data_ptr() == a.ctypes.data
    # Tensor(torch_tensor): the reference's import direction
    back = T.module().Tensor(torch.from_numpy(a))
    assert back.ptr == a.ctypes.data and back.dims == [2, 3, 4] and back.type == T.element_type(dtype)


def test_tensor_limits():
    mp = T.module()
    with pytest.raises(Exception):
        mp.Tensor.from_pointer(0, mp.ElementType.Float32, [1] * 17, -1)
    with pytest.raises(Exception):
        mp.Tensor(torch.zeros(4, dtype=torch.float64))         # not one of the reference's element types
    with pytest.raises(Exception):
        mp.Tensor(torch.zeros(4, 4)[:, ::2])                   # not contiguous


@pytest.mark.gpu
def test_exported_column_tensor_and_cuda_sync(cuda_lib):
    from gpu_ecs_madrona_b200 import phys, scenes
    sc = scenes.collisions64(16)
    ex = phys.PhysExecutor(sc)
    t = ex.tensor(phys.COL_POSITION)
    assert t.is_on_gpu and t.__dlpack_device__() == (2, 0) and t.ptr == ex.column_ptr(phys.COL_POSITION)
    pos = t.to_torch()
    assert pos.is_cuda and tuple(pos.shape) == (16, ex.row_stride(), 3)
    sync = ex.cuda_sync()
    ex.run_async(3)                      # enqueued on the executor's stream, no host sync
    sync.wait(torch.cuda.current_stream().cuda_stream)
    got = pos[:, :sc.bodies_per_world].clone()      # torch's stream, ordered after the steps
    torch.cuda.current_stream().synchronize()
    ex.sync()
    assert np.array_equal(got.cpu().numpy(), ex.read_f32(phys.COL_POSITION))
    ex.close()
