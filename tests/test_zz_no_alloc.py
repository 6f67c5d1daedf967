"""The allocation-free entry point `mdr_sort_pairs_static` (include/madrona_b200.h; no counterpart in the reference):
the key sort with its scratch in a fixed device array of the library.
Checked through the C ABI against the existing entry point `mdr_sort_run` on the same keys, bit for bit (sorted
keys and permutation), on two calls in a row that share the scratch array (the second one smaller and with fewer
digit passes, so stale contents of the first would show). Named test_zz_* so that pytest collects it last."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _lib():
    from gpu_ecs_madrona_b200 import phys
    lib = phys.load_library()
    lib.mdr_sort_pairs_static.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]
    lib.mdr_sort_pairs_static.restype = C.c_int32
    lib.mdr_sort_static_max_rows.restype = C.c_int32
    lib.mdr_sort_static_scratch_bytes.restype = C.c_int64
    return lib


def _existing(keys, num_worlds):
    """mdr_sort_run (scratch from mdr_sort_create) on [Entity, key, iota]: sorted keys and permutation."""
    import torch
    from gpu_ecs_madrona_b200 import table_sort
    n = len(keys)
    ent = np.stack([np.zeros(n, np.int32), np.arange(n, dtype=np.int32)], 1)
    cols = [ent, keys.view(np.int32), np.arange(n, dtype=np.int32)]
    s = table_sort.TableSort(n, [8, 4, 4], 1, num_worlds)
    d_in = [torch.from_numpy(np.ascontiguousarray(c)).cuda() for c in cols]
    d_out = [torch.empty_like(c) for c in d_in]
    s.run(d_in, d_out, n)
    out = d_out[1].cpu().numpy().view(np.uint32), d_out[2].cpu().numpy()
    s.close()
    return out


def _static(lib, keys, key_bits):
    import torch
    n = len(keys)
    d_in = torch.from_numpy(keys.view(np.int32).copy()).cuda()
    d_keys = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    d_perm = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    torch.cuda.synchronize()
    rc = lib.mdr_sort_pairs_static(0, d_in.data_ptr(), d_keys.data_ptr(), d_perm.data_ptr(), n, key_bits, None)
    assert rc == 0, rc
    torch.cuda.synchronize()            # the call only enqueues
    return d_keys.cpu().numpy().view(np.uint32), d_perm.cpu().numpy()


def test_static_sort_matches_existing_entry_point_twice(cuda_lib):
    lib = _lib()
    cap = lib.mdr_sort_static_max_rows()
    assert cap == 65536 and 0 < lib.mdr_sort_static_scratch_bytes() < (1 << 20)
    rng = np.random.default_rng(11)
    # call 1: the largest table it takes, full 32-bit keys with runs of equal keys (stability shows), 4 passes
    k1 = rng.integers(0, 1 << 32, cap, dtype=np.uint64).astype(np.uint32)
    k1[rng.random(cap) < 0.3] = k1[0]
    # call 2, right after, on the same scratch: fewer rows, WorldID keys of 300 worlds with dead rows (-1), 2 passes
    k2 = np.repeat(np.arange(300, dtype=np.int32), 33)
    rng.shuffle(k2)
    k2[rng.random(len(k2)) < 0.1] = -1
    k2 = k2.view(np.uint32).copy()
    for keys, bits, num_worlds in ((k1, 32, 0), (k2, 16, 300), (k1[:4097], 32, 0)):
        want_k, want_p = _existing(keys, num_worlds)
        got_k, got_p = _static(lib, keys, bits)
        assert np.array_equal(got_k, want_k)
        assert np.array_equal(got_p, want_p)


def test_static_sort_rejects_what_it_cannot_hold(cuda_lib):
    import torch
    lib = _lib()
    n = lib.mdr_sort_static_max_rows() + 1
    a = torch.zeros(n, dtype=torch.int32, device="cuda")
    b, c = torch.zeros_like(a), torch.zeros_like(a)
    assert lib.mdr_sort_pairs_static(0, a.data_ptr(), b.data_ptr(), c.data_ptr(), n, 32, None) == -3     # MDR_ECAPACITY
    assert lib.mdr_sort_pairs_static(0, a.data_ptr(), a.data_ptr(), c.data_ptr(), 16, 32, None) == -1    # in place
    assert lib.mdr_sort_pairs_static(0, a.data_ptr(), b.data_ptr(), c.data_ptr(), 16, 33, None) == -1
    assert lib.mdr_sort_pairs_static(0, a.data_ptr(), b.data_ptr(), c.data_ptr(), 0, 32, None) == 0


def test_tables_that_do_not_fit_are_an_error_code_not_an_abort(cuda_lib):
    """cudaMalloc failures used to abort() the host process. A batch that cannot fit the device answers
    MDR_ECAPACITY from mdr_phys_create / mdr_ecs_build (mdr_sort_create likewise), frees what it took, and the process goes on."""
    from gpu_ecs_madrona_b200 import phys, scenes
    from gpu_ecs_madrona_b200.ecs import EcsRegistry
    sc = scenes.collisions64(4)
    sc.num_worlds = 400_000_000                    # the first column alone is 217 GB: fails at once
    with pytest.raises(phys.MadronaError) as ei:
        phys.PhysExecutor(sc)
    assert ei.value.code == phys.MDR_ECAPACITY
    r = EcsRegistry(1 << 20)
    c = r.register_component(1024)
    r.register_archetype([c], 1 << 15)             # 2^20 worlds x 32768 rows: the Entity column alone is 275 GB
    with pytest.raises(phys.MadronaError) as ei:
        r.build()
    assert ei.value.code == phys.MDR_ECAPACITY
    r.close()
    # and the device is still usable
    ex = phys.PhysExecutor(scenes.collisions64(8))
    ex.run(2)
    assert np.isfinite(ex.state()[0]).all()
    ex.close()
