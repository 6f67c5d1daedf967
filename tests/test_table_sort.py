"""Global archetype sort / compaction (src/mw/device/sort_archetype.cpp:960-1392):
the numpy oracle against a literal pass-by-pass LSD walk (CPU), and the CUDA operator through the
mdr_sort_* C ABI against the oracle, bit-exact (keys, permutation of every column, new row count,
entity -> row map). Parity is restated from the specification (oracle/sort_oracle.py: unpinned)."""
import numpy as np
import pytest

from gpu_ecs_madrona_b200 import table_sort
from oracle import sort_oracle

PHYS_COLS = [8, 4, 12, 16, 12, 24, 4, 4, 28, 28, 24, 12, 12, 4]   # Entity, WorldID, phys::Cols


@pytest.mark.parametrize("num_worlds", [1, 3, 254, 255, 256, 65534, 65535, 70000, 1 << 20])
def test_pass_count_matches_spec(num_worlds):
    bits = 32 - (32 - int(num_worlds + 1).bit_length())      # 32 - clz(num_worlds + 1)
    assert sort_oracle.num_passes(num_worlds) == -(-bits // 8)


@pytest.mark.parametrize("num_worlds,rows", [(5, 7), (300, 11), (70000, 1)])
def test_oracle_equals_literal_lsd_passes(num_worlds, rows):
    cols = table_sort.churned_table(num_worlds, rows, [8, 4, 12], seed=num_worlds)
    passes = sort_oracle.num_passes(num_worlds)
    perm = sort_oracle.lsd_reference(cols[1], passes)
    out, new_rows, _ = sort_oracle.sort_table(cols, 1, num_worlds)
    for c, o in zip(cols, out):
        assert np.array_equal(c[perm], o)
    w = out[1]
    assert (w[:new_rows] >= 0).all() and (w[new_rows:] == -1).all()
    assert (np.diff(w[:new_rows]) >= 0).all()


def _device_columns(cols, torch):
    return [torch.from_numpy(np.ascontiguousarray(c)).cuda() for c in cols]


@pytest.mark.gpu
@pytest.mark.parametrize("num_worlds,rows_per_world,in_place", [
    (1, 1, False), (3, 5, False), (64, 65, False), (64, 65, True), (300, 33, False),
    (4096, 17, True), (70000, 3, False), (1000, 257, False)])
def test_cuda_world_sort_matches_oracle(cuda_lib, num_worlds, rows_per_world, in_place):
    import torch
    cols = table_sort.churned_table(num_worlds, rows_per_world, PHYS_COLS, seed=num_worlds + rows_per_world)
    n = len(cols[1])
    ent_row = np.full(n, -7, np.int32)
    want, want_rows, want_map = sort_oracle.sort_table(cols, 1, num_worlds, ent_row)
    d_in = _device_columns(cols, torch)
    d_out = d_in if in_place else [torch.empty_like(c) for c in d_in]
    d_map = torch.from_numpy(ent_row).cuda()
    s = table_sort.TableSort(n, PHYS_COLS, 1, num_worlds)
    assert s.passes == sort_oracle.num_passes(num_worlds)
    got_rows = s.run(d_in, d_out, n, d_map)
    assert got_rows == want_rows
    for c, (g, w) in enumerate(zip(d_out, want)):
        assert np.array_equal(g.cpu().numpy(), w), f"column {c}"
    assert np.array_equal(d_map.cpu().numpy(), want_map)
    # idempotence: sorting the compacted table again changes nothing
    d2 = [torch.empty_like(c) for c in d_out]
    assert s.run(d_out, d2, got_rows) == got_rows
    for c, (a, b) in enumerate(zip(d_out, d2)):
        assert torch.equal(a[:got_rows], b[:got_rows]), f"column {c} after re-sort"
    s.close()


@pytest.mark.gpu
def test_cuda_generic_key_sort_and_edge_cases(cuda_lib):
    """A full 32-bit key (4 passes, nothing dropped), ragged tile sizes, empty table, bad arguments."""
    import torch
    rng = np.random.default_rng(5)
    for n in (1, 31, 4095, 4096, 4097, 100003):
        keys = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
        keys[rng.random(n) < 0.3] = keys[0]           # runs of equal keys: stability shows
        ent = np.stack([np.zeros(n, np.int32), np.arange(n, dtype=np.int32)], 1)
        pay = rng.integers(0, 1 << 32, (n, 3), dtype=np.uint64).astype(np.uint32)
        cols = [ent, keys, pay]
        want, want_rows, _ = sort_oracle.sort_table(cols, 1, 0)
        s = table_sort.TableSort(n, [8, 4, 12], 1, 0)
        d_in = _device_columns(cols, torch)
        d_out = [torch.empty_like(c) for c in d_in]
        assert s.run(d_in, d_out, n) == want_rows == n
        for g, w in zip(d_out, want):
            assert np.array_equal(g.cpu().numpy(), w), n
        assert s.run(d_in, d_out, 0) == 0
        from gpu_ecs_madrona_b200 import phys
        with pytest.raises(phys.MadronaError):
            s.run(d_in, d_out, n + 1)
        s.close()
    with pytest.raises(phys.MadronaError):
        table_sort.TableSort(16, [8, 6], 1, 0)          # key column must be 4 bytes


@pytest.mark.gpu
def test_cuda_sort_full_size_properties(cuda_lib):
    """BASELINE.json's collisions-64 table as one shared table (16 384 worlds x 65 rows): sortedness,
    every surviving entity exactly once, payload still attached to its entity, map consistent."""
    import torch
    W, R = 16384, 65
    cols = table_sort.churned_table(W, R, PHYS_COLS, seed=1)
    n = W * R
    d_in = _device_columns(cols, torch)
    d_out = [torch.empty_like(c) for c in d_in]
    d_map = torch.full((n,), -1, dtype=torch.int32, device="cuda")
    s = table_sort.TableSort(n, PHYS_COLS, 1, W)
    rows = s.run(d_in, d_out, n, d_map)
    alive = int((cols[1] >= 0).sum())
    assert rows == alive
    w = d_out[1][:rows]
    assert bool((w[1:] >= w[:-1]).all()) and int(w.min()) >= 0
    assert bool((d_out[1][rows:] == -1).all())
    ids = d_out[0][:rows, 1].long()
    assert int(torch.unique(ids).numel()) == rows
    assert bool((d_map[ids] == torch.arange(rows, device="cuda", dtype=torch.int32)).all())
    first_payload = d_out[2][:rows, 0].cpu().numpy().view(np.uint32)
    assert np.array_equal(first_payload, (ids.cpu().numpy().astype(np.uint32) * np.uint32(2654435761)))
    s.close()


# ---- pinned on cub::DeviceRadixSort::SortPairs (tests/golden/sort_cub.npz, generated on a B200 by
# tests/golden/make_sort_golden.py through oracle/_ref/cub_sort_ref): the third-party algorithm the reference's
# sort nodes adapt (src/mw/device/sort_archetype.cpp:960-1392) ------------------------------------------------
import hashlib
import os

from tests.golden import make_sort_golden as msg

CUB = np.load(os.path.join(os.path.dirname(__file__), "golden", "sort_cub.npz"))


def _sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8)


def _check_against_cub(name, keys, sorted_keys, perm):
    if name + "_perm" in CUB:
        assert np.array_equal(keys, CUB[name + "_keys"]), "fixture keys"
        assert np.array_equal(sorted_keys, CUB[name + "_sorted"]), name
        assert np.array_equal(perm, CUB[name + "_perm"]), name
    else:
        assert np.array_equal(_sha(sorted_keys), CUB[name + "_sorted_sha"]), name
        assert np.array_equal(_sha(np.asarray(perm, np.int32)), CUB[name + "_perm_sha"]), name


def test_numpy_oracle_pinned_on_cub():
    """oracle/sort_oracle.py (stable sort by the low end_bit bits) == CUB, keys and permutation, every fixture."""
    for name, (keys, end_bit, _) in msg.cases().items():
        assert int(CUB[name + "_end_bit"]) == end_bit
        mask = np.uint32(0xFFFFFFFF if end_bit == 32 else (1 << end_bit) - 1)
        perm = np.argsort(keys & mask, kind="stable").astype(np.int32)
        _check_against_cub(name, keys, keys[perm], perm)
        if len(keys) <= 5000:
            assert np.array_equal(sort_oracle.lsd_reference(keys, end_bit // 8), perm)


@pytest.mark.gpu
def test_cuda_global_sort_pinned_on_cub(cuda_lib):
    """mdr_sort_run on the fixture keys: sorted keys and the permutation (an iota payload column) == CUB."""
    import torch
    for name, (keys, end_bit, _) in msg.cases().items():
        n = len(keys)
        if n == 0:
            continue
        num_worlds = 0
        if name.startswith("world_"):
            num_worlds = int(name.split("_")[1].split("x")[0])
            assert 8 * sort_oracle.num_passes(num_worlds) == end_bit
        ent = np.stack([np.zeros(n, np.int32), np.arange(n, dtype=np.int32)], 1)
        iota = np.arange(n, dtype=np.int32)
        cols = [ent, keys.view(np.int32), iota]
        s = table_sort.TableSort(n, [8, 4, 4], 1, num_worlds)
        d_in = _device_columns(cols, torch)
        d_out = [torch.empty_like(c) for c in d_in]
        s.run(d_in, d_out, n)
        _check_against_cub(name, keys, d_out[1].cpu().numpy().view(np.uint32), d_out[2].cpu().numpy())
        s.close()


@pytest.mark.gpu
def test_megakernel_sort_node_pinned_on_cub(cuda_lib):
    """The per-world SortArchetypeNode inside the task-graph megakernel (csrc ecs sortArchetype): one world per
    fixture key set, sorted keys and permutation == CUB."""
    from gpu_ecs_madrona_b200.ecs import EcsRegistry
    names = [n for n, (k, e, full) in msg.cases().items() if full and e == 32 and 0 < len(k) <= 4097]
    sets = [msg.cases()[n][0] for n in names]
    cap = max(len(k) for k in sets)
    r = EcsRegistry(len(sets))
    key, pay = r.register_component(4), r.register_component(4)
    arch = r.register_archetype([key, pay], cap)
    r.sort_archetype(arch, key, remap_entities=True)
    r.build()
    keys = np.zeros((len(sets), cap), np.uint32)
    iota = np.tile(np.arange(cap, dtype=np.int32), (len(sets), 1))
    from oracle import eorc
    ops = []
    for w, k in enumerate(sets):
        keys[w, :len(k)] = k
        o = np.zeros(len(k), eorc.OP_DTYPE)
        o["op"], o["a"] = eorc.OP_CREATE, arch
        ops.append(o)
    r.script(ops)                       # len(k) rows in world w (immediate-mode creates)
    r.write_column_all(arch, 1, keys)
    r.write_column_all(arch, 2, iota)
    r.run(1)
    got_k = r.read_column_all(arch, 1).view(np.uint32)[..., 0]
    got_p = r.read_column_all(arch, 2).view(np.int32)[..., 0]
    for w, (name, k) in enumerate(zip(names, sets)):
        _check_against_cub(name, k, got_k[w, :len(k)], got_p[w, :len(k)])
    # entities followed their rows: the id map points at the sorted positions
    ents = r.read_column_all(arch, 0).view(np.int32)
    for w, k in enumerate(sets[:3]):
        a, b, gen, st = r.idmap(w)
        ids = ents[w, :len(k), 1]
        assert np.array_equal(b[ids], np.arange(len(k)))
    r.close()
