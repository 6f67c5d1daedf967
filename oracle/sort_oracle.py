"""TEST INFRASTRUCTURE ONLY: numpy restatement of the reference's archetype sort
(src/mw/device/sort_archetype.cpp:960-1392). That file is an orphan of the snapshot (no build target
compiles it, it needs CUB and a GPU) and the reference holds no test or golden vector for it.
PARITY PINNED on the third-party algorithm it adapts: cub::DeviceRadixSort::SortPairs (the CUB of CUDA 12.9;
the reference's vendored external/cub @ 09686d2 does not compile against this toolkit), run on a B200 by
oracle/_ref/cub_sort_ref; keys -> (sorted keys, permutation) fixtures in tests/golden/sort_cub.npz
(tests/golden/make_sort_golden.py), checked in tests/test_table_sort.py. What is restated:
  * sortSetup / addToGraph :1290-1299: passes = ceil(bits(num_worlds + 1) / 8) for a WorldID
    sort, else 4;
  * histogram + binScan + onesweep per pass :1083-1214: a least-significant-digit radix sort of
    (key, row index) pairs over `passes` 8-bit digits, each pass stable = a stable sort by the
    low 8 * passes bits of the key (`lsd_reference` below walks the passes literally, the fast
    path uses one stable argsort; tests/test_table_sort.py checks they agree);
  * resizeTable :1216-1221: new row count = exclusive prefix of digit 255 in the last pass;
  * stageColumn / rearrangeColumn :1228-1241,1265-1288: every column gathered by the final indices;
  * rearrangeEntities :1243-1263: rows with WorldID != -1 remap their entity to the new row.
Imported by tests/ and bench_sort.py's cpu_baseline leg only."""
import numpy as np


def num_passes(num_worlds):
    if num_worlds <= 0:
        return 4
    return (int(num_worlds + 1).bit_length() + 7) // 8


def lsd_reference(keys, passes):
    """The passes one by one (counting sort per digit, as the Onesweep nodes do)."""
    keys = np.asarray(keys).view(np.uint32)
    idx = np.arange(len(keys), dtype=np.int64)
    k = keys.copy()
    for p in range(passes):
        d = (k >> np.uint32(8 * p)) & np.uint32(255)
        order = np.argsort(d, kind="stable")
        k, idx = k[order], idx[order]
    return idx


def sort_table(columns, key_column, num_worlds, entity_row=None):
    """Returns (sorted columns, new_row_count, entity_row)."""
    keys = np.ascontiguousarray(columns[key_column]).view(np.uint32).reshape(-1)
    passes = num_passes(num_worlds)
    mask = np.uint32(0xFFFFFFFF if passes == 4 else (1 << (8 * passes)) - 1)
    perm = np.argsort(keys & mask, kind="stable")
    out = [np.ascontiguousarray(c)[perm] for c in columns]
    new_rows = len(keys)
    if num_worlds > 0:
        top = (keys >> np.uint32(8 * (passes - 1))) & np.uint32(255)
        new_rows = int((top < 255).sum())
    if entity_row is not None:
        entity_row = entity_row.copy()
        ids = out[0][:new_rows, 1]
        ok = (ids >= 0) & (ids < len(entity_row))
        entity_row[ids[ok]] = np.arange(new_rows, dtype=np.int32)[ok]
    return out, new_rows, entity_row
