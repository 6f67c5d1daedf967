#!/usr/bin/env python
"""Generates tests/golden/sort_cub.npz ON A GPU BOX: keys -> (sorted keys, permutation) from
cub::DeviceRadixSort::SortPairs (oracle/_ref/cub_sort_ref, built by `make -C oracle _ref/cub_sort_ref`), the
algorithm the reference's sort nodes adapt (src/mw/device/sort_archetype.cpp:960-1392).

  gpurun -- 'python tests/golden/make_sort_golden.py && cp tests/golden/sort_cub.npz gpurun_out/'

Small cases keep keys and permutation; large ones keep a SHA-256 of the permutation and of the sorted keys (the
keys are regenerated from the seed by cases() below, which the tests import)."""
import hashlib
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
HARNESS = os.path.join(ROOT, "oracle", "_ref", "cub_sort_ref")
OUT = os.path.join(HERE, "sort_cub.npz")


def num_passes(num_worlds):
    return 4 if num_worlds <= 0 else (int(num_worlds + 1).bit_length() + 7) // 8


def cases():
    """name -> (keys u32 [n], end_bit, keep_full). WorldID keys carry -1 (destroyed rows) as 0xffffffff."""
    out = {}
    rng = np.random.default_rng(20261105)
    for n in (0, 1, 2, 31, 32, 33, 255, 256, 257, 1000, 4095, 4096, 4097):
        k = rng.integers(0, 1 << 32, n, dtype=np.uint64).astype(np.uint32)
        if n > 3:
            k[rng.random(n) < 0.3] = k[0]             # runs of equal keys: stability shows
        out[f"u32_{n}"] = (k, 32, True)
    for W, R in ((3, 5), (64, 65), (300, 33), (1000, 7)):
        k = np.repeat(np.arange(W, dtype=np.int32), R)
        rng.shuffle(k)
        k[rng.random(len(k)) < 0.1] = -1
        out[f"world_{W}x{R}"] = (k.view(np.uint32).copy(), 8 * num_passes(W), True)
    # the per-world sort node: few hundred small keys per world, many equal
    for w in range(6):
        n = int(rng.integers(1, 400))
        out[f"node_w{w}"] = (rng.integers(0, 1 << int(rng.integers(3, 20)), n, dtype=np.uint64).astype(np.uint32), 32, True)
    # full size: BASELINE collisions-64 table as one shared table, and 2^20 random keys
    W, R = 16384, 65
    k = np.repeat(np.arange(W, dtype=np.int32), R)
    np.random.default_rng(7).shuffle(k)
    k[np.random.default_rng(8).random(len(k)) < 0.1] = -1
    out["world_16384x65"] = (k.view(np.uint32).copy(), 8 * num_passes(W), False)
    out["u32_1048576"] = (np.random.default_rng(9).integers(0, 1 << 32, 1 << 20, dtype=np.uint64).astype(np.uint32), 32, False)
    return out


def run_cub(keys, end_bit):
    with tempfile.TemporaryDirectory() as d:
        fin, fout = os.path.join(d, "in.bin"), os.path.join(d, "out.bin")
        with open(fin, "wb") as f:
            f.write(np.array([len(keys), end_bit], np.int32).tobytes())
            f.write(np.ascontiguousarray(keys, np.uint32).tobytes())
        subprocess.run([HARNESS, fin, fout], check=True)
        raw = np.fromfile(fout, np.uint32)
    n = len(keys)
    return raw[:n].copy(), raw[n:].view(np.int32).copy()


def digest(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), np.uint8).copy()


def main():
    data = {}
    for name, (keys, end_bit, full) in cases().items():
        sk, perm = run_cub(keys, end_bit)
        if full:
            data[name + "_keys"], data[name + "_sorted"], data[name + "_perm"] = keys, sk, perm
        else:
            data[name + "_sorted_sha"], data[name + "_perm_sha"] = digest(sk), digest(perm)
        data[name + "_end_bit"] = np.int32(end_bit)
        print(name, len(keys), "end_bit", end_bit)
    np.savez_compressed(OUT, **data)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
