"""Generates tests/golden/static_map.npz from the UNMODIFIED reference StaticIntegerMap<128>
(include/madrona/hashmap.hpp, src/common/hashmap.cpp) through oracle/_ref/libmadrona_ecs_ref.so
(`eorc_static_map_build`, oracle/ecs_ref_harness.cpp). Run in the build container:
`python tests/golden/make_static_map_golden.py`. Key sets: the reference's own test (80 identity pairs,
tests/static_map.cpp:15-40), the column maps of the built-in worlds' archetypes, seeded random sets."""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(ROOT, "oracle", "_ref", "libmadrona_ecs_ref.so")


def distinct(rng, hi, n):
    """n distinct keys below hi (no permutation of the whole range)."""
    keys = np.unique(rng.randint(0, hi, size=4 * n + 8, dtype=np.int64))
    rng.shuffle(keys)
    return keys[:n].astype(np.uint32)


def cases():
    out = {"reference_test_80": (np.arange(80, dtype=np.uint32), np.arange(80, dtype=np.uint32)),
           "empty": (np.zeros(0, np.uint32), np.zeros(0, np.uint32)),
           "single": (np.array([7], np.uint32), np.array([3], np.uint32)),
           # Entity + 12 physics columns (phys::Cols, component ids as the physics world registers them)
           "rigid_body": (np.array([0, 4, 5, 6, 9, 10, 11, 12, 13, 14, 15, 16, 17], np.uint32),
                          np.arange(13, dtype=np.uint32))}
    rng = np.random.RandomState(3)
    for n in (2, 5, 13, 14, 15, 29, 30, 31, 45):
        for rep in range(3):
            # component ids are small and dense in practice; also sparse and large ones
            hi = (64, 4096, 2 ** 31)[rep]
            keys = distinct(rng, hi, n)
            out[f"random_n{n}_hi{hi}"] = (keys, rng.randint(0, 128, size=n).astype(np.uint32))
    return out


def ref_build(keys, values):
    lib = ctypes.CDLL(REF_LIB)
    lib.eorc_static_map_build.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint32,
                                          ctypes.c_void_p, ctypes.c_void_p]
    ks, vs = np.zeros(128, np.uint32), np.zeros(128, np.uint32)
    k, v = np.ascontiguousarray(keys, np.uint32), np.ascontiguousarray(values, np.uint32)
    lib.eorc_static_map_build(k.ctypes.data, v.ctypes.data, len(k), ks.ctypes.data, vs.ctypes.data)
    return ks, vs


def main():
    import subprocess
    subprocess.run(["make", "-C", os.path.join(ROOT, "oracle"), "ref"], check=True, capture_output=True)
    out = {}
    from gpu_ecs_madrona_b200 import ecs
    for name, (k, v) in cases().items():
        try:
            ecs.static_map_build(k, v)
        except ecs.MadronaError:
            # no perfect hash within 128 slots: the reference FATALs (aborts this process) on such a set
            print("skipped (unhashable):", name)
            continue
        ks, vs = ref_build(k, v)
        out[name + "__keys"], out[name + "__values"] = k, v
        out[name + "__key_storage"], out[name + "__value_storage"] = ks, vs
    np.savez_compressed(os.path.join(HERE, "static_map.npz"), **out)
    print("wrote", len(out) // 4, "maps")


if __name__ == "__main__":
    main()
