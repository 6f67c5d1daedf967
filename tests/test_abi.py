"""The C-ABI library loads and exports every symbol include/*.h declares.
No compute call is made here (no GPU on the CPU test machine)."""
import ctypes
import glob
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    names = set()
    for h in glob.glob(os.path.join(ROOT, "include", "*.h")):
        text = re.sub(r"/\*.*?\*/", "", open(h).read(), flags=re.S)
        if 'extern "C"' not in text:
            continue   # madrona_b200_apps.h: static inline workload arithmetic, no exported symbols
        names.update(re.findall(r"\b(mdr_[a-z0-9_]+)\s*\(", text))
    return sorted(names)


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    return ctypes.CDLL(g.build_cuda())


def test_header_declares_entry_points():
    names = declared_symbols()
    assert "mdr_phys_create" in names and "mdr_phys_step" in names
    assert "mdr_ecs_create" in names and "mdr_ecs_step" in names and "mdr_app_setup" in names
    assert len(names) >= 50


def test_library_exports_all_declared_symbols(lib):
    missing = [n for n in declared_symbols() if not hasattr(lib, n)]
    assert not missing, missing


def test_no_device_is_an_error_not_a_fallback(lib):
    """Without a GPU creation must fail with MDR_ENODEVICE; with one, bad
    arguments must be rejected before any device work."""
    from gpu_ecs_madrona_b200 import phys, scenes
    phys.bind(lib)
    lib.mdr_device_count.restype = ctypes.c_int32
    sc = scenes.collisions64(2)
    if lib.mdr_device_count() == 0:
        with pytest.raises(phys.MadronaError) as ei:
            phys.PhysExecutor(sc, lib=lib)
        assert ei.value.code == phys.MDR_ENODEVICE
    cfg = phys.MdrPhysConfig(0, 65, 4, 640, 1 / 60, (ctypes.c_float * 3)(0, 0, -9.8), 0, 0, 0)
    h = ctypes.c_void_p()
    assert lib.mdr_phys_create(ctypes.byref(cfg), None, 0, None, None, ctypes.byref(h)) == phys.MDR_EINVAL


def test_product_package_does_not_touch_the_oracle():
    pkg = os.path.join(ROOT, "gpu_ecs_madrona_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".hpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "oracle" not in text.replace("oracle/port.cpp", ""), f
                assert "libmdr_emu" not in text and "tests/emu" not in text.replace(
                    "tests/emu can run", "").replace("(tests/emu only)", ""), f
