import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def port_oracle():
    """Builds oracle/libmadrona_port.so (plain g++)."""
    from oracle import orc
    orc.build("port")
    return orc


@pytest.fixture(scope="session")
def ref_oracle():
    """The compiled reference; only where /root/reference is mounted or the
    prebuilt oracle/_ref travelled with the repo."""
    from oracle import orc
    if os.path.isdir("/root/reference/src/physics"):
        orc.build("ref")
    if not orc.available("reference"):
        pytest.skip("compiled reference not available on this machine")
    return orc


@pytest.fixture(scope="session")
def emu_lib():
    """tests/emu: the device source compiled for the host (test tool only)."""
    import ctypes
    from gpu_ecs_madrona_b200 import phys
    emu = os.path.join(ROOT, "tests", "emu")
    out = os.path.join(emu, "_build", "libmdr_emu.so")
    os.makedirs(os.path.dirname(out), exist_ok=True)
    srcs = [os.path.join(emu, "emu.cpp")] + [
        os.path.join(ROOT, "gpu_ecs_madrona_b200", "csrc", f)
        for f in os.listdir(os.path.join(ROOT, "gpu_ecs_madrona_b200", "csrc"))]
    if not os.path.exists(out) or any(os.path.getmtime(s) > os.path.getmtime(out) for s in srcs):
        tmp = "%s.%d.tmp" % (out, os.getpid())
        subprocess.run(["g++", "-std=c++17", "-O2", "-march=x86-64-v3", "-ffp-contract=off", "-fPIC",
                        "-shared", "-x", "c++", os.path.join(emu, "emu.cpp"), "-o", tmp], check=True)
        os.replace(tmp, out)    # atomic: concurrent xdist workers never load a half-written file
    return phys.bind(ctypes.CDLL(out))


@pytest.fixture(scope="session")
def cuda_lib():
    from gpu_ecs_madrona_b200 import phys
    lib = phys.load_library()
    if lib.mdr_device_count() < 1:
        pytest.fail("gpu test selected but no CUDA device is visible")
    return lib
