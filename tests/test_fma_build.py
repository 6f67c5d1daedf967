"""SURVEY H1 / VERDICT item 2: the FMA-contracted build (libmadrona_b200_fma.so, the same sources with
-fmad=true). What it keeps exact and what it does not, enforced:
  * integer state after one step from the same inputs is bit-exact: candidate-pair lists (the AABB math is
    never contracted, csrc/phys_math.cuh xmul/xadd), pair counts, contact counts;
  * positions and rotations after one step: <= 1e-5 relative (north_star's bound);
  * velocities after one step: <= 2.4e-3 -- NOT 1e-5. A velocity is (x - x_prev) / h with h = 1/240 s: one ulp
    of a position of magnitude 4 (4.8e-7) is already 1.1e-4 in the velocity, so no build whose positions
    differ in the last bit can meet 1e-5 on velocities. The bound is 240 x the position bound; this is why
    the bit-exact build stays the default and the headline (DESIGN.md "Float parity");
  * after 100 steps the worlds of a chaotic scene have diverged (contacts made or missed by an ulp):
    the stated drift bound is on the resting scene only, collisions-64: |dx| <= 0.05 body sizes.
Relative error = |a - b| / max(|b|, 1) per component, b = the oracle."""
import ctypes as C
import os

import numpy as np
import pytest

from gpu_ecs_madrona_b200 import phys, scenes

pytestmark = pytest.mark.gpu
FMA = os.path.join(os.path.dirname(phys.LIB_PATH), "libmadrona_b200_fma.so")


def rel(a, b):
    a, b = a.astype(np.float64), b.astype(np.float64)
    return float(np.max(np.abs(a - b) / np.maximum(np.abs(b), 1.0)))


@pytest.fixture(scope="module")
def fma_lib(cuda_lib):
    if not os.path.exists(FMA):
        pytest.fail(f"{FMA} is not built (python -c 'import __graft_entry__ as g; g.build()')")
    return phys.bind(C.CDLL(FMA))


@pytest.mark.parametrize("name,make", [("collisions64", lambda: scenes.collisions64(64)),
                                       ("tumble", lambda: scenes.tumble(48)),
                                       ("hull_pile", lambda: scenes.hull_pile(32))])
def test_fma_build_one_step(fma_lib, port_oracle, name, make):
    sc = make()
    o = port_oracle.Oracle(sc, "port")
    ex = phys.PhysExecutor(sc)
    fm = phys.PhysExecutor(sc, lib=fma_lib)
    for w in (0, sc.num_worlds - 1):
        assert np.array_equal(ex.probe_step(w)["pairs"], fm.probe_step(w)["pairs"]) if w == 0 else True
    o.step(1, 8)
    # probe_step stepped world batches once already (it runs one step with the probe on)
    (pe, ce), (pf, cf) = ex.stats(), fm.stats()
    assert np.array_equal(pe, pf), "candidate-pair counts"
    assert np.array_equal(ce, cf), "contact counts"
    for got, want in zip(ex.state(), o.state()):
        assert np.array_equal(got.view(np.uint32), want.view(np.uint32))       # the default build is bit-exact
    pos, rot, vel = fm.state()
    opos, orot, ovel = o.state()
    assert rel(pos, opos) <= 1e-5
    assert rel(rot, orot) <= 1e-5
    assert rel(vel, ovel) <= 2.4e-3
    ex.close(); fm.close(); o.close()


def test_fma_build_drift_on_the_resting_scene(fma_lib, port_oracle):
    sc = scenes.collisions64(64)
    o = port_oracle.Oracle(sc, "port")
    fm = phys.PhysExecutor(sc, lib=fma_lib)
    o.step(100, 8)
    fm.run(100)
    pos, _, _ = fm.state()
    opos, _, _ = o.state()
    assert float(np.abs(pos.astype(np.float64) - opos).max()) <= 0.05
    assert np.isfinite(pos).all()
    fm.close(); o.close()
