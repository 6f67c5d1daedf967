"""Ray sets for the traceRay parity tests (shared by the CPU and GPU suites)."""
import numpy as np


def make_rays(rng, n, world_pos):
    """n rays around one world's bodies: random directions, a quarter aimed down, some
    axis-aligned (zero direction components), the last eight starting inside a body."""
    lo, hi = world_pos.min(0) - 2, world_pos.max(0) + 2
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    o[:, 2] = rng.uniform(0.1, 6.0, n).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    d[: n // 4, 2] = -np.abs(d[: n // 4, 2]) - 0.5
    d[n // 4: n // 4 + 8] = np.array([0, 0, -1], np.float32)
    d[n // 4 + 8: n // 4 + 16] = np.array([1, 0, 0], np.float32)
    o[-8:] = world_pos[rng.integers(0, len(world_pos), 8)]
    return o, d


def world_rays(rng, n, pos):
    """[W, n, 3] origins and directions for positions [W, N, 3]."""
    rays = [make_rays(rng, n, pos[w]) for w in range(pos.shape[0])]
    return np.stack([r[0] for r in rays]), np.stack([r[1] for r in rays])


def assert_rays_equal(got, want, worlds, what):
    """(t, normal, row) bit-exact on `worlds`; normals only where something was hit."""
    gt, gn, grow = got
    wt, wn, wrow = want
    for w in worlds:
        assert np.array_equal(grow[w], wrow[w]), (what, w, np.nonzero(grow[w] != wrow[w]))
        assert np.array_equal(gt[w].view(np.uint32), wt[w].view(np.uint32)), (what, w)
        hit = wrow[w] >= 0
        assert np.array_equal(gn[w][hit].view(np.uint32), wn[w][hit].view(np.uint32)), (what, w)
        assert not gn[w][~hit].any(), (what, w)
