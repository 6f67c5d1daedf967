import os

import numpy as np

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def bits(a):
    return np.ascontiguousarray(a).view(np.uint32)


def worlds_equal(state_a, state_b):
    """Per-world bitwise equality of (pos, rot, vel) tuples -> bool[W]."""
    W = state_a[0].shape[0]
    eq = np.ones(W, bool)
    for x, y in zip(state_a, state_b):
        eq &= (bits(x).reshape(W, -1) == bits(y).reshape(W, -1)).all(axis=1)
    return eq


def assert_state_equal(got, want, mask=None, what=""):
    eq = worlds_equal(got, want)
    if mask is not None:
        eq = eq | ~mask
    assert eq.all(), f"{what}: worlds {np.nonzero(~eq)[0][:16].tolist()} differ bitwise"


def contacts_equal(a, b):
    """Bitwise comparison of two contact lists ignoring lambda_n (solver output)."""
    if len(a) != len(b):
        return False
    for f in ("ref_row", "alt_row", "num_points"):
        if not np.array_equal(a[f], b[f]):
            return False
    return (np.array_equal(bits(a["points"]), bits(b["points"]))
            and np.array_equal(bits(a["normal"]), bits(b["normal"])))


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))
