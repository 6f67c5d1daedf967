"""Workload definitions of the ECS / task-graph configs (BASELINE.json configs[0], [1], [3];
SURVEY.md section 8d configs 1, 2, 4): the initial column values of every world, generated
from the GLOBAL world index with a counter hash so that results do not depend on how worlds
are sharded over GPUs. The per-row arithmetic of the systems lives in
include/madrona_b200_apps.h; this file only makes input data (numpy, host side).
"""
import numpy as np

APP_SCRIPT, APP_TASKGRAPH, APP_FANOUT, APP_FVS, APP_CHURN = range(5)   # churn: out-of-tree world (examples/churn)
APP_NAMES = {APP_SCRIPT: "script", APP_TASKGRAPH: "simple_taskgraph", APP_FANOUT: "simple_jobs_fanout",
             APP_FVS: "fantasy_vs"}
DEFAULT_PARAMS = {APP_SCRIPT: [1], APP_TASKGRAPH: [100], APP_FANOUT: [100], APP_FVS: [50, 200]}

BOUNDS_MIN = np.array([-10.0, -10.0, 0.0], np.float32)
BOUNDS_MAX = np.array([10.0, 10.0, 10.0], np.float32)


def _fmix(h):
    h = h.astype(np.uint32)
    h ^= h >> np.uint32(16)
    h *= np.uint32(0x85EBCA6B)
    h ^= h >> np.uint32(13)
    h *= np.uint32(0xC2B2AE35)
    h ^= h >> np.uint32(16)
    return h


def _u01(world, item, stream):
    """[W, n] uniform floats in [0, 1) keyed by (global world, item, stream)."""
    with np.errstate(over="ignore"):
        h = _fmix(world[:, None].astype(np.uint32) * np.uint32(0x9E3779B1) + np.uint32(stream))
        h = _fmix(h ^ (item[None, :].astype(np.uint32) * np.uint32(0x85EBCA77) + np.uint32(0x165667B1)))
    return (h >> np.uint32(8)).astype(np.float32) * np.float32(1.0 / 16777216.0)


def _positions(worlds, n, stream):
    item = np.arange(n, dtype=np.uint32)
    u = np.stack([_u01(worlds, item, stream + k) for k in range(3)], axis=-1)
    return (BOUNDS_MIN + (BOUNDS_MAX - BOUNDS_MIN) * u).astype(np.float32)


def _identity_rot(W, n):
    q = np.zeros((W, n, 4), np.float32)
    q[..., 0] = 1.0
    return q


def initial_columns(app, num_worlds, params=None, first_world=0):
    """{(archetype, column): array [num_worlds, rows, ...]} for the world constructors."""
    params = list(params if params is not None else DEFAULT_PARAMS[app])
    worlds = np.arange(first_world, first_world + num_worlds, dtype=np.uint32)
    W = num_worlds
    item = lambda n: np.arange(n, dtype=np.uint32)
    if app == APP_SCRIPT:
        return {}
    if app == APP_TASKGRAPH:
        n = params[0]
        return {(0, 1): _positions(worlds, n, 100), (0, 2): _identity_rot(W, n),
                (1, 1): _positions(worlds, 1, 200), (1, 2): _identity_rot(W, 1)}
    if app == APP_FANOUT:
        n = params[0]
        ang = _u01(worlds, item(n), 300) * np.float32(2 * np.pi)
        axis = np.stack([_u01(worlds, item(n), 301 + k) - np.float32(0.5) for k in range(3)], axis=-1)
        axis /= np.maximum(np.linalg.norm(axis, axis=-1, keepdims=True), 1e-6)
        q = np.concatenate([np.cos(ang / 2)[..., None], np.sin(ang / 2)[..., None] * axis], axis=-1).astype(np.float32)
        q /= np.linalg.norm(q, axis=-1, keepdims=True)
        return {(0, 1): _positions(worlds, n, 310), (0, 2): q.astype(np.float32),
                (0, 3): np.zeros((W, n, 6), np.float32), (0, 4): np.zeros((W, n), np.int32)}
    if app == APP_FVS:
        nd, nk = params[0], params[1]
        return {
            (0, 1): _positions(worlds, nd, 400), (0, 2): np.full((W, nd), 300, np.int32),
            (0, 3): np.zeros((W, nd), np.float32), (0, 4): (_u01(worlds, item(nd), 410) * np.float32(50)).astype(np.float32),
            (1, 1): _positions(worlds, nk, 420), (1, 2): np.full((W, nk), 100, np.int32),
            (1, 3): np.zeros((W, nk), np.float32),
            (1, 4): (20 + (_u01(worlds, item(nk), 430) * np.float32(21)).astype(np.int32)).astype(np.int32),
        }
    raise ValueError(app)


def random_script(rng, num_ops, max_live=400, p_clear=0.01):
    """A random immediate-mode op sequence for the script app (tests/state.cpp style churn)."""
    ops = np.zeros(num_ops, np.dtype([("op", "<i4"), ("a", "<i4"), ("b", "<i4")]))
    handles, live = 0, []
    mode, left = 1, 0
    for i in range(num_ops):
        if left == 0:
            mode = int(rng.integers(1, 3))
            left = int(rng.integers(1, 90))
        left -= 1
        r = rng.random()
        if r < p_clear:
            ops[i] = (3, int(rng.integers(0, 3)), 0)
        elif r < 2 * p_clear:
            ops[i] = (5, 0, 0)
        elif r < 0.08:
            ops[i] = (4, 0, int(rng.integers(0, 1 << 20)))
        elif mode == 2 and handles > 0:
            # mostly live handles, sometimes a stale one (destroyEntityNow of a dead entity is a no-op)
            h = live.pop(int(rng.integers(0, len(live)))) if live and rng.random() < 0.9 else int(rng.integers(0, handles))
            ops[i] = (2, h, 0)
        else:
            ops[i] = (1, int(rng.integers(0, 3)), int(rng.integers(0, 1 << 20)))
            live.append(handles)
            handles += 1
            if len(live) > max_live:
                mode, left = 2, int(rng.integers(20, 200))
    return ops
