"""Deterministic synthetic scenes for the collisions hot path (numpy only).

The reference ships no runnable physics example, so the
workloads are defined here once and fed bit-identically to the CUDA executor
and to the CPU checkers. `collisions64` is BASELINE.json configs[2] as
SURVEY §8(d) "Config 3" defines it; `tumble` is a denser, rotated and moving
variant used by the parity tests to reach hull-hull face, edge and >4-point
manifold paths (src/physics/narrowphase.cpp:678-864).
"""
from dataclasses import dataclass, field

import numpy as np

HULL = 2   # CollisionPrimitive::Type::Hull  (include/madrona/physics.hpp:251-256)
PLANE = 4  # CollisionPrimitive::Type::Plane
DYNAMIC, KINEMATIC, STATIC = 0, 1, 2  # ResponseType (physics.hpp:162-166)

# Unit cube, half-extent 0.5: 8 vertices, 6 outward-facing quads (CCW seen from outside).
CUBE_VERTS = np.array([
    [-0.5, -0.5, -0.5], [0.5, -0.5, -0.5], [0.5, 0.5, -0.5], [-0.5, 0.5, -0.5],
    [-0.5, -0.5, 0.5], [0.5, -0.5, 0.5], [0.5, 0.5, 0.5], [-0.5, 0.5, 0.5],
], dtype=np.float32)
CUBE_POLYS = [
    [0, 3, 2, 1],  # -z
    [4, 5, 6, 7],  # +z
    [0, 1, 5, 4],  # -y
    [2, 3, 7, 6],  # +y
    [0, 4, 7, 3],  # -x
    [1, 2, 6, 5],  # +x
]


# JointConstraint (include/madrona/physics.hpp:196-238) with the entities as rows:
# the layout of mdr_joint (include/madrona_b200.h).
JOINT_FIXED, JOINT_HINGE = 0, 1
JOINT_DTYPE = np.dtype([("type", "<i4"), ("row1", "<i4"), ("row2", "<i4"),
                        ("r1", "<f4", (3,)), ("r2", "<f4", (3,)), ("data", "<f4", (12,))])
assert JOINT_DTYPE.itemsize == 84


@dataclass
class ObjectDef:
    """One shared collision object (reference: ObjectManager row,
    include/madrona/physics.hpp:243-290)."""
    type: int
    inv_inertia: tuple = (0.0, 0.0, 0.0)
    inv_mass: float = 0.0
    mu_s: float = 0.5
    mu_d: float = 0.5
    aabb_min: tuple = (0.0, 0.0, 0.0)
    aabb_max: tuple = (0.0, 0.0, 0.0)
    verts: np.ndarray = field(default_factory=lambda: np.zeros((0, 3), np.float32))
    polys: list = field(default_factory=list)

    def poly_words(self):
        """FastPolygonList encoding [n, i0..in-1]* (physics.hpp:12-35)."""
        out = []
        for p in self.polys:
            out.append(len(p))
            out.extend(p)
        return np.asarray(out, dtype=np.uint32)


def cube_object(half=0.5, inv_mass=1.0, mu=0.5):
    # solid cube: I = m s^2 / 6 with s = 2*half -> invI = 6 invM / s^2
    s = 2.0 * half
    inv_i = 6.0 * inv_mass / (s * s)
    return ObjectDef(HULL, (inv_i,) * 3, inv_mass, mu, mu,
                     (-half,) * 3, (half,) * 3,
                     (CUBE_VERTS * np.float32(2.0 * half)).astype(np.float32),
                     [list(p) for p in CUBE_POLYS])


def box_object(hx, hy, hz, inv_mass=1.0, mu=0.5):
    v = (CUBE_VERTS * np.array([2 * hx, 2 * hy, 2 * hz], np.float32)).astype(np.float32)
    m = 1.0 / inv_mass
    ix = m * ((2 * hy) ** 2 + (2 * hz) ** 2) / 12.0
    iy = m * ((2 * hx) ** 2 + (2 * hz) ** 2) / 12.0
    iz = m * ((2 * hx) ** 2 + (2 * hy) ** 2) / 12.0
    return ObjectDef(HULL, (1.0 / ix, 1.0 / iy, 1.0 / iz), inv_mass, mu, mu,
                     (-hx, -hy, -hz), (hx, hy, hz), v, [list(p) for p in CUBE_POLYS])


def plane_object(mu=0.5):
    # A finite box stands in for the half-space: FLT_MAX overflows in
    # AABB::applyTRS.
    return ObjectDef(PLANE, (0.0, 0.0, 0.0), 0.0, mu, mu,
                     (-100.0, -100.0, -1.0), (100.0, 100.0, 0.0))


@dataclass
class Scene:
    num_worlds: int
    bodies_per_world: int
    num_substeps: int
    max_contacts: int
    delta_t: float
    gravity: tuple
    objects: list
    pos: np.ndarray        # [W, N, 3] f32
    rot: np.ndarray        # [W, N, 4] f32 (w, x, y, z)
    scale: np.ndarray      # [W, N, 3] f32
    vel: np.ndarray        # [W, N, 6] f32 (linear, angular)
    obj_ids: np.ndarray    # [W, N] i32
    response: np.ndarray   # [W, N] u32
    ext_force: np.ndarray  # [W, N, 3] f32
    ext_torque: np.ndarray # [W, N, 3] f32
    name: str = ""
    joints: np.ndarray = None        # [W, J] JOINT_DTYPE (ConstraintData rows), optional
    joint_counts: np.ndarray = None  # [W] i32


def _lcg_uniform(seeds, count):
    """u[k] of the scene LCG: s = s*1664525 + 1013904223 (mod 2^32),
    u = (s >> 8) / 2^24, `count` draws per seed. Returns [len(seeds), count] f64."""
    s = seeds.astype(np.uint64) & np.uint64(0xFFFFFFFF)
    out = np.empty((len(seeds), count), dtype=np.float64)
    for k in range(count):
        s = (s * np.uint64(1664525) + np.uint64(1013904223)) & np.uint64(0xFFFFFFFF)
        out[:, k] = (s >> np.uint64(8)).astype(np.float64) / float(1 << 24)
    return out


def collisions64(num_worlds, world_offset=0, num_cubes=64, num_substeps=4,
                 max_contacts=640):
    """BASELINE.json configs[2] / SURVEY §8(d) Config 3. Row 0 is the static
    plane, rows 1..64 the dynamic unit cubes. `world_offset` is the global index
    of this shard's first world, so shards are independent of the GPU count
    (SURVEY §8e)."""
    W, N = num_worlds, num_cubes + 1
    side = int(np.ceil(np.sqrt(num_cubes)))
    gidx = np.arange(world_offset, world_offset + W, dtype=np.uint64)
    seeds = (gidx * np.uint64(2654435761) + np.uint64(1)) & np.uint64(0xFFFFFFFF)
    u = _lcg_uniform(seeds, 3 * num_cubes).reshape(W, num_cubes, 3)
    k = np.arange(num_cubes)
    base = np.stack([
        (k % side - (side - 1) / 2.0) * 1.25,
        (k // side - (side - 1) / 2.0) * 1.25,
        0.5 + 0.05 + 1.5 * (k % 3),
    ], axis=-1)
    cubes = base[None] + (u * 2.0 - 1.0) * 0.1
    pos = np.zeros((W, N, 3), np.float32)
    pos[:, 1:] = cubes.astype(np.float32)
    rot = np.zeros((W, N, 4), np.float32)
    rot[..., 0] = 1.0
    obj_ids = np.zeros((W, N), np.int32)
    obj_ids[:, 0] = 1
    response = np.zeros((W, N), np.uint32)
    response[:, 0] = STATIC
    return Scene(W, N, num_substeps, max_contacts, 1.0 / 60.0, (0.0, 0.0, -9.8),
                 [cube_object(), plane_object()],
                 pos, rot, np.ones((W, N, 3), np.float32),
                 np.zeros((W, N, 6), np.float32), obj_ids, response,
                 np.zeros((W, N, 3), np.float32), np.zeros((W, N, 3), np.float32),
                 name=f"collisions-{num_cubes}")


def tumble(num_worlds, num_bodies=24, seed=1, num_substeps=4, max_contacts=512,
           spread=1.6, with_plane=True):
    """Dense pile of randomly rotated, moving cubes and boxes (two hull objects,
    non-unit scales) over a plane: exercises hull-hull face/edge contacts,
    clipping to >4 points, restitution and friction. Parity-test scene only."""
    W = num_worlds
    N = num_bodies + (1 if with_plane else 0)
    seeds = (np.arange(W, dtype=np.uint64) * np.uint64(747796405)
             + np.uint64(seed)) & np.uint64(0xFFFFFFFF)
    u = _lcg_uniform(seeds, 16 * num_bodies).reshape(W, num_bodies, 16)
    p = np.empty((W, num_bodies, 3))
    p[..., 0] = (u[..., 0] * 2 - 1) * spread
    p[..., 1] = (u[..., 1] * 2 - 1) * spread
    p[..., 2] = 0.8 + u[..., 2] * 2.5
    q = u[..., 3:7] * 2 - 1
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    v = np.concatenate([(u[..., 7:10] * 2 - 1) * 2.0, (u[..., 10:13] * 2 - 1) * 3.0], -1)
    sc = 0.6 + 0.8 * u[..., 13:16]
    first = 1 if with_plane else 0
    pos = np.zeros((W, N, 3), np.float32)
    rot = np.zeros((W, N, 4), np.float32)
    rot[..., 0] = 1.0
    scale = np.ones((W, N, 3), np.float32)
    vel = np.zeros((W, N, 6), np.float32)
    pos[:, first:] = p.astype(np.float32)
    rot[:, first:] = q.astype(np.float32)
    scale[:, first:] = sc.astype(np.float32)
    vel[:, first:] = v.astype(np.float32)
    obj_ids = np.zeros((W, N), np.int32)
    obj_ids[:, first:] = (np.arange(num_bodies) % 2)[None]
    response = np.zeros((W, N), np.uint32)
    if with_plane:
        obj_ids[:, 0] = 2
        response[:, 0] = STATIC
    f = np.zeros((W, N, 3), np.float32)
    t = np.zeros((W, N, 3), np.float32)
    f[:, first::3, 0] = 0.5
    t[:, first::4, 2] = 0.25
    return Scene(W, N, num_substeps, max_contacts, 1.0 / 60.0, (0.0, 0.0, -9.8),
                 [cube_object(), box_object(0.5, 0.3, 0.2, inv_mass=2.0, mu=0.3), plane_object()],
                 pos, rot, scale, vel, obj_ids, response, f, t, name=f"tumble-{num_bodies}")


HULL_SHAPES = ("cube", "dodeca", "frustum", "hex_prism", "oct_frustum", "octa", "pyramid", "rock",
               "tetra", "tri_prism")


def hull_pile(num_worlds, num_bodies=20, seed=11, num_substeps=4, max_contacts=512, spread=1.5,
              hull_dir=None, shapes=HULL_SHAPES):
    """Non-cube convex hulls loaded from OBJ files through the asset pipeline (SURVEY 8f rank 4:
    gpu_ecs_madrona_b200.assets -> mdr_load_hull_obj_file), dropped in a pile over a plane:
    triangles to octagons as faces, 4 to 20 vertices, non-parallel faces, different masses and
    friction, non-uniform scales. Exercises the general hull paths of the narrowphase (face
    queries over unequal face counts, edge-edge over unequal edge counts, clipping of n-gons,
    manifold reduction). Parity-test scene only; the OBJ files live in gpu_ecs_madrona_b200/hulls
    (written by tests/golden/make_asset_golden.py)."""
    import os
    from . import assets
    if hull_dir is None:
        hull_dir = os.path.join(os.path.dirname(os.path.abspath(__file__)), "hulls")
    objects = [assets.hull_object(os.path.join(hull_dir, n + ".obj"), inv_mass=0.5 + 0.25 * (k % 4),
                                  mu=0.3 + 0.1 * (k % 3)) for k, n in enumerate(shapes)]
    objects.append(plane_object())
    H = len(shapes)
    W, N = num_worlds, num_bodies + 1
    seeds = (np.arange(W, dtype=np.uint64) * np.uint64(3935559000) + np.uint64(seed)) & np.uint64(0xFFFFFFFF)
    u = _lcg_uniform(seeds, 16 * num_bodies).reshape(W, num_bodies, 16)
    p = np.empty((W, num_bodies, 3))
    p[..., 0] = (u[..., 0] * 2 - 1) * spread
    p[..., 1] = (u[..., 1] * 2 - 1) * spread
    p[..., 2] = 0.7 + u[..., 2] * 2.5
    q = u[..., 3:7] * 2 - 1
    q /= np.linalg.norm(q, axis=-1, keepdims=True)
    v = np.concatenate([(u[..., 7:10] * 2 - 1) * 2.0, (u[..., 10:13] * 2 - 1) * 3.0], -1)
    sc = 0.7 + 0.6 * u[..., 13:16]
    pos = np.zeros((W, N, 3), np.float32)
    rot = np.zeros((W, N, 4), np.float32)
    rot[..., 0] = 1.0
    scale = np.ones((W, N, 3), np.float32)
    vel = np.zeros((W, N, 6), np.float32)
    pos[:, 1:] = p.astype(np.float32)
    rot[:, 1:] = q.astype(np.float32)
    scale[:, 1:] = sc.astype(np.float32)
    vel[:, 1:] = v.astype(np.float32)
    obj_ids = np.zeros((W, N), np.int32)
    # every world starts the cycle of shapes somewhere else
    obj_ids[:, 1:] = (np.arange(num_bodies)[None] + np.arange(W)[:, None]) % H
    obj_ids[:, 0] = H
    response = np.zeros((W, N), np.uint32)
    response[:, 0] = STATIC
    z = np.zeros((W, N, 3), np.float32)
    return Scene(W, N, num_substeps, max_contacts, 1.0 / 60.0, (0.0, 0.0, -9.8), objects,
                 pos, rot, scale, vel, obj_ids, response, z, z.copy(), name=f"hull-pile-{num_bodies}")


def jointed(num_worlds, seed=7, num_substeps=4, max_contacts=256, loose=6, hinges=False):
    """Joint constraints (JointConstraint::setupFixed / setupHinge,
    include/madrona/physics.inl:151-191) next to contacts. Per world, over a static plane:
      * a pendulum: a static anchor cube in the air and three cubes hanging below it,
        linked by hinges about a common axis (static-dynamic and dynamic-dynamic joints);
      * a rigid dumbbell: two boxes held apart by a fixed joint with rotated attach
        frames and a non-zero separation, dropped so that it lands on the plane;
      * a hinged pair lying on the plane from the start (joints and contacts on the
        same bodies in the same substep);
      * `loose` free cubes thrown at the rest.
    Worlds differ by LCG jitter of poses, velocities and joint frames; every second
    world leaves its last joint out (ragged joint counts). Parity-test scene only.

    `hinges=False` swaps every hinge for a fixed joint with the same anchors. In this
    snapshot of the reference the hinge pushes its bodies AWAY from the constraint
    (handleJointConstraint hands applyPositionalUpdate `r2 - r1`, physics.cpp:617, where the
    fixed joint hands it the negated error, :594-605; applyJointAxisConstraint has the same
    sign and a debug printf, :521-535), so hinged bodies leave to infinity within ~10
    steps: hinge parity is checked bit for bit on the first steps, while the state is finite."""
    W = num_worlds
    nb = 4 + 2 + 2 + loose
    N = 1 + nb
    seeds = (np.arange(W, dtype=np.uint64) * np.uint64(2891336453) + np.uint64(seed)) & np.uint64(0xFFFFFFFF)
    u = _lcg_uniform(seeds, 16 * nb + 32).reshape(W, -1)
    ub, uj = u[:, :16 * nb].reshape(W, nb, 16), u[:, 16 * nb:]
    pos = np.zeros((W, N, 3))
    rot = np.zeros((W, N, 4))
    rot[..., 0] = 1.0
    scale = np.ones((W, N, 3), np.float32)
    vel = np.zeros((W, N, 6))
    obj_ids = np.zeros((W, N), np.int32)
    response = np.zeros((W, N), np.uint32)
    obj_ids[:, 0] = 2
    response[:, 0] = STATIC
    jit = (ub[..., 0:3] * 2 - 1)
    # pendulum: rows 1 (static anchor) .. 4
    for k in range(4):
        pos[:, 1 + k] = np.array([-3.0, 0.0, 6.0 - 1.2 * k]) + (0.05 * jit[:, k] if k else 0.0)
        vel[:, 1 + k, 0] = 1.5 * k * (ub[:, k, 3] * 2 - 1)
        vel[:, 1 + k, 1] = 0.5 * k * (ub[:, k, 4] * 2 - 1)
    response[:, 1] = STATIC
    # dumbbell: rows 5, 6 (boxes, object 1)
    obj_ids[:, 5:7] = 1
    pos[:, 5] = np.array([0.0, -0.6, 1.2]) + 0.05 * jit[:, 4]
    pos[:, 6] = np.array([0.0, 0.6, 1.3]) + 0.05 * jit[:, 5]
    vel[:, 5:7, 3:6] = (ub[:, 4:6, 5:8] * 2 - 1) * 2.0
    # hinged pair on the ground: rows 7, 8
    pos[:, 7] = np.array([3.0, -0.55, 0.5]) + 0.02 * jit[:, 6] * np.array([1, 1, 0])
    pos[:, 8] = np.array([3.0, 0.55, 0.5]) + 0.02 * jit[:, 7] * np.array([1, 1, 0])
    vel[:, 8, 0] = 1.0
    # loose cubes
    for k in range(loose):
        b = 8 + k
        pos[:, 1 + b] = np.array([-3.0 + 1.5 * k, -2.0, 2.0 + 0.7 * (k % 3)]) + 0.3 * jit[:, b]
        q = ub[:, b, 8:12] * 2 - 1
        rot[:, 1 + b] = q / np.linalg.norm(q, axis=-1, keepdims=True)
        vel[:, 1 + b, 0:3] = np.array([0.0, 3.0, 0.0]) + (ub[:, b, 12:15] * 2 - 1)
    J = 6
    joints = np.zeros((W, J), JOINT_DTYPE)
    counts = np.full(W, J, np.int32)
    counts[1::2] = J - 1
    axis = np.array([0.0, 1.0, 0.0])
    for k in range(3):                      # anchor-1, 1-2, 2-3
        j = joints[:, k]
        j["type"], j["row1"], j["row2"] = JOINT_HINGE, 1 + k, 2 + k
        j["r1"], j["r2"] = (0.0, 0.0, -0.6), (0.0, 0.0, 0.6)
        a2 = axis[None] + 0.05 * (uj[:, 3 * k:3 * k + 3] * 2 - 1)
        a2 /= np.linalg.norm(a2, axis=-1, keepdims=True)
        j["data"][:, 0:3] = axis
        j["data"][:, 3:6] = a2
        j["data"][:, 6:9] = (1.0, 0.0, 0.0)
        j["data"][:, 9:12] = (1.0, 0.0, 0.0)
    j = joints[:, 3]                        # dumbbell
    j["type"], j["row1"], j["row2"] = JOINT_FIXED, 5, 6
    j["r1"], j["r2"] = (0.0, 0.3, 0.0), (0.0, -0.3, 0.0)
    q1 = np.array([1.0, 0.0, 0.0, 0.0])[None] + 0.2 * (uj[:, 9:13] * 2 - 1)
    q1 /= np.linalg.norm(q1, axis=-1, keepdims=True)
    q2 = np.array([1.0, 0.0, 0.0, 0.0])[None] + 0.2 * (uj[:, 13:17] * 2 - 1)
    q2 /= np.linalg.norm(q2, axis=-1, keepdims=True)
    j["data"][:, 0:4], j["data"][:, 4:8] = q1, q2
    j["data"][:, 8] = 0.5 + 0.2 * uj[:, 17]
    j = joints[:, 4]                        # ground pair
    j["type"], j["row1"], j["row2"] = JOINT_HINGE, 7, 8
    j["r1"], j["r2"] = (0.0, 0.55, 0.0), (0.0, -0.55, 0.0)
    j["data"][:, 0:3], j["data"][:, 3:6] = (0.0, 0.0, 1.0), (0.0, 0.0, 1.0)
    j = joints[:, 5]                        # a loose cube tied to the pendulum's tip (odd worlds: absent)
    j["type"], j["row1"], j["row2"] = JOINT_FIXED, 4, 9
    j["r1"], j["r2"] = (0.5, 0.0, 0.0), (-0.5, 0.0, 0.0)
    j["data"][:, 0], j["data"][:, 4] = 1.0, 1.0
    j["data"][:, 8] = 0.0
    if not hinges:
        for k in (0, 1, 2, 4):
            j = joints[:, k]
            j["type"] = JOINT_FIXED
            j["data"] = 0.0
            j["data"][:, 0], j["data"][:, 4] = 1.0, 1.0
    for w in np.nonzero(counts < J)[0]:
        joints[w, counts[w]:] = np.zeros((), JOINT_DTYPE)
    z = np.zeros((W, N, 3), np.float32)
    return Scene(W, N, num_substeps, max_contacts, 1.0 / 60.0, (0.0, 0.0, -9.8),
                 [cube_object(), box_object(0.5, 0.3, 0.2, inv_mass=2.0, mu=0.3), plane_object()],
                 pos.astype(np.float32), rot.astype(np.float32), scale, vel.astype(np.float32),
                 obj_ids, response, z, z.copy(), name="jointed-hinges" if hinges else "jointed",
                 joints=joints, joint_counts=counts)
