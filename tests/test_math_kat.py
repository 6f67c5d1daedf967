"""The reference's own known-answer tests for the path's arithmetic, restated: tests/math.cpp:26-51
(Quaternions.Math, four answers at 1e-4), evaluated by the library's math source (csrc/phys_math.cuh through
mdr_math_quat) -- plus identities the step relies on."""
import ctypes as C
import math

import numpy as np

from gpu_ecs_madrona_b200 import phys


def quat(op, a, b=None):
    lib = phys.load_library()
    lib.mdr_math_quat.argtypes = [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
    a = np.ascontiguousarray(a, np.float32)
    b = None if b is None else np.ascontiguousarray(b, np.float32)
    out = np.zeros(4, np.float32)
    assert lib.mdr_math_quat(op, a.ctypes.data, None if b is None else b.ctypes.data, out.ctypes.data) == 0
    return out


def angle_axis(deg, axis):
    return quat(0, [math.radians(deg), *axis])


def test_reference_quaternion_known_answers():
    q1 = angle_axis(0, (0, 1, 0))
    assert np.allclose(q1, [1, 0, 0, 0], atol=1e-4)
    q2 = angle_axis(45, (0, 1, 0))
    assert np.allclose(q2, [0.9238795, 0, 0.3826834, 0], atol=1e-4)
    q3 = angle_axis(45, (1, 0, 0))
    assert np.allclose(q3, [0.9238795, 0.3826834, 0, 0], atol=1e-4)
    m1 = quat(1, q2, q3)
    assert np.allclose(m1, [0.853553, 0.353553, 0.353553, -0.146447], atol=1e-4)


def test_rotate_and_inverse_identities():
    q = quat(3, [0.3, -0.5, 0.7, 0.2])
    assert abs(float((q.astype(np.float64) ** 2).sum()) - 1.0) < 1e-6
    v = np.float32([1.5, -2.0, 0.25])
    r = quat(2, q, v)[1:]
    back = quat(2, quat(4, q), r)[1:]
    assert np.allclose(back, v, atol=1e-5)
    assert np.array_equal(quat(2, [1, 0, 0, 0], v)[1:], v)          # the identity leaves finite vectors bit for bit
    # 90 degrees about z takes x to y
    assert np.allclose(quat(2, angle_axis(90, (0, 0, 1)), [1, 0, 0])[1:], [0, 1, 0], atol=1e-6)
