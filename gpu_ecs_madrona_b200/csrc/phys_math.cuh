// fp32 vector / quaternion / AABB arithmetic of the world step.
//
// The association order of every expression is part of the parity contract with
// the reference's CPU path (include/madrona/math.hpp:195-476 Vector3, :478-730
// Quat, :807-905 Mat3x3, :985-1128 AABB): this translation unit is compiled with
// -fmad=false, IEEE division and square root, no flush-to-zero, so each
// operation rounds exactly like the reference built by g++ (-ffp-contract=off).
//
// Everything is MDR_HD so that tests/emu can run the same source on the host
// (test infrastructure only; the product library is device code).
#pragma once
#include <stdint.h>
#include <math.h>
#include <float.h>

#if defined(__CUDACC__)
#define MDR_HD __host__ __device__ __forceinline__
#else
#define MDR_HD inline
#endif

namespace mdr {

// Arithmetic that is never contracted into an FMA, whatever -fmad says: the AABB math
// below decides the broadphase pair list (integer state, bit-exact in every build), so
// the FMA-contracted tolerance build (libmadrona_b200_fma.so, -fmad=true) keeps it exact.
#if defined(__CUDA_ARCH__)
MDR_HD float xmul(float a, float b) { return __fmul_rn(a, b); }
MDR_HD float xadd(float a, float b) { return __fadd_rn(a, b); }
MDR_HD float xsub(float a, float b) { return __fsub_rn(a, b); }
#else
MDR_HD float xmul(float a, float b) { return a * b; }
MDR_HD float xadd(float a, float b) { return a + b; }
MDR_HD float xsub(float a, float b) { return a - b; }
#endif

struct V3 { float x, y, z; };
struct Q4 { float w, x, y, z; };
struct Box { V3 lo, hi; };
struct M3 { V3 c[3]; };
struct Plane { V3 n; float d; };

MDR_HD V3 mk3(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
MDR_HD Q4 mk4(float w, float x, float y, float z) { Q4 r; r.w = w; r.x = x; r.y = y; r.z = z; return r; }

MDR_HD V3 operator+(V3 a, V3 b) { return mk3(a.x + b.x, a.y + b.y, a.z + b.z); }
MDR_HD V3 operator-(V3 a, V3 b) { return mk3(a.x - b.x, a.y - b.y, a.z - b.z); }
MDR_HD V3 operator-(V3 a) { return mk3(-a.x, -a.y, -a.z); }
MDR_HD V3 operator*(V3 a, float s) { return mk3(a.x * s, a.y * s, a.z * s); }
MDR_HD V3 operator*(float s, V3 a) { return mk3(a.x * s, a.y * s, a.z * s); }
// math.hpp:345-350: vector / scalar multiplies by the reciprocal.
MDR_HD V3 operator/(V3 a, float s) { float inv = 1.f / s; return a * inv; }
MDR_HD float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
MDR_HD V3 cross(V3 a, V3 b)
{
    return mk3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
MDR_HD float len2(V3 a) { return a.x * a.x + a.y * a.y + a.z * a.z; }
MDR_HD float len(V3 a) { return sqrtf(len2(a)); }
// math.hpp:239-247, CPU branch (1.f / length(), not rsqrtf).
MDR_HD V3 normalize(V3 a) { return a * (1.f / len(a)); }
MDR_HD V3 mulDiag(V3 d, V3 v) { return mk3(d.x * v.x, d.y * v.y, d.z * v.z); }
MDR_HD float comp(const V3 &v, int i) { return i == 0 ? v.x : (i == 1 ? v.y : v.z); }
MDR_HD void setComp(V3 &v, int i, float f) { if (i == 0) v.x = f; else if (i == 1) v.y = f; else v.z = f; }

// libm fminf/fmaxf as the reference's host build resolves them (x86-64 glibc,
// measured: the FIRST operand is returned on ties, fminf(+0,-0) = +0): differs
// from PTX min/max only in the sign of a zero result, which bit-exact tests see.
MDR_HD float fminRef(float x, float y) { return y < x ? y : x; }
MDR_HD float fmaxRef(float x, float y) { return y > x ? y : x; }
MDR_HD V3 vmin(V3 a, V3 b) { return mk3(fminRef(a.x, b.x), fminRef(a.y, b.y), fminRef(a.z, b.z)); }
MDR_HD V3 vmax(V3 a, V3 b) { return mk3(fmaxRef(a.x, b.x), fmaxRef(a.y, b.y), fmaxRef(a.z, b.z)); }

MDR_HD Q4 operator+(Q4 a, Q4 b) { return mk4(a.w + b.w, a.x + b.x, a.y + b.y, a.z + b.z); }
MDR_HD Q4 operator-(Q4 a, Q4 b) { return mk4(a.w - b.w, a.x - b.x, a.y - b.y, a.z - b.z); }
// math.hpp:707-715
MDR_HD Q4 operator*(Q4 a, Q4 b)
{
    return mk4((a.w * b.w - a.x * b.x - a.y * b.y - a.z * b.z),
               (a.w * b.x + a.x * b.w + a.y * b.z - a.z * b.y),
               (a.w * b.y - a.x * b.z + a.y * b.w + a.z * b.x),
               (a.w * b.z + a.x * b.y - a.y * b.x + a.z * b.w));
}
MDR_HD Q4 qinv(Q4 q) { return mk4(q.w, -q.x, -q.y, -q.z); }
// math.hpp:550-561
MDR_HD Q4 qangleAxis(float angle, V3 normal)
{
    const float coshalf = cosf(angle / 2.f), sinhalf = sinf(angle / 2.f);
    return mk4(coshalf, normal.x * sinhalf, normal.y * sinhalf, normal.z * sinhalf);
}
// math.hpp:500-520
MDR_HD Q4 qnormalize(Q4 q)
{
    float inv = 1.f / sqrtf(q.w * q.w + q.x * q.x + q.y * q.y + q.z * q.z);
    return mk4(q.w * inv, q.x * inv, q.y * inv, q.z * inv);
}
// math.hpp:539-548: v + 2((p x v) w + p x (p x v))
MDR_HD V3 rotate(Q4 q, V3 v)
{
    V3 pure = mk3(q.x, q.y, q.z);
    V3 pxv = cross(pure, v);
    V3 ppxv = cross(pure, pxv);
    return v + 2.f * ((pxv * q.w) + ppxv);
}
MDR_HD Q4 fromAngular(V3 v) { return mk4(0.f, v.x, v.y, v.z); }
MDR_HD bool qBitsEqual(Q4 a, Q4 b) { return a.w == b.w && a.x == b.x && a.y == b.y && a.z == b.z; }

// math.hpp:810-834
MDR_HD M3 matFromQuat(Q4 r)
{
    float x2 = r.x * r.x, y2 = r.y * r.y, z2 = r.z * r.z;
    float xz = r.x * r.z, xy = r.x * r.y, yz = r.y * r.z;
    float wx = r.w * r.x, wy = r.w * r.y, wz = r.w * r.z;
    M3 m;
    m.c[0] = mk3(1.f - 2.f * (y2 + z2), 2.f * (xy + wz), 2.f * (xz - wy));
    m.c[1] = mk3(2.f * (xy - wz), 1.f - 2.f * (x2 + z2), 2.f * (yz + wx));
    m.c[2] = mk3(2.f * (xz + wy), 2.f * (yz - wx), 1.f - 2.f * (x2 + y2));
    return m;
}
// math.hpp:835-866 (only the AABB transform uses it: never contracted, see xmul)
MDR_HD M3 matFromRS(Q4 r, V3 s)
{
    float x2 = xmul(r.x, r.x), y2 = xmul(r.y, r.y), z2 = xmul(r.z, r.z);
    float xz = xmul(r.x, r.z), xy = xmul(r.x, r.y), yz = xmul(r.y, r.z);
    float wx = xmul(r.w, r.x), wy = xmul(r.w, r.y), wz = xmul(r.w, r.z);
    V3 ds = mk3(xmul(s.x, 2.f), xmul(s.y, 2.f), xmul(s.z, 2.f));
    M3 m;
    m.c[0] = mk3(xsub(s.x, xmul(ds.x, xadd(y2, z2))), xmul(ds.x, xadd(xy, wz)), xmul(ds.x, xsub(xz, wy)));
    m.c[1] = mk3(xmul(ds.y, xsub(xy, wz)), xsub(s.y, xmul(ds.y, xadd(x2, z2))), xmul(ds.y, xadd(yz, wx)));
    m.c[2] = mk3(xmul(ds.z, xadd(xz, wy)), xmul(ds.z, xsub(yz, wx)), xsub(s.z, xmul(ds.z, xadd(x2, y2))));
    return m;
}
MDR_HD V3 matMul(const M3 &m, V3 v) { return m.c[0] * v.x + m.c[1] * v.y + m.c[2] * v.z; }
MDR_HD M3 matMulDiag(const M3 &m, V3 d)
{
    M3 r;
    r.c[0] = m.c[0] * d.x; r.c[1] = m.c[1] * d.y; r.c[2] = m.c[2] * d.z;
    return r;
}

// math.hpp:999-1007 (strict inequalities)
MDR_HD bool overlaps(const Box &a, const Box &b)
{
    return a.lo.x < b.hi.x && b.lo.x < a.hi.x &&
           a.lo.y < b.hi.y && b.lo.y < a.hi.y &&
           a.lo.z < b.hi.z && b.lo.z < a.hi.z;
}
// math.hpp:1071-1103: AABB of a transformed AABB, column by column.
MDR_HD Box applyTRS(const Box &b, V3 t, Q4 r, V3 s)
{
    M3 m = matFromRS(r, s);
    Box o;
    float lo[3], hi[3];
    const float tt[3] = { t.x, t.y, t.z };
    const float bl[3] = { b.lo.x, b.lo.y, b.lo.z };
    const float bh[3] = { b.hi.x, b.hi.y, b.hi.z };
#pragma unroll
    for (int i = 0; i < 3; i++) {
        lo[i] = hi[i] = tt[i];
#pragma unroll
        for (int j = 0; j < 3; j++) {
            float mji = comp(m.c[j], i);
            float e = xmul(mji, bl[j]);
            float f = xmul(mji, bh[j]);
            if (e < f) { lo[i] = xadd(lo[i], e); hi[i] = xadd(hi[i], f); }
            else { lo[i] = xadd(lo[i], f); hi[i] = xadd(hi[i], e); }
        }
    }
    o.lo = mk3(lo[0], lo[1], lo[2]);
    o.hi = mk3(hi[0], hi[1], hi[2]);
    return o;
}

// src/physics/broadphase.cpp:435-459
MDR_HD Box expandWithMotion(Box b, V3 v, float vf, float af)
{
    float lo[3] = { b.lo.x, b.lo.y, b.lo.z }, hi[3] = { b.hi.x, b.hi.y, b.hi.z };
    const float vv[3] = { v.x, v.y, v.z };
#pragma unroll
    for (int i = 0; i < 3; i++) {
        float pos_delta = xmul(vf, vv[i]);
        float min_delta = xsub(pos_delta, af);
        float max_delta = xadd(pos_delta, af);
        if (min_delta < 0.f) lo[i] = xadd(lo[i], min_delta);
        if (max_delta > 0.f) hi[i] = xadd(hi[i], max_delta);
    }
    b.lo = mk3(lo[0], lo[1], lo[2]);
    b.hi = mk3(hi[0], hi[1], hi[2]);
    return b;
}

}  // namespace mdr
