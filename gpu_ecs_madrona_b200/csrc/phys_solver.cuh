// XPBD substep: integrate, positional contact solve, velocity derivation,
// restitution + friction velocity solve (src/physics/physics.cpp:79-164,
// 166-476, 673-714, 716-1008), one thread per body / per contact.
//
// Exact shortcuts for a STABLE STATIC body s (inverse mass and inertia zero, no
// -0 component in its pose, quaternion a fixed point of normalize; finite data):
//   * its generalized inverse mass w = 0 + dot(ta, 0*ta) is +0, so
//     w1 + w2 (+ 0) is the other body's w bit for bit;
//   * the positional and velocity updates add +-0 to its state and renormalize
//     a fixed point: the state does not change, so it is neither computed nor
//     written (which is also what lets contacts on one static body commute);
//   * its world contact point p = rotate(q, r) + x is constant through the
//     solve and equals its previous-pose point, so (p - p_hat) is the +0 vector;
//   * its velocity and pre-solve velocity are +0, so v + cross(omega, r) is +0.
// The expressions below keep the reference's operand order with those zeros
// substituted (e.g. `zero - x`, not `-x`), so signs of zero come out the same.
#pragma once
#include "phys_narrow.cuh"

namespace mdr {

struct BodyMass { float inv_m; V3 inv_I; float mu_s, mu_d; };
// Static bodies take zero inverse mass / inertia (physics.cpp:404-416,889-901).
MDR_HD BodyMass bodyMass(const WorldView &v, const StepParams &p, int row)
{
    const uint32_t meta = v.meta[row];
    const ObjDev *o = p.a.objs + (meta & 0xffffu);
    BodyMass m;
    m.inv_m = MDR_LDG(&o->inv_mass);
    m.inv_I = mk3(MDR_LDG(&o->inv_inertia[0]), MDR_LDG(&o->inv_inertia[1]), MDR_LDG(&o->inv_inertia[2]));
    m.mu_s = MDR_LDG(&o->mu_s);
    m.mu_d = MDR_LDG(&o->mu_d);
    if (meta & META_STATIC) { m.inv_m = 0.f; m.inv_I = mk3(0.f, 0.f, 0.f); }
    return m;
}

// solver::substepRigidBodies, physics.cpp:79-164. External force / torque are
// read straight from HBM (read-only, once per substep); PreSolvePositional is
// only an output of the step (the solver keeps local arms instead) and goes
// straight to HBM.
MDR_HD void substepBodies(const BlockCtx &c, WorldView &v, const StepParams &p, int w, int sub)
{
    const float h = p.h;
    const V3 g = mk3(p.g[0], p.g[1], p.g[2]);
    const size_t r0 = (size_t)w * p.t.stride;
    const int rows = sub > 0 ? v.misc[3] : v.n;
    for (int k = c.tid(); k < rows; k += c.nthreads()) {
        const int i = v.row_perm[k];
        V3 x = ld3(v.pos, i);
        Q4 q = ld4(v.rot, i);
        V3 vl = ldVelLin(v.vel, i), omega = ldVelAng(v.vel, i);
        stPose(v.prev, i, x, q);
        if (v.meta[i] & META_STATIC) {
            stPose(p.t.prex, r0 + i, x, q);
            stVel(v.prev_vel, i, mk3(0.f, 0.f, 0.f), mk3(0.f, 0.f, 0.f));
            continue;
        }
        const ObjDev *o = p.a.objs + (v.meta[i] & 0xffffu);
        const float inv_m = MDR_LDG(&o->inv_mass);
        const V3 inv_I = mk3(MDR_LDG(&o->inv_inertia[0]), MDR_LDG(&o->inv_inertia[1]), MDR_LDG(&o->inv_inertia[2]));
        const float *pf = p.t.ext_f + 3 * (r0 + i), *pt = p.t.ext_t + 3 * (r0 + i);
        const V3 ext_f = mk3(MDR_LDG(pf), MDR_LDG(pf + 1), MDR_LDG(pf + 2));
        const V3 ext_t = mk3(MDR_LDG(pt), MDR_LDG(pt + 1), MDR_LDG(pt + 2));
        if (MDR_LDG(p.t.resp + r0 + i) == 0u) vl = vl + h * g;
        vl = vl + (h * inv_m) * ext_f;
        x = x + h * vl;
        // inv_I == 0 ? 0 : 1 / inv_I per axis (physics.cpp:134-138), divided once per object on the host (ObjDev)
        V3 I = mk3(MDR_LDG(&o->inertia[0]), MDR_LDG(&o->inertia[1]), MDR_LDG(&o->inertia[2]));
        Q4 to_local = qinv(q);
        V3 tau_local = rotate(to_local, ext_t);
        V3 omega_local = rotate(to_local, omega);
        V3 I_omega = mulDiag(I, omega_local);
        omega_local = omega_local + h * mulDiag(inv_I, tau_local - cross(omega_local, I_omega));
        omega = rotate(q, omega_local);
        Q4 apply = fromAngular((0.5f * h) * omega);
        q = q + apply * q;
        q = qnormalize(q);
        st3(v.pos, i, x);
        st4(v.rot, i, q);
        stPose(p.t.prex, r0 + i, x, q);
        stVel(v.prev_vel, i, vl, omega);
    }
}

// One side of applyPositionalUpdate, physics.cpp:185-211. sign = +1 for the
// first body (x += ..., q += ...), -1 for the second (x -= ..., q -= ...).
template <int kSign>
MDR_HD void applyPositionalSide(V3 &x, Q4 &q, V3 ra, float inv_m, V3 n, float dl)
{
    V3 dx = (dl * inv_m) * n;
    x = kSign > 0 ? x + dx : x - dx;
    float half = 0.5f * dl;
    V3 a = rotate(q, half * ra);
    Q4 dq = fromAngular(a) * q;
    q = kSign > 0 ? q + dq : q - dq;
    q = qnormalize(q);
}

// Generalized inverse mass of one body along a direction (physics.cpp:166-183):
// inv_m + dot(ta, inv_I ta); ra is returned for the update.
MDR_HD float generalizedInvMass(const BodyMass &m, V3 ta, V3 &ra)
{
    ra = mulDiag(m.inv_I, ta);
    return m.inv_m + dot(ta, ra);
}

MDR_HD V3 ldArm(const float (*r)[3], int i) { return mk3(r[i][0], r[i][1], r[i][2]); }

// handleContact + handleContactConstraint, physics.cpp:281-363,387-476.
// kS1 / kS2: body 1 / body 2 is a stable static (compile-time specialization:
// the skipped arithmetic must not even be issued).
template <bool kS1, bool kS2>
MDR_HD void solveContactPositionsT(WorldView &v, const StepParams &p, SolverContact *ct)
{
    const uint32_t hdr = ct->hdr;
    const int a = hdr & 1023, b = (hdr >> 10) & 1023, npts = (hdr >> 20) & 7;
    constexpr bool s1 = kS1, s2 = kS2;
    const V3 nrm = mk3(ct->normal[0], ct->normal[1], ct->normal[2]);
    const V3 zero = mk3(0.f, 0.f, 0.f);
    V3 x1 = ld3(v.pos, a), x2 = ld3(v.pos, b);
    Q4 q1 = ld4(v.rot, a), q2 = ld4(v.rot, b);
    const BodyMass m1 = bodyMass(v, p, a), m2 = bodyMass(v, p, b);
    const float avg_mu_s = 0.5f * (m1.mu_s + m2.mu_s);
    const V3 prev_x1 = ldPoseX(v.prev, a), prev_x2 = ldPoseX(v.prev, b);
    const Q4 prev_q1 = ldPoseQ(v.prev, a), prev_q2 = ldPoseQ(v.prev, b);

    for (int i = 0; i < npts; i++) {
        const V3 r1 = ldArm(ct->r1, i), r2 = ldArm(ct->r2, i);   // world point for a stable static
        float lambda_n = 0.f;
        V3 p1 = s1 ? r1 : rotate(q1, r1) + x1;
        V3 p2 = s2 ? r2 : rotate(q2, r2) + x2;
        float d = dot(p1 - p2, nrm);
        if (d > 0) {
            {
                V3 ra1 = zero, ra2 = zero;
                float w1 = 0.f, w2 = 0.f;
                if (!s1) w1 = generalizedInvMass(m1, cross(r1, rotate(qinv(q1), nrm)), ra1);
                if (!s2) w2 = generalizedInvMass(m2, cross(r2, rotate(qinv(q2), nrm)), ra2);
                lambda_n = -d / (w1 + w2 + 0.f);
                if (!s1) applyPositionalSide<1>(x1, q1, ra1, m1.inv_m, nrm, lambda_n);
                if (!s2) applyPositionalSide<-1>(x2, q2, ra2, m2.inv_m, nrm, lambda_n);
            }
            V3 d1 = zero, d2 = zero;
            if (!s1) d1 = (rotate(q1, r1) + x1) - (rotate(prev_q1, r1) + prev_x1);
            if (!s2) d2 = (rotate(q2, r2) + x2) - (rotate(prev_q2, r2) + prev_x2);
            V3 delta_p = d1 - d2;
            V3 delta_p_t = delta_p - dot(delta_p, nrm) * nrm;
            float tang = len(delta_p_t);
            if (tang > 0.f) {
                V3 t = delta_p_t / tang;
                V3 fra1 = zero, fra2 = zero;
                float w1 = 0.f, w2 = 0.f;
                if (!s1) w1 = generalizedInvMass(m1, cross(r1, rotate(qinv(q1), t)), fra1);
                if (!s2) w2 = generalizedInvMass(m2, cross(r2, rotate(qinv(q2), t)), fra2);
                float lambda_t = -tang / (w1 + w2 + 0.f);
                float threshold = lambda_n * avg_mu_s;
                if (lambda_t > threshold) {
                    if (!s1) applyPositionalSide<1>(x1, q1, fra1, m1.inv_m, t, lambda_t);
                    if (!s2) applyPositionalSide<-1>(x2, q2, fra2, m2.inv_m, t, lambda_t);
                }
            }
        }
        ct->lambda[i] = lambda_n;
    }
    // Written back unless the body does not order contacts (stable static:
    // unchanged, and possibly shared by other contacts of this level).
    if (!s1) { st3(v.pos, a, x1); st4(v.rot, a, q1); }
    if (!s2) { st3(v.pos, b, x2); st4(v.rot, b, q2); }
}

MDR_HD void solveContactPositions(WorldView &v, const StepParams &p, SolverContact *ct)
{
    const uint32_t hdr = ct->hdr;
    if ((hdr >> 24) & 1) solveContactPositionsT<true, false>(v, p, ct);
    else if ((hdr >> 25) & 1) solveContactPositionsT<false, true>(v, p, ct);
    else solveContactPositionsT<false, false>(v, p, ct);
}

// ------------------------------------------------------------- joints -------
// computeAngularUpdate + applyAngularUpdate (physics.cpp:247-279) behind the
// part applyJointOrientationConstraint and applyJointAxisConstraint (:478-535)
// share: normalise the error vector, take it to both body frames, rotate both.
MDR_HD void jointAngularConstraint(Q4 &q1, Q4 &q2, V3 delta_q, V3 inv_I1, V3 inv_I2)
{
    const float theta = len(delta_q);
    if (theta > 0) {
        delta_q = delta_q / theta;
        const V3 n1 = rotate(qinv(q1), delta_q), n2 = rotate(qinv(q2), delta_q);
        const V3 ra1 = mulDiag(inv_I1, n1), ra2 = mulDiag(inv_I2, n2);
        const float w1 = dot(n1, ra1), w2 = dot(n2, ra2);
        const float half = 0.5f * (-theta / (w1 + w2 + 0.f));
        const Q4 u1 = fromAngular(rotate(q1, half * ra1));
        const Q4 u2 = fromAngular(rotate(q2, half * ra2));
        q1 = qnormalize(q1 + u1 * q1);
        q2 = qnormalize(q2 - u2 * q2);
    }
}

// handleJointConstraint, physics.cpp:537-648. Static bodies take part with zero
// inverse mass / inertia and are written back like the reference does (a stable
// static's pose comes back bit for bit: +-0 added, a fixed point renormalised).
#if defined(__CUDA_ARCH__)
__device__ __noinline__
#else
inline
#endif
void solveJoint(WorldView &v, const StepParams &p, const JointDev *j)
{
    const int a = MDR_LDG(&j->row1), b = MDR_LDG(&j->row2);
    V3 x1 = ld3(v.pos, a), x2 = ld3(v.pos, b);
    Q4 q1 = ld4(v.rot, a), q2 = ld4(v.rot, b);
    const BodyMass m1 = bodyMass(v, p, a), m2 = bodyMass(v, p, b);
    const V3 r1 = mk3(MDR_LDG(&j->r1[0]), MDR_LDG(&j->r1[1]), MDR_LDG(&j->r1[2]));
    const V3 r2 = mk3(MDR_LDG(&j->r2[0]), MDR_LDG(&j->r2[1]), MDR_LDG(&j->r2[2]));
    float d[12];
#pragma unroll
    for (int k = 0; k < 12; k++) d[k] = MDR_LDG(&j->data[k]);
    V3 corr;
    if (MDR_LDG(&j->type) == 0) {
        const Q4 at1 = mk4(d[0], d[1], d[2], d[3]), at2 = mk4(d[4], d[5], d[6], d[7]);
        const Q4 o1 = qnormalize(q1 * at1), o2 = qnormalize(q2 * at2);
        const Q4 diff = o1 * qinv(o2);
        jointAngularConstraint(q1, q2, 2.f * mk3(diff.x, diff.y, diff.z), m1.inv_I, m2.inv_I);
        const V3 delta_r = (rotate(q2, r2) + x2) - (rotate(q1, r1) + x1);
        const Q4 axes = qnormalize(q1 * at1);
        const V3 a1 = rotate(axes, mk3(0.f, 1.f, 0.f));      // math::fwd
        const V3 b1 = rotate(axes, mk3(1.f, 0.f, 0.f));      // math::right
        const V3 c1 = cross(a1, b1);
        corr = mk3(0.f, 0.f, 0.f);
        corr = corr - (dot(delta_r, a1) - d[8]) * a1;
        corr = corr - dot(delta_r, b1) * b1;
        corr = corr - dot(delta_r, c1) * c1;
    } else {
        const V3 ax1 = rotate(q1, mk3(d[0], d[1], d[2])), ax2 = rotate(q2, mk3(d[3], d[4], d[5]));
        jointAngularConstraint(q1, q2, cross(ax1, ax2), m1.inv_I, m2.inv_I);
        corr = (rotate(q2, r2) + x2) - (rotate(q1, r1) + x1);
    }
    const float mag = len(corr);
    if (mag > 0.f) {
        corr = corr / mag;
        V3 ra1, ra2;
        const float w1 = generalizedInvMass(m1, cross(r1, rotate(qinv(q1), corr)), ra1);
        const float w2 = generalizedInvMass(m2, cross(r2, rotate(qinv(q2), corr)), ra2);
        const float lambda = -mag / (w1 + w2 + 0.f);
        applyPositionalSide<1>(x1, q1, ra1, m1.inv_m, corr, lambda);
        applyPositionalSide<-1>(x2, q2, ra2, m2.inv_m, corr, lambda);
    }
    st3(v.pos, a, x1); st3(v.pos, b, x2);
    st4(v.rot, a, q1); st4(v.rot, b, q2);
}

// The joints of one world, after the contacts of the substep, one after the other
// in table order (solvePositions, physics.cpp:663-668). `joint_stride` is uniform
// over the grid: a batch without joints pays one predictable branch.
MDR_HD void solveJoints(const BlockCtx &c, WorldView &v, const StepParams &p, int w)
{
    if (p.joint_stride == 0) return;
    if (c.tid() == 0 && v.active) {
        const int nj = MDR_LDG(p.joint_count + w);
        const JointDev *js = p.joints + (size_t)w * p.joint_stride;
        for (int k = 0; k < nj; k++) solveJoint(v, p, js + k);
    }
    c.sync();
}

// solver::setVelocities, physics.cpp:673-714
MDR_HD void setVelocities(const BlockCtx &c, WorldView &v, const StepParams &p, int sub)
{
    const float h = p.h;
    const int rows = sub > 0 ? v.misc[3] : v.n;
    for (int k = c.tid(); k < rows; k += c.nthreads()) {
        const int i = v.row_perm[k];
        Q4 q = ld4(v.rot, i), qp = ldPoseQ(v.prev, i);
        Q4 dq;
        if (q.w != qp.w || q.x != qp.x || q.y != qp.y || q.z != qp.z) dq = q * qinv(qp);
        else dq = mk4(1.f, 0.f, 0.f, 0.f);
        V3 new_omega = (2.f / h) * mk3(dq.x, dq.y, dq.z);
        V3 lin = (ld3(v.pos, i) - ldPoseX(v.prev, i)) / h;
        stVel(v.vel, i, lin, dq.w > 0.f ? new_omega : -new_omega);
    }
}

// One side of applyVelocityUpdate (physics.cpp:730-750) once the magnitude has
// been divided by the generalized inverse masses.
template <int kSign>
MDR_HD void applyVelSide(V3 &vl, V3 &om, Q4 q, V3 ra, float inv_m, V3 dv, float mag)
{
    V3 dvl = (mag * inv_m) * dv;
    V3 dom = rotate(q, mag * ra);
    vl = kSign > 0 ? vl + dvl : vl - dvl;
    om = kSign > 0 ? om + dom : om - dom;
}

// Relative velocity of the two contact points (physics.cpp:716-728); a stable
// static body's term is the zero vector.
template <bool kS1, bool kS2>
MDR_HD V3 relVelAt(V3 v1, V3 o1, V3 r1w, V3 v2, V3 o2, V3 r2w)
{
    V3 va = mk3(0.f, 0.f, 0.f), vb = mk3(0.f, 0.f, 0.f);
    if (!kS1) va = v1 + cross(o1, r1w);
    if (!kS2) vb = v2 + cross(o2, r2w);
    return va - vb;
}

// solveVelocitiesForContact, physics.cpp:865-993 (+ :752-863).
// The world arm of a point, rotate(q, r), is the same in all three sweeps (the velocity solve does not move poses).
// When the other side is a stable static, that side's slot in the contact (its world point, last read by the
// positional solve; the narrowphase rewrites the contact next substep) is free: the first sweep leaves the arm
// there and the other two read it back -- the same value, 3 loads instead of 30 instructions.
template <bool kS1, bool kS2>
MDR_HD void solveContactVelocitiesT(WorldView &v, const StepParams &p, SolverContact *ct)
{
    const float h = p.h;
    const uint32_t hdr = ct->hdr;
    const int a = hdr & 1023, b = (hdr >> 10) & 1023, npts = (hdr >> 20) & 7;
    constexpr bool s1 = kS1, s2 = kS2;
    const V3 nrm = mk3(ct->normal[0], ct->normal[1], ct->normal[2]);
    const V3 zero = mk3(0.f, 0.f, 0.f);
    V3 v1 = ldVelLin(v.vel, a), o1 = ldVelAng(v.vel, a);
    V3 v2 = ldVelLin(v.vel, b), o2 = ldVelAng(v.vel, b);
    const Q4 q1 = ld4(v.rot, a), q2 = ld4(v.rot, b);
    const BodyMass m1 = bodyMass(v, p, a), m2 = bodyMass(v, p, b);
    const float mu_d = 0.5f * (m1.mu_d + m2.mu_d);
    V3 n1 = zero, n2 = zero;
    if (!s1) n1 = rotate(qinv(q1), nrm);
    if (!s2) n2 = rotate(qinv(q2), nrm);

    // restitution: two sweeps over the points, then dynamic friction: one sweep
    for (int pass = 0; pass < 3; pass++) {
        for (int i = 0; i < npts; i++) {
            // dynamic friction of a point that took no normal impulse: the magnitude is zero and the reference leaves
            // the point alone after computing a tangential velocity it then does not use (physics.cpp:775-787);
            // here the test comes first (a NaN magnitude is != 0 on both sides and goes on)
            float fric = 0.f;
            if (pass == 2) {
                fric = mu_d * fabsf(ct->lambda[i]) / h;
                if (fric == 0.f) continue;
            }
            const V3 r1 = ldArm(ct->r1, i), r2 = ldArm(ct->r2, i);
            V3 r1w = zero, r2w = zero;
            if (!s1 && s2) {
                if (pass == 0) { r1w = rotate(q1, r1); ct->r2[i][0] = r1w.x; ct->r2[i][1] = r1w.y; ct->r2[i][2] = r1w.z; }
                else r1w = r2;
            } else if (!s2 && s1) {
                if (pass == 0) { r2w = rotate(q2, r2); ct->r1[i][0] = r2w.x; ct->r1[i][1] = r2w.y; ct->r1[i][2] = r2w.z; }
                else r2w = r1;
            } else {
                if (!s1) r1w = rotate(q1, r1);
                if (!s2) r2w = rotate(q2, r2);
            }
            V3 vr = relVelAt<kS1, kS2>(v1, o1, r1w, v2, o2, r2w);
            float vn = dot(nrm, vr);
            V3 dir, ta1 = zero, ta2 = zero;
            float mag;
            if (pass < 2) {
                const float vn_bar = ct->vn_bar[i];
                float e = 0.3f;
                if (fabsf(vn_bar) <= p.restitution_threshold) e = 0.f;
                mag = fminRef(-e * vn_bar, 0.f) - vn;
                dir = nrm;
                if (!s1) ta1 = cross(r1, n1);
                if (!s2) ta2 = cross(r2, n2);
            } else {
                V3 vt = vr - nrm * vn;
                float vt_len = len(vt);
                if (!(vt_len != 0 && fric != 0.f)) continue;
                mag = -fminRef(fric, vt_len);
                dir = vt / vt_len;
                if (!s1) ta1 = cross(r1, rotate(qinv(q1), dir));
                if (!s2) ta2 = cross(r2, rotate(qinv(q2), dir));
            }
            V3 ra1 = zero, ra2 = zero;
            float w1 = 0.f, w2 = 0.f;
            if (!s1) w1 = generalizedInvMass(m1, ta1, ra1);
            if (!s2) w2 = generalizedInvMass(m2, ta2, ra2);
            mag *= 1.f / (w1 + w2);
            if (!s1) applyVelSide<1>(v1, o1, q1, ra1, m1.inv_m, dir, mag);
            if (!s2) applyVelSide<-1>(v2, o2, q2, ra2, m2.inv_m, dir, mag);
        }
    }
    if (!s1) stVel(v.vel, a, v1, o1);
    if (!s2) stVel(v.vel, b, v2, o2);
}

MDR_HD void solveContactVelocities(WorldView &v, const StepParams &p, SolverContact *ct)
{
    const uint32_t hdr = ct->hdr;
    if ((hdr >> 24) & 1) solveContactVelocitiesT<true, false>(v, p, ct);
    else if ((hdr >> 25) & 1) solveContactVelocitiesT<false, true>(v, p, ct);
    else solveContactVelocitiesT<false, false>(v, p, ct);
}

// One Gauss-Seidel sweep in the reference's contact order: level by level,
// the contacts of one level in parallel. `max_level` is uniform over the block.
template <bool kVelocity>
MDR_HD void solveLevels(const BlockCtx &c, WorldView &v, const StepParams &p, int max_level)
{
    const int np = v.misc[0];
    for (int lv = 1; lv <= max_level; lv++) {
        for (int i = c.tid(); i < np; i += c.nthreads()) {
            if (pairLevel(v.pairs[i]) != lv) continue;
            SolverContact *ct = contactSlot(v, i);
            if (ct->hdr == 0) continue;
            if (kVelocity) solveContactVelocities(v, p, ct);
            else solveContactPositions(v, p, ct);
        }
        c.sync();
    }
}

}  // namespace mdr
