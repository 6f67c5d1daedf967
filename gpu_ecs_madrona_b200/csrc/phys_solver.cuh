// XPBD substep: integrate, positional contact solve, velocity derivation,
// restitution + friction velocity solve (src/physics/physics.cpp:79-164,
// 166-476, 673-714, 716-1008), one thread per body / per contact.
#pragma once
#include "phys_world.cuh"

namespace mdr {

MDR_HD ContactDev *contactSlot(const WorldView &v, int p)
{
    return p < v.smem_contacts ? v.contacts + p : v.overflow + (p - v.smem_contacts);
}

struct BodyMass { float inv_m; V3 inv_I; float mu_s, mu_d; };
// Static bodies take zero inverse mass / inertia (physics.cpp:404-416,889-901).
MDR_HD BodyMass bodyMass(const WorldView &v, const StepParams &p, int row)
{
    const ObjDev *o = p.a.objs + v.obj[row];
    BodyMass m;
    m.inv_m = MDR_LDG(&o->inv_mass);
    m.inv_I = mk3(MDR_LDG(&o->inv_inertia[0]), MDR_LDG(&o->inv_inertia[1]), MDR_LDG(&o->inv_inertia[2]));
    m.mu_s = MDR_LDG(&o->mu_s);
    m.mu_d = MDR_LDG(&o->mu_d);
    if (v.resp[row] == 2u) { m.inv_m = 0.f; m.inv_I = mk3(0.f, 0.f, 0.f); }
    return m;
}

// solver::substepRigidBodies, physics.cpp:79-164. External force / torque are
// read straight from HBM (read-only, once per substep).
MDR_HD void substepBodies(const BlockCtx &c, WorldView &v, const StepParams &p, int w)
{
    const float h = p.h;
    const V3 g = mk3(p.g[0], p.g[1], p.g[2]);
    const size_t r0 = (size_t)w * p.t.stride;
    for (int i = c.tid(); i < v.n; i += c.nthreads()) {
        V3 x = ld3(v.pos, i);
        Q4 q = ld4(v.rot, i);
        V3 vl = ldVelLin(v.vel, i), omega = ldVelAng(v.vel, i);
        const uint32_t rt = v.resp[i];
        stPose(v.prev, i, x, q);
        if (rt == 2u) {
            stPose(v.prex, i, x, q);
            stVel(v.prev_vel, i, mk3(0.f, 0.f, 0.f), mk3(0.f, 0.f, 0.f));
            continue;
        }
        const ObjDev *o = p.a.objs + v.obj[i];
        const float inv_m = MDR_LDG(&o->inv_mass);
        const V3 inv_I = mk3(MDR_LDG(&o->inv_inertia[0]), MDR_LDG(&o->inv_inertia[1]), MDR_LDG(&o->inv_inertia[2]));
        const V3 ext_f = mk3(MDR_LDG(p.t.ext_f + 3 * (r0 + i)), MDR_LDG(p.t.ext_f + 3 * (r0 + i) + 1), MDR_LDG(p.t.ext_f + 3 * (r0 + i) + 2));
        const V3 ext_t = mk3(MDR_LDG(p.t.ext_t + 3 * (r0 + i)), MDR_LDG(p.t.ext_t + 3 * (r0 + i) + 1), MDR_LDG(p.t.ext_t + 3 * (r0 + i) + 2));
        if (rt == 0u) vl = vl + h * g;
        vl = vl + (h * inv_m) * ext_f;
        x = x + h * vl;
        V3 I = mk3(inv_I.x == 0 ? 0.0f : 1.0f / inv_I.x,
                   inv_I.y == 0 ? 0.0f : 1.0f / inv_I.y,
                   inv_I.z == 0 ? 0.0f : 1.0f / inv_I.z);
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
        stPose(v.prex, i, x, q);
        stVel(v.prev_vel, i, vl, omega);
    }
}

// physics.cpp:185-211
MDR_HD void applyPositional(V3 &x1, V3 &x2, Q4 &q1, Q4 &q2, V3 ra1, V3 ra2,
                            float inv_m1, float inv_m2, V3 n, float dl)
{
    x1 = x1 + (dl * inv_m1) * n;
    x2 = x2 - (dl * inv_m2) * n;
    float half = 0.5f * dl;
    V3 a1 = rotate(q1, half * ra1);
    V3 a2 = rotate(q2, half * ra2);
    q1 = q1 + fromAngular(a1) * q1;
    q2 = q2 - fromAngular(a2) * q2;
    q1 = qnormalize(q1);
    q2 = qnormalize(q2);
}

// physics.cpp:166-183
MDR_HD float positionalLambda(V3 ta1, V3 ta2, V3 ra1, V3 ra2,
                              float inv_m1, float inv_m2, float cval, float alpha)
{
    float w1 = inv_m1 + dot(ta1, ra1);
    float w2 = inv_m2 + dot(ta2, ra2);
    return -cval / (w1 + w2 + alpha);
}

// physics.cpp:365-382
MDR_HD void localContacts(V3 pre_x1, Q4 pre_q1, V3 pre_x2, Q4 pre_q2,
                          const ContactDev *ct, V3 nrm, int i, V3 &r1, V3 &r2)
{
    V3 c1 = mk3(ct->pts[i][0], ct->pts[i][1], ct->pts[i][2]);
    float depth = ct->pts[i][3];
    V3 c2 = c1 - nrm * depth;
    r1 = rotate(qinv(pre_q1), c1 - pre_x1);
    r2 = rotate(qinv(pre_q2), c2 - pre_x2);
}

// handleContact + handleContactConstraint, physics.cpp:281-363,387-476.
MDR_HD void solveContactPositions(WorldView &v, const StepParams &p, ContactDev *ct)
{
    const int a = ct->ref, b = ct->alt;
    const int npts = ct->num_points;
    const V3 nrm = mk3(ct->normal[0], ct->normal[1], ct->normal[2]);
    V3 x1 = ld3(v.pos, a), x2 = ld3(v.pos, b);
    Q4 q1 = ld4(v.rot, a), q2 = ld4(v.rot, b);
    const BodyMass m1 = bodyMass(v, p, a), m2 = bodyMass(v, p, b);
    const float avg_mu_s = 0.5f * (m1.mu_s + m2.mu_s);
    const V3 pre_x1 = ldPoseX(v.prex, a), pre_x2 = ldPoseX(v.prex, b);
    const Q4 pre_q1 = ldPoseQ(v.prex, a), pre_q2 = ldPoseQ(v.prex, b);
    const V3 prev_x1 = ldPoseX(v.prev, a), prev_x2 = ldPoseX(v.prev, b);
    const Q4 prev_q1 = ldPoseQ(v.prev, a), prev_q2 = ldPoseQ(v.prev, b);

    for (int i = 0; i < npts; i++) {
        V3 r1, r2;
        localContacts(pre_x1, pre_q1, pre_x2, pre_q2, ct, nrm, i, r1, r2);
        float lambda_n = 0.f;
        V3 p1 = rotate(q1, r1) + x1;
        V3 p2 = rotate(q2, r2) + x2;
        float d = dot(p1 - p2, nrm);
        if (d > 0) {
            {
                V3 n1 = rotate(qinv(q1), nrm), n2 = rotate(qinv(q2), nrm);
                V3 ta1 = cross(r1, n1), ta2 = cross(r2, n2);
                V3 ra1 = mulDiag(m1.inv_I, ta1), ra2 = mulDiag(m2.inv_I, ta2);
                lambda_n = positionalLambda(ta1, ta2, ra1, ra2, m1.inv_m, m2.inv_m, d, 0.f);
                applyPositional(x1, x2, q1, q2, ra1, ra2, m1.inv_m, m2.inv_m, nrm, lambda_n);
            }
            V3 p1_hat = rotate(prev_q1, r1) + prev_x1;
            V3 p2_hat = rotate(prev_q2, r2) + prev_x2;
            p1 = rotate(q1, r1) + x1;
            p2 = rotate(q2, r2) + x2;
            V3 delta_p = (p1 - p1_hat) - (p2 - p2_hat);
            V3 delta_p_t = delta_p - dot(delta_p, nrm) * nrm;
            float tang = len(delta_p_t);
            if (tang > 0.f) {
                V3 t = delta_p_t / tang;
                V3 t1 = rotate(qinv(q1), t), t2 = rotate(qinv(q2), t);
                V3 fta1 = cross(r1, t1), fta2 = cross(r2, t2);
                V3 fra1 = mulDiag(m1.inv_I, fta1), fra2 = mulDiag(m2.inv_I, fta2);
                float lambda_t = positionalLambda(fta1, fta2, fra1, fra2, m1.inv_m, m2.inv_m, tang, 0.f);
                float threshold = lambda_n * avg_mu_s;
                if (lambda_t > threshold) {
                    applyPositional(x1, x2, q1, q2, fra1, fra2, m1.inv_m, m2.inv_m, t, lambda_t);
                }
            }
        }
        ct->lambda_n[i] = lambda_n;
    }
    // A stable static body is written back unchanged by the reference; several
    // contacts of one level may share it, so it is not written here.
    if (v.dep[a]) { st3(v.pos, a, x1); st4(v.rot, a, q1); }
    if (v.dep[b]) { st3(v.pos, b, x2); st4(v.rot, b, q2); }
}

// solver::setVelocities, physics.cpp:673-714
MDR_HD void setVelocities(const BlockCtx &c, WorldView &v, const StepParams &p)
{
    const float h = p.h;
    for (int i = c.tid(); i < v.n; i += c.nthreads()) {
        Q4 q = ld4(v.rot, i), qp = ldPoseQ(v.prev, i);
        Q4 dq;
        if (q.w != qp.w || q.x != qp.x || q.y != qp.y || q.z != qp.z) dq = q * qinv(qp);
        else dq = mk4(1.f, 0.f, 0.f, 0.f);
        V3 new_omega = (2.f / h) * mk3(dq.x, dq.y, dq.z);
        V3 lin = (ld3(v.pos, i) - ldPoseX(v.prev, i)) / h;
        stVel(v.vel, i, lin, dq.w > 0.f ? new_omega : -new_omega);
    }
}

// physics.cpp:716-750
MDR_HD V3 relVel(V3 v1, V3 v2, V3 o1, V3 o2, V3 d1, V3 d2)
{
    return (v1 + cross(o1, d1)) - (v2 + cross(o2, d2));
}
MDR_HD void applyVelUpdate(V3 &v1, V3 &v2, V3 &o1, V3 &o2, Q4 q1, Q4 q2, V3 ta1, V3 ta2,
                           const BodyMass &m1, const BodyMass &m2, V3 dv, float mag)
{
    V3 ra1 = mulDiag(m1.inv_I, ta1), ra2 = mulDiag(m2.inv_I, ta2);
    float w1 = m1.inv_m + dot(ta1, ra1);
    float w2 = m2.inv_m + dot(ta2, ra2);
    mag *= 1.f / (w1 + w2);
    v1 = v1 + (mag * m1.inv_m) * dv;
    v2 = v2 - (mag * m2.inv_m) * dv;
    o1 = o1 + rotate(q1, mag * ra1);
    o2 = o2 - rotate(q2, mag * ra2);
}

// solveVelocitiesForContact, physics.cpp:865-993 (+ :752-863).
MDR_HD void solveContactVelocities(WorldView &v, const StepParams &p, const ContactDev *ct)
{
    const float h = p.h;
    const int a = ct->ref, b = ct->alt;
    const int npts = ct->num_points;
    const V3 nrm = mk3(ct->normal[0], ct->normal[1], ct->normal[2]);
    V3 v1 = ldVelLin(v.vel, a), o1 = ldVelAng(v.vel, a);
    V3 v2 = ldVelLin(v.vel, b), o2 = ldVelAng(v.vel, b);
    const Q4 q1 = ld4(v.rot, a), q2 = ld4(v.rot, b);
    const BodyMass m1 = bodyMass(v, p, a), m2 = bodyMass(v, p, b);
    const float mu_d = 0.5f * (m1.mu_d + m2.mu_d);
    const V3 pre_x1 = ldPoseX(v.prex, a), pre_x2 = ldPoseX(v.prex, b);
    const Q4 pre_q1 = ldPoseQ(v.prex, a), pre_q2 = ldPoseQ(v.prex, b);
    const V3 pv1 = ldVelLin(v.prev_vel, a), po1 = ldVelAng(v.prev_vel, a);
    const V3 pv2 = ldVelLin(v.prev_vel, b), po2 = ldVelAng(v.prev_vel, b);
    const V3 n1 = rotate(qinv(q1), nrm), n2 = rotate(qinv(q2), nrm);

    V3 r1l[4], r2l[4], r1w[4], r2w[4];
    float vn_bars[4];
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i >= npts) continue;
        V3 r1, r2;
        localContacts(pre_x1, pre_q1, pre_x2, pre_q2, ct, nrm, i, r1, r2);
        V3 r1p = rotate(pre_q1, r1), r2p = rotate(pre_q2, r2);
        V3 v_bar = relVel(pv1, pv2, po1, po2, r1p, r2p);
        vn_bars[i] = dot(nrm, v_bar);
        r1l[i] = r1; r2l[i] = r2;
        r1w[i] = rotate(q1, r1); r2w[i] = rotate(q2, r2);
    }
    for (int it = 0; it < 2; it++) {
#pragma unroll
        for (int i = 0; i < 4; i++) {
            if (i >= npts) continue;
            V3 vr = relVel(v1, v2, o1, o2, r1w[i], r2w[i]);
            float vn = dot(nrm, vr);
            float e = 0.3f;
            if (fabsf(vn_bars[i]) <= p.restitution_threshold) e = 0.f;
            float mag = fminRef(-e * vn_bars[i], 0.f) - vn;
            applyVelUpdate(v1, v2, o1, o2, q1, q2, cross(r1l[i], n1), cross(r2l[i], n2),
                           m1, m2, nrm, mag);
        }
    }
#pragma unroll
    for (int i = 0; i < 4; i++) {
        if (i >= npts) continue;
        V3 vr = relVel(v1, v2, o1, o2, r1w[i], r2w[i]);
        float fric = mu_d * fabsf(ct->lambda_n[i]) / h;
        float vn = dot(nrm, vr);
        V3 vt = vr - nrm * vn;
        float vt_len = len(vt);
        if (vt_len != 0 && fric != 0.f) {
            float corrected = -fminRef(fric, vt_len);
            V3 dw = vt / vt_len;
            V3 d1 = rotate(qinv(q1), dw), d2 = rotate(qinv(q2), dw);
            applyVelUpdate(v1, v2, o1, o2, q1, q2, cross(r1l[i], d1), cross(r2l[i], d2),
                           m1, m2, dw, corrected);
        }
    }
    if (v.dep[a]) stVel(v.vel, a, v1, o1);
    if (v.dep[b]) stVel(v.vel, b, v2, o2);
}

// One Gauss-Seidel sweep in the reference's contact order: level by level,
// the contacts of one level in parallel.
template <bool kVelocity>
MDR_HD void solveLevels(const BlockCtx &c, WorldView &v, const StepParams &p)
{
    const int np = v.misc[0], max_level = v.misc[1];
    for (int lv = 1; lv <= max_level; lv++) {
        for (int i = c.tid(); i < np; i += c.nthreads()) {
            if (v.level[i] != lv) continue;
            ContactDev *ct = contactSlot(v, i);
            if (ct->num_points <= 0) continue;
            if (kVelocity) solveContactVelocities(v, p, ct);
            else solveContactPositions(v, p, ct);
        }
        c.sync();
    }
}

}  // namespace mdr
