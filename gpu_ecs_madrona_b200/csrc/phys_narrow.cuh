// Narrowphase: hull-hull and hull-plane SAT, clipping, manifold reduction
// (src/physics/narrowphase.cpp:139-235, 309-393, 418-576, 578-661, 678-1121,
// 1182-1364, 1515-1728), one thread per candidate pair, hull vertices and face
// planes pre-transformed once per body per substep into shared memory.
#pragma once
#include "phys_world.cuh"
#include "phys_solver.cuh"

namespace mdr {

#ifndef MDR_MAX_POLY
#define MDR_MAX_POLY 24     // vertices of a clipped contact polygon
#endif

// World-space hull of one body, in SMEM.
struct HullRef {
    const float *v;        // xyz packed
    const float *pl;       // nx ny nz d
    const ObjDev *o;
    const HEdgeDev *he;
    const uint32_t *edges;
    const uint32_t *polys;
    int nv, nf, ne;
    V3 center;
};

MDR_HD V3 hv(const HullRef &h, int i) { return mk3(h.v[3 * i], h.v[3 * i + 1], h.v[3 * i + 2]); }
MDR_HD Plane hp(const HullRef &h, int f)
{
    Plane p;
    p.n = mk3(h.pl[4 * f], h.pl[4 * f + 1], h.pl[4 * f + 2]);
    p.d = h.pl[4 * f + 3];
    return p;
}
MDR_HD HEdgeDev ldHE(const HullRef &h, uint32_t i)
{
    HEdgeDev e;
    e.next = MDR_LDG(&h.he[i].next); e.twin = MDR_LDG(&h.he[i].twin);
    e.root = MDR_LDG(&h.he[i].root); e.poly = MDR_LDG(&h.he[i].poly);
    return e;
}

MDR_HD HullRef hullRef(const WorldView &v, const StepParams &p, int row)
{
    HullRef h;
    const ObjDev *o = p.a.objs + v.obj[row];
    h.o = o;
    h.v = v.hverts + (size_t)row * v.hv_stride;
    h.pl = v.hplanes + (size_t)row * v.hp_stride;
    h.he = p.a.hedges + MDR_LDG(&o->hedge_off);
    h.edges = p.a.edges + MDR_LDG(&o->edge_off);
    h.polys = p.a.polys + MDR_LDG(&o->face_off);
    h.nv = MDR_LDG(&o->num_verts);
    h.nf = MDR_LDG(&o->num_faces);
    h.ne = MDR_LDG(&o->num_edges);
    h.center = ld3(v.pos, row);
    return h;
}

// Per body, per substep: world AABB of the object box (the narrowphase's
// cheap reject, narrowphase.cpp:1583-1596) and makeHullState (:139-212).
MDR_HD void transformBodies(const BlockCtx &c, WorldView &v, const StepParams &p)
{
    for (int i = c.tid(); i < v.n; i += c.nthreads()) {
        const ObjDev *o = p.a.objs + v.obj[i];
        const V3 t = ld3(v.pos, i);
        const Q4 r = ld4(v.rot, i);
        const V3 s = ld3(v.scale, i);
        stBox(v.box, i, applyTRS(objAABB(o), t, r, s));
        if (MDR_LDG(&o->type) != 2) continue;
        M3 rot = matFromQuat(r);
        M3 vtx = matMulDiag(rot, s);
        M3 nrm = matMulDiag(rot, mk3(1.f / s.x, 1.f / s.y, 1.f / s.z));
        const float *src_v = p.a.verts + 3 * MDR_LDG(&o->vert_off);
        const float *src_p = p.a.planes + 4 * MDR_LDG(&o->face_off);
        float *dv = v.hverts + (size_t)i * v.hv_stride;
        float *dp = v.hplanes + (size_t)i * v.hp_stride;
        const int nv = MDR_LDG(&o->num_verts), nf = MDR_LDG(&o->num_faces);
        for (int k = 0; k < nv; k++) {
            V3 lv = mk3(MDR_LDG(src_v + 3 * k), MDR_LDG(src_v + 3 * k + 1), MDR_LDG(src_v + 3 * k + 2));
            V3 wv = matMul(vtx, lv) + t;
            dv[3 * k] = wv.x; dv[3 * k + 1] = wv.y; dv[3 * k + 2] = wv.z;
        }
        for (int f = 0; f < nf; f++) {
            V3 ln = mk3(MDR_LDG(src_p + 4 * f), MDR_LDG(src_p + 4 * f + 1), MDR_LDG(src_p + 4 * f + 2));
            float ld = MDR_LDG(src_p + 4 * f + 3);
            V3 origin = matMul(vtx, ln * ld) + t;
            V3 wn = normalize(matMul(nrm, ln));
            dp[4 * f] = wn.x; dp[4 * f + 1] = wn.y; dp[4 * f + 2] = wn.z;
            dp[4 * f + 3] = dot(wn, origin);
        }
    }
}

// narrowphase.cpp:309-350
MDR_HD float hullDistanceFromPlane(const Plane &pl, const HullRef &h)
{
    float m = FLT_MAX;
    for (int i = 0; i < h.nv; i++) {
        float d = dot(pl.n, hv(h, i));
        if (d < m) m = d;
    }
    return m - pl.d;
}

struct FaceQuery { float sep; int face; Plane plane; };
// narrowphase.cpp:352-378: first strict maximum, early out once separated.
MDR_HD FaceQuery queryFaces(const HullRef &a, const HullRef &b)
{
    FaceQuery r;
    r.sep = -FLT_MAX; r.face = -1; r.plane.n = mk3(0.f, 0.f, 0.f); r.plane.d = 0.f;
    for (int f = 0; f < a.nf; f++) {
        Plane pl = hp(a, f);
        float d = hullDistanceFromPlane(pl, b);
        if (d > r.sep) {
            r.sep = d; r.face = f; r.plane = pl;
            if (r.sep > 0) break;
        }
    }
    return r;
}

// narrowphase.cpp:380-393
MDR_HD bool minkowskiFace(V3 a, V3 b, V3 cc, V3 d)
{
    V3 bxa = cross(b, a), dxc = cross(d, cc);
    float cba = dot(cc, bxa), dba = dot(d, bxa), adc = dot(a, dxc), bdc = dot(b, dxc);
    return cba * dba < 0.0f && adc * bdc < 0.0f && cba * bdc > 0.0f;
}

struct EdgeQuery { float sep; V3 normal; int ea, eb; };
// queryEdgeDirections (CPU branch), narrowphase.cpp:474-576 with
// buildsMinkowskiFace :418-426 and edgeDistance :433-472.
MDR_HD EdgeQuery queryEdges(const HullRef &a, const HullRef &b)
{
    EdgeQuery r;
    r.sep = -FLT_MAX; r.normal = mk3(0.f, 0.f, 0.f); r.ea = 0; r.eb = 0;
    for (int ia = 0; ia < a.ne; ia++) {
        const uint32_t hea = MDR_LDG(a.edges + ia);
        const HEdgeDev ea = ldHE(a, hea);
        const V3 an1 = hp(a, ea.poly).n, an2 = hp(a, MDR_LDG(&a.he[ea.twin].poly)).n;
        const V3 a1 = hv(a, ea.root), a2 = hv(a, MDR_LDG(&a.he[ea.next].root));
        for (int ib = 0; ib < b.ne; ib++) {
            const uint32_t heb = MDR_LDG(b.edges + ib);
            const HEdgeDev eb = ldHE(b, heb);
            V3 bn1 = hp(b, eb.poly).n, bn2 = hp(b, MDR_LDG(&b.he[eb.twin].poly)).n;
            float sep = -FLT_MAX;
            V3 normal = mk3(0.f, 0.f, 0.f);
            if (minkowskiFace(an1, an2, -bn1, -bn2)) {
                V3 b1 = hv(b, eb.root), b2 = hv(b, MDR_LDG(&b.he[eb.next].root));
                V3 da = a2 - a1, db = b2 - b1;
                V3 cr = cross(da, db);
                float l2 = len2(cr);
                if (l2 != 0) {
                    float inv = 1.f / sqrtf(l2);
                    normal = cr * inv;
                    if (dot(normal, a1 - a.center) < 0.0f) normal = -normal;
                    sep = dot(normal, b1 - a1);
                }
            }
            if (sep > r.sep) {
                r.sep = sep; r.normal = normal; r.ea = (int)hea; r.eb = (int)heb;
                if (r.sep > 0) return r;
            }
        }
    }
    return r;
}

// narrowphase.cpp:578-624: most anti-parallel face, first strict minimum.
MDR_HD int incidentFace(const HullRef &h, V3 ref_n)
{
    float m = FLT_MAX;
    int best = -1;
    for (int f = 0; f < h.nf; f++) {
        float d = dot(hp(h, f).n, ref_n);
        if (d < m) { m = d; best = f; }
    }
    return best;
}

MDR_HD float planeDist(const Plane &pl, V3 a) { return dot(a, pl.n) - pl.d; }

// clipPolygon, narrowphase.cpp:626-661; planeIntersection :246-250
MDR_HD int clipPolygon(V3 *dst, Plane cp, const V3 *in, int n_in)
{
    int n = 0;
    V3 v1 = in[n_in - 1];
    float d1 = planeDist(cp, v1);
    for (int i = 0; i < n_in; i++) {
        V3 v2 = in[i];
        float d2 = planeDist(cp, v2);
        if (d1 <= 0.0f && d2 <= 0.0f) {
            if (n < MDR_MAX_POLY) dst[n++] = v2;
        } else if (d1 <= 0.0f && d2 > 0.0f) {
            float dist = planeDist(cp, v1);
            if (n < MDR_MAX_POLY) dst[n++] = v1 + (v2 - v1) * (-dist / dot(cp.n, v2 - v1));
        } else if (d2 <= 0.0f && d1 > 0.0f) {
            float dist = planeDist(cp, v1);
            if (n < MDR_MAX_POLY) dst[n++] = v1 + (v2 - v1) * (-dist / dot(cp.n, v2 - v1));
            if (n < MDR_MAX_POLY) dst[n++] = v2;
        }
        v1 = v2;
        d1 = d2;
    }
    return n;
}

// Writes the manifold of a face contact into `ct` (buildFaceContactManifold,
// narrowphase.cpp:790-864). The >4-point reduction reproduces the reference as
// written: `largestArea` is never updated, so point 2 is the LAST candidate of
// positive signed area and point 3 the LAST of negative area; when
// none qualifies the reference reads an uninitialised slot -- here the slot
// repeats point 0, like oracle/port.cpp, and such worlds are excluded from
// bit-exact comparison by the tests.
MDR_HD void buildManifold(ContactDev *ct, V3 n, V3 *contacts, float *depths, int num)
{
    V3 pts[4];
    float dps[4];
    int cnt;
    if (num <= 4) {
        cnt = num;
        for (int i = 0; i < num; i++) { pts[i] = contacts[i]; dps[i] = depths[i]; }
    } else {
        cnt = 4;
        pts[0] = contacts[0];
        dps[0] = depths[0];
        V3 p0 = pts[0];
        float largest_d2 = 0.0f;
        int far_idx = 0;
        pts[1] = mk3(0.f, 0.f, 0.f); dps[1] = 0.f;
        for (int i = 1; i < num; i++) {
            float d2 = len2(p0 - contacts[i]);
            if (d2 > largest_d2) {
                largest_d2 = d2;
                pts[1] = contacts[i];
                dps[1] = depths[i];
                far_idx = i;
            }
        }
        contacts[far_idx] = pts[0];
        V3 diff0 = pts[1] - p0;
        int area_idx = 0;
        pts[2] = pts[3] = pts[0];
        dps[2] = dps[3] = dps[0];
        for (int i = 1; i < num; i++) {
            float area = dot(n, cross(diff0, contacts[i] - p0));
            if (area > 0.0f) { pts[2] = contacts[i]; dps[2] = depths[i]; area_idx = i; }
        }
        contacts[area_idx] = pts[0];
        for (int i = 1; i < num; i++) {
            float area = dot(n, cross(diff0, contacts[i] - p0));
            if (area < 0.0f) { pts[3] = contacts[i]; dps[3] = depths[i]; }
        }
    }
    // Manifold is expressed in world space through an identity transform
    // (narrowphase.cpp:855-863): kept, it turns -0 into +0.
    const Q4 ident = mk4(1.f, 0.f, 0.f, 0.f);
    const V3 zero = mk3(0.f, 0.f, 0.f);
    for (int i = 0; i < 4; i++) {
        V3 q = i < cnt ? rotate(ident, pts[i]) + zero : zero;
        ct->pts[i][0] = q.x; ct->pts[i][1] = q.y; ct->pts[i][2] = q.z;
        ct->pts[i][3] = i < cnt ? dps[i] : 0.f;
    }
    V3 wn = rotate(ident, n);
    ct->normal[0] = wn.x; ct->normal[1] = wn.y; ct->normal[2] = wn.z;
    ct->num_points = cnt;
}

// createFaceContact, narrowphase.cpp:866-972
MDR_HD void faceContact(ContactDev *ct, Plane ref_plane, int ref_face, int inc_face,
                        const HullRef &ref, const HullRef &other)
{
    V3 buf_a[MDR_MAX_POLY], buf_b[MDR_MAX_POLY];
    V3 *in = buf_a, *dst = buf_b;
    int n = 0;
    {
        uint32_t he = MDR_LDG(other.polys + inc_face), start = he;
        do {
            HEdgeDev e = ldHE(other, he);
            he = e.next;
            if (n < MDR_MAX_POLY) in[n++] = hv(other, e.root);
        } while (he != start);
    }
    uint32_t he = MDR_LDG(ref.polys + ref_face), start = he;
    V3 cur = hv(ref, MDR_LDG(&ref.he[he].root));
    do {
        he = MDR_LDG(&ref.he[he].next);
        V3 next = hv(ref, MDR_LDG(&ref.he[he].root));
        V3 edge = next - cur;
        Plane cp;
        cp.n = cross(edge, ref_plane.n);
        cp.d = dot(cp.n, cur);
        cur = next;
        n = n > 0 ? clipPolygon(dst, cp, in, n) : 0;
        V3 *tmp = in; in = dst; dst = tmp;
    } while (he != start);

    float depths[MDR_MAX_POLY];
    int below = 0;
    for (int i = 0; i < n; i++) {
        V3 vv = in[i];
        float d = planeDist(ref_plane, vv);
        if (d < 0.0f) {
            in[below] = vv - d * ref_plane.n;
            depths[below] = -d;
            below++;
        }
    }
    buildManifold(ct, ref_plane.n, in, depths, below);
}

// createFacePlaneContact, narrowphase.cpp:974-1017
MDR_HD void facePlaneContact(ContactDev *ct, Plane plane, int inc_face, const HullRef &h)
{
    V3 pts[MDR_MAX_POLY];
    float depths[MDR_MAX_POLY];
    uint32_t he = MDR_LDG(h.polys + inc_face), start = he;
    int n = 0;
    do {
        HEdgeDev e = ldHE(h, he);
        he = e.next;
        V3 vv = hv(h, e.root);
        float d = planeDist(plane, vv);
        if (d < 0.0f && n < MDR_MAX_POLY) {
            pts[n] = vv - d * plane.n;
            depths[n] = -d;
            n++;
        }
    } while (he != start);
    buildManifold(ct, plane.n, pts, depths, n);
}

// shortestSegmentBetween + createEdgeContact, narrowphase.cpp:1020-1121
MDR_HD void edgeContact(ContactDev *ct, V3 normal, float sep, int ea, int eb,
                        const HullRef &a, const HullRef &b)
{
    V3 a1 = hv(a, MDR_LDG(&a.he[ea].root)), a2 = hv(a, MDR_LDG(&a.he[MDR_LDG(&a.he[ea].next)].root));
    V3 b1 = hv(b, MDR_LDG(&b.he[eb].root)), b2 = hv(b, MDR_LDG(&b.he[MDR_LDG(&b.he[eb].next)].root));
    V3 v1 = a2 - a1, v2 = b2 - b1, v21 = b1 - a1;
    float d22 = dot(v2, v2), d11 = dot(v1, v1), d21 = dot(v2, v1);
    float d211 = dot(v21, v1), d212 = dot(v21, v2);
    float denom = d21 * d21 - d22 * d11;
    float s;
    if (fabsf(denom) < 0.00001f) s = 0.0f;
    else s = (d212 * d21 - d22 * d211) / denom;
    s = fmaxRef(fminRef(s, 1.0f), 0.0f);
    V3 contact = a1 + s * v1;
    const Q4 ident = mk4(1.f, 0.f, 0.f, 0.f);
    const V3 zero = mk3(0.f, 0.f, 0.f);
    V3 q = rotate(ident, contact) + zero;
    for (int i = 0; i < 4; i++) for (int k = 0; k < 4; k++) ct->pts[i][k] = 0.f;
    ct->pts[0][0] = q.x; ct->pts[0][1] = q.y; ct->pts[0][2] = q.z; ct->pts[0][3] = -sep;
    V3 wn = rotate(ident, normal);
    ct->normal[0] = wn.x; ct->normal[1] = wn.y; ct->normal[2] = wn.z;
    ct->num_points = 1;
}

// runNarrowphase (CPU branch) + narrowphaseDispatch + generateContacts
// (narrowphase.cpp:1515-1728,1182-1364,1366-1513), doSAT :678-758, doSATPlane
// :760-788, for pair slot `pi`. A slot without contact gets num_points = 0.
MDR_HD void narrowphasePair(WorldView &v, const StepParams &p, int pi)
{
    ContactDev *ct = contactSlot(v, pi);
    ct->num_points = 0;
    const uint32_t pr = v.pairs[pi];
    int a = pr & 0xffff, b = pr >> 16;
    int ta = MDR_LDG(&p.a.objs[v.obj[a]].type), tb = MDR_LDG(&p.a.objs[v.obj[b]].type);
    if ((uint32_t)ta > (uint32_t)tb) { int t = a; a = b; b = t; t = ta; ta = tb; tb = t; }
    if (!overlaps(ldBox(v.box, a), ldBox(v.box, b))) return;
    const int test = ta | tb;
    ct->ref_pad = 0; ct->alt_pad = 0;
    ct->lambda_n[0] = ct->lambda_n[1] = ct->lambda_n[2] = ct->lambda_n[3] = 0.f;
    if (test == 2) {
        HullRef ha = hullRef(v, p, a), hb = hullRef(v, p, b);
        FaceQuery fa = queryFaces(ha, hb);
        if (fa.sep > 0.0f) return;
        FaceQuery fb = queryFaces(hb, ha);
        if (fb.sep > 0.0f) return;
        EdgeQuery eq = queryEdges(ha, hb);
        if (eq.sep > 0.0f) return;
        bool face_a = fa.sep > eq.sep, face_b = fb.sep > eq.sep;
        if (face_a || face_b) {
            bool a_ref = fa.sep >= fb.sep;
            Plane rp = a_ref ? fa.plane : fb.plane;
            int rf = a_ref ? fa.face : fb.face;
            if (a_ref) {
                faceContact(ct, rp, rf, incidentFace(hb, rp.n), ha, hb);
                ct->ref = a; ct->alt = b;
            } else {
                faceContact(ct, rp, rf, incidentFace(ha, rp.n), hb, ha);
                ct->ref = b; ct->alt = a;
            }
        } else {
            edgeContact(ct, eq.normal, eq.sep, eq.ea, eq.eb, ha, hb);
            ct->ref = a; ct->alt = b;
        }
    } else if (test == 6) {
        HullRef ha = hullRef(v, p, a);
        Plane plane;
        plane.n = rotate(ld4(v.rot, b), mk3(0.f, 0.f, 1.f));
        plane.d = dot(plane.n, ld3(v.pos, b));
        float sep = hullDistanceFromPlane(plane, ha);
        if (sep > 0.0f) return;
        facePlaneContact(ct, plane, incidentFace(ha, plane.n), ha);
        ct->ref = b; ct->alt = a;
    }
    // Other primitive combinations assert in the reference (narrowphase.cpp:
    // 1197-1225,1268-1313): no contact.
}

}  // namespace mdr
