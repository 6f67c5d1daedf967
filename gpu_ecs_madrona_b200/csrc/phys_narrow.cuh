// Narrowphase: hull-hull and hull-plane SAT, clipping, manifold reduction
// (src/physics/narrowphase.cpp:139-235, 309-393, 418-576, 578-661, 678-1121,
// 1182-1364, 1515-1728), one thread per candidate pair. Hull vertices and face
// planes are transformed to world space on the fly (makeHullState, :139-212):
// streamed for the hull-plane test, into per-thread arrays for hull-hull.
// Contacts leave in solver form (SolverContact).
#pragma once
#include "phys_world.cuh"

namespace mdr {

#ifndef MDR_MAX_POLY
#define MDR_MAX_POLY 24     // vertices of a clipped contact polygon
#endif

MDR_HD SolverContact *contactSlot(const WorldView &v, int p)
{
    return p < v.smem_contacts ? v.contacts + p : v.overflow + (p - v.smem_contacts);
}

// Object-to-world transform of one body (makeHullState's matrices).
struct BodyXf {
    V3 t;
    M3 vtx, nrm;
    const ObjDev *o;
};

MDR_HD BodyXf bodyXf(const WorldView &v, const StepParams &p, int row)
{
    BodyXf x;
    x.o = p.a.objs + (v.meta[row] & 0xffffu);
    x.t = ld3(v.pos, row);
    const V3 s = ld3(v.scale, row);
    M3 rot = matFromQuat(ld4(v.rot, row));
    if (s.x == 1.f && s.y == 1.f && s.z == 1.f) {
        // unit scale: every product below is x * 1.f (and 1.f / 1.f = 1.f), the same bits as x
        x.vtx = rot;
        x.nrm = rot;
        return x;
    }
    x.vtx = matMulDiag(rot, s);
    x.nrm = matMulDiag(rot, mk3(1.f / s.x, 1.f / s.y, 1.f / s.z));
    return x;
}

MDR_HD V3 worldVert(const BodyXf &x, const float *verts, int k)
{
    V3 lv = mk3(MDR_LDG(verts + 3 * k), MDR_LDG(verts + 3 * k + 1), MDR_LDG(verts + 3 * k + 2));
    return matMul(x.vtx, lv) + x.t;
}

MDR_HD V3 worldNormal(const BodyXf &x, const float *planes, int f)
{
    V3 ln = mk3(MDR_LDG(planes + 4 * f), MDR_LDG(planes + 4 * f + 1), MDR_LDG(planes + 4 * f + 2));
    return normalize(matMul(x.nrm, ln));
}

MDR_HD Plane worldPlane(const BodyXf &x, const float *planes, int f)
{
    V3 ln = mk3(MDR_LDG(planes + 4 * f), MDR_LDG(planes + 4 * f + 1), MDR_LDG(planes + 4 * f + 2));
    float ld = MDR_LDG(planes + 4 * f + 3);
    V3 origin = matMul(x.vtx, ln * ld) + x.t;
    Plane pl;
    pl.n = normalize(matMul(x.nrm, ln));
    pl.d = dot(pl.n, origin);
    return pl;
}

// World-space hull of one body in per-thread arrays (hull-hull only).
struct HullRef {
    const V3 *v;
    const Plane *pl;
    const HEdgeDev *he;
    const uint32_t *edges;
    const uint32_t *polys;
    int nv, nf, ne;
    V3 center;
};

MDR_HD V3 hv(const HullRef &h, int i) { return h.v[i]; }
MDR_HD Plane hp(const HullRef &h, int f) { return h.pl[f]; }
MDR_HD HEdgeDev ldHE(const HEdgeDev *he, uint32_t i)
{
    HEdgeDev e;
    e.next = MDR_LDG(&he[i].next); e.twin = MDR_LDG(&he[i].twin);
    e.root = MDR_LDG(&he[i].root); e.poly = MDR_LDG(&he[i].poly);
    return e;
}
MDR_HD HEdgeDev ldHE(const HullRef &h, uint32_t i) { return ldHE(h.he, i); }

MDR_HD HullRef makeHull(const BodyXf &x, const StepParams &p, V3 *dst_v, Plane *dst_p)
{
    HullRef h;
    const ObjDev *o = x.o;
    const float *src_v = p.a.verts + 3 * MDR_LDG(&o->vert_off);
    const float *src_p = p.a.planes + 4 * MDR_LDG(&o->face_off);
    h.nv = MDR_LDG(&o->num_verts);
    h.nf = MDR_LDG(&o->num_faces);
    h.ne = MDR_LDG(&o->num_edges);
    for (int k = 0; k < h.nv; k++) dst_v[k] = worldVert(x, src_v, k);
    for (int f = 0; f < h.nf; f++) dst_p[f] = worldPlane(x, src_p, f);
    h.v = dst_v;
    h.pl = dst_p;
    h.he = p.a.hedges + MDR_LDG(&o->hedge_off);
    h.edges = p.a.edges + MDR_LDG(&o->edge_off);
    h.polys = p.a.polys + MDR_LDG(&o->face_off);
    h.center = x.t;
    return h;
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


// A contact manifold before it is handed to the solver.
struct Manifold { V3 pts[4]; float depths[4]; int n; V3 normal; };

// The manifold goes to world space through an identity transform
// (narrowphase.cpp:855-863): point = rotate(identity, p) + (0,0,0), normal =
// rotate(identity, n). For finite p, rotate(identity, p) is p plus a signed
// zero, and adding the +0 translation leaves p when p != 0 and +0 when p is
// either zero: exactly `p + 0.f`. The normal has no translation, so the sign of
// a zero component depends on the zeros inside rotate(): it is evaluated as is.
MDR_HD void finishManifold(Manifold &m, V3 n)
{
#pragma unroll
    for (int i = 0; i < 4; i++) {       // unrolled: the manifold stays in registers
        if (i < m.n) m.pts[i] = mk3(m.pts[i].x + 0.f, m.pts[i].y + 0.f, m.pts[i].z + 0.f);
    }
    m.normal = rotate(mk4(1.f, 0.f, 0.f, 0.f), n);
}

// Register-resident insertion of the first four manifold points.
MDR_HD void setManifoldPoint(Manifold &m, int idx, V3 pt, float depth)
{
    if (idx == 0) { m.pts[0] = pt; m.depths[0] = depth; }
    else if (idx == 1) { m.pts[1] = pt; m.depths[1] = depth; }
    else if (idx == 2) { m.pts[2] = pt; m.depths[2] = depth; }
    else if (idx == 3) { m.pts[3] = pt; m.depths[3] = depth; }
}

// buildFaceContactManifold, narrowphase.cpp:790-864. The >4-point reduction
// reproduces the reference as written: `largestArea` is never updated, so point
// 2 is the LAST candidate of positive signed area and point 3 the LAST of
// negative area; when none qualifies the reference reads a slot
// of its local manifold that it never wrote. The slot is zero here: the value
// the unmodified reference computes once its locals are defined (compiled with
// -ftrivial-auto-var-init=zero -fno-elide-constructors), so
// every world compares bit for bit with that build.
MDR_HD void buildManifold(Manifold &m, V3 n, V3 *contacts, float *depths, int num)
{
    const V3 zero = mk3(0.f, 0.f, 0.f);
    for (int i = 0; i < 4; i++) { m.pts[i] = zero; m.depths[i] = 0.f; }
    if (num <= 4) {
        m.n = num;
        for (int i = 0; i < num; i++) { m.pts[i] = contacts[i]; m.depths[i] = depths[i]; }
    } else {
        m.n = 4;
        m.pts[0] = contacts[0];
        m.depths[0] = depths[0];
        V3 p0 = m.pts[0];
        float largest_d2 = 0.0f;
        int far_idx = 0;
        for (int i = 1; i < num; i++) {
            float d2 = len2(p0 - contacts[i]);
            if (d2 > largest_d2) {
                largest_d2 = d2;
                m.pts[1] = contacts[i];
                m.depths[1] = depths[i];
                far_idx = i;
            }
        }
        contacts[far_idx] = m.pts[0];
        V3 diff0 = m.pts[1] - p0;
        int area_idx = 0;
        for (int i = 1; i < num; i++) {
            float area = dot(n, cross(diff0, contacts[i] - p0));
            if (area > 0.0f) { m.pts[2] = contacts[i]; m.depths[2] = depths[i]; area_idx = i; }
        }
        contacts[area_idx] = m.pts[0];
        for (int i = 1; i < num; i++) {
            float area = dot(n, cross(diff0, contacts[i] - p0));
            if (area < 0.0f) { m.pts[3] = contacts[i]; m.depths[3] = depths[i]; }
        }
    }
    finishManifold(m, n);
}

// Where contact generation takes hull vertices from: the per-thread arrays of
// the serial path, or recomputed from the body transform (same bits) on the
// warp-cooperative path, which keeps no per-thread copy of the hull.
struct ArrayVerts {
    const V3 *v;
    MDR_HD V3 at(int k) const { return v[k]; }
};
struct XfVerts {
    BodyXf x;
    const float *src;
    MDR_HD V3 at(int k) const { return worldVert(x, src, k); }
};
struct HullTopo {
    const HEdgeDev *he;
    const uint32_t *polys;
};

// createFaceContact, narrowphase.cpp:866-972
template <typename VR, typename VO>
MDR_HD void faceContact(Manifold &m, Plane ref_plane, int ref_face, int inc_face,
                        const HullTopo &ref, const VR &ref_v, const HullTopo &other, const VO &other_v)
{
    V3 buf_a[MDR_MAX_POLY], buf_b[MDR_MAX_POLY];
    V3 *in = buf_a, *dst = buf_b;
    int n = 0;
    {
        uint32_t he = MDR_LDG(other.polys + inc_face), start = he;
        do {
            HEdgeDev e = ldHE(other.he, he);
            he = e.next;
            if (n < MDR_MAX_POLY) in[n++] = other_v.at((int)e.root);
        } while (he != start);
    }
    uint32_t he = MDR_LDG(ref.polys + ref_face), start = he;
    V3 cur = ref_v.at((int)MDR_LDG(&ref.he[he].root));
    do {
        he = MDR_LDG(&ref.he[he].next);
        V3 next = ref_v.at((int)MDR_LDG(&ref.he[he].root));
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
    buildManifold(m, ref_plane.n, in, depths, below);
}

// shortestSegmentBetween + createEdgeContact, narrowphase.cpp:1020-1121
template <typename VA, typename VB>
MDR_HD void edgeContact(Manifold &m, V3 normal, float sep, int ea, int eb,
                        const HullTopo &a, const VA &a_v, const HullTopo &b, const VB &b_v)
{
    V3 a1 = a_v.at((int)MDR_LDG(&a.he[ea].root)), a2 = a_v.at((int)MDR_LDG(&a.he[MDR_LDG(&a.he[ea].next)].root));
    V3 b1 = b_v.at((int)MDR_LDG(&b.he[eb].root)), b2 = b_v.at((int)MDR_LDG(&b.he[MDR_LDG(&b.he[eb].next)].root));
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
    for (int i = 0; i < 4; i++) { m.pts[i] = zero; m.depths[i] = 0.f; }
    m.pts[0] = mk3(contact.x + 0.f, contact.y + 0.f, contact.z + 0.f);   // see finishManifold
    m.depths[0] = -sep;
    m.n = 1;
    m.normal = rotate(ident, normal);
}

// doSATPlane + createFacePlaneContact (narrowphase.cpp:760-788,974-1017) with
// the hull streamed: incident face = most anti-parallel normal (first strict
// minimum), contact points = its vertices below the plane, projected onto it.
//
// The reference first takes the separation, min over ALL vertices of dot(n, v) - d, and stops when it is positive
// (doSATPlane :766-773). That scan decides nothing the incident face does not decide, so it is not run:
//  * a pair yields a contact iff the manifold has a point (:1510), i.e. iff a vertex of the incident face has
//    dot(n, v) - d < 0 -- the very subtraction createFacePlaneContact makes (:997), same operands;
//  * if such a vertex exists, the minimum over all vertices is at most its dot, so the separation is negative and
//    the reference does not stop: it builds exactly this manifold;
//  * if none exists the reference either stops (no contact) or builds an empty manifold (no contact).
// NaN poses agree too: comparisons with NaN are false on both sides (no minimum, no point below the plane), and a
// hull whose normals are all NaN has no incident face here (the reference has stopped at its vertex scan: every
// d < min is false, FLT_MAX - d is positive). 320 of the ~1200 instructions of a resting cube's pair.
MDR_HD bool hullPlaneContact(Manifold &m, const BodyXf &x, const StepParams &p, Plane plane)
{
    const ObjDev *o = x.o;
    const float *src_v = p.a.verts + 3 * MDR_LDG(&o->vert_off);
    const float *src_p = p.a.planes + 4 * MDR_LDG(&o->face_off);
    const int nf = MDR_LDG(&o->num_faces);
    float best = FLT_MAX;
    int inc_face = -1;
    if (MDR_LDG(&o->face_pairs)) {
        // Faces 2k, 2k + 1 with opposite local normals (phys_assets.hpp): every step of worldNormal and of the dot
        // product -- products, sums, the square root of a sum of squares, the division -- commutes with negation
        // under round-to-nearest, so the second face's value is the first one's negated, bit for bit (up to the sign
        // of a zero, which no comparison sees). Same scan order, half the transforms.
        for (int f = 0; f < nf; f += 2) {
            const float d = dot(worldNormal(x, src_p, f), plane.n);
            if (d < best) { best = d; inc_face = f; }
            const float d2 = -d;
            if (d2 < best) { best = d2; inc_face = f + 1; }
        }
    } else {
        for (int f = 0; f < nf; f++) {
            float d = dot(worldNormal(x, src_p, f), plane.n);
            if (d < best) { best = d; inc_face = f; }
        }
    }
    if (inc_face < 0) return false;
    // Up to four points stay in registers (every quad face); a fifth sends the
    // whole list through the generic reduction.
    const V3 zero = mk3(0.f, 0.f, 0.f);
    for (int i = 0; i < 4; i++) { m.pts[i] = zero; m.depths[i] = 0.f; }
    V3 more[MDR_MAX_POLY - 4];
    float more_d[MDR_MAX_POLY - 4];
    const HEdgeDev *hes = p.a.hedges + MDR_LDG(&o->hedge_off);
    uint32_t he = MDR_LDG(p.a.polys + MDR_LDG(&o->face_off) + inc_face), start = he;
    int n = 0;
    do {
        HEdgeDev e = ldHE(hes, he);
        he = e.next;
        V3 vv = worldVert(x, src_v, (int)e.root);
        float d = planeDist(plane, vv);
        if (d < 0.0f && n < MDR_MAX_POLY) {
            V3 pt = vv - d * plane.n;
            if (n < 4) setManifoldPoint(m, n, pt, -d);
            else { more[n - 4] = pt; more_d[n - 4] = -d; }
            n++;
        }
    } while (he != start);
    if (n <= 4) {
        m.n = n;
        finishManifold(m, plane.n);
    } else {
        V3 pts[MDR_MAX_POLY];
        float depths[MDR_MAX_POLY];
#pragma unroll
        for (int i = 0; i < 4; i++) { pts[i] = m.pts[i]; depths[i] = m.depths[i]; }
        for (int i = 4; i < n; i++) { pts[i] = more[i - 4]; depths[i] = more_d[i - 4]; }
        buildManifold(m, plane.n, pts, depths, n);
    }
    return true;
}

// doSAT + generateContacts for two hulls, narrowphase.cpp:678-758,1366-1513.
// Returns 0 (separated), 1 (contact, a is the reference body) or 2 (b is).
MDR_HD int hullHullContact(Manifold &m, const BodyXf &xa, const BodyXf &xb, const StepParams &p)
{
    V3 va[MDR_MAX_HULL_VERTS], vb[MDR_MAX_HULL_VERTS];
    Plane pa[MDR_MAX_HULL_FACES], pb[MDR_MAX_HULL_FACES];
    HullRef ha = makeHull(xa, p, va, pa), hb = makeHull(xb, p, vb, pb);
    FaceQuery fa = queryFaces(ha, hb);
    if (fa.sep > 0.0f) return 0;
    FaceQuery fb = queryFaces(hb, ha);
    if (fb.sep > 0.0f) return 0;
    EdgeQuery eq = queryEdges(ha, hb);
    if (eq.sep > 0.0f) return 0;
    bool face_a = fa.sep > eq.sep, face_b = fb.sep > eq.sep;
    if (face_a || face_b) {
        bool a_ref = fa.sep >= fb.sep;
        Plane rp = a_ref ? fa.plane : fb.plane;
        int rf = a_ref ? fa.face : fb.face;
        HullTopo ta, tb;
        ta.he = ha.he; ta.polys = ha.polys; tb.he = hb.he; tb.polys = hb.polys;
        ArrayVerts ava, avb;
        ava.v = va; avb.v = vb;
        if (a_ref) { faceContact(m, rp, rf, incidentFace(hb, rp.n), ta, ava, tb, avb); return 1; }
        faceContact(m, rp, rf, incidentFace(ha, rp.n), tb, avb, ta, ava);
        return 2;
    }
    {
        HullTopo ta, tb;
        ta.he = ha.he; ta.polys = ha.polys; tb.he = hb.he; tb.polys = hb.polys;
        ArrayVerts ava, avb;
        ava.v = va; avb.v = vb;
        edgeContact(m, eq.normal, eq.sep, eq.ea, eq.eb, ta, ava, tb, avb);
    }
    return 1;
}

// Hands a manifold to the solver: local arms (getLocalSpaceContacts,
// physics.cpp:365-382) and pre-solve normal velocities (:902-934) computed once.
// At this point of the substep pos/rot ARE the pre-solve poses (nothing has
// moved since substepRigidBodies).
MDR_HD void emitContact(WorldView &v, const StepParams &p, SolverContact *sc, const Manifold &m,
                        int ref, int alt)
{
    const bool s1 = !(v.meta[ref] & META_DEP), s2 = !(v.meta[alt] & META_DEP);
    sc->hdr = (uint32_t)ref | ((uint32_t)alt << 10) | ((uint32_t)m.n << 20) |
              ((uint32_t)s1 << 24) | ((uint32_t)s2 << 25);
    sc->normal[0] = m.normal.x; sc->normal[1] = m.normal.y; sc->normal[2] = m.normal.z;
    const V3 x1 = ld3(v.pos, ref), x2 = ld3(v.pos, alt);
    const Q4 q1 = ld4(v.rot, ref), q2 = ld4(v.rot, alt);
    const V3 pv1 = ldVelLin(v.prev_vel, ref), po1 = ldVelAng(v.prev_vel, ref);
    const V3 pv2 = ldVelLin(v.prev_vel, alt), po2 = ldVelAng(v.prev_vel, alt);
    const V3 zero = mk3(0.f, 0.f, 0.f);
    // A stable static whose rotation is the identity (the ground plane). All the solver keeps of such a side is
    // the world point o = rotate(q, rotate(inv q, u)) + x, u = c - x. rotate(identity, u) is u plus a signed zero
    // (see finishManifold): a non-zero component of u comes back bit for bit; a zero component comes back as a
    // zero of either sign, and adding x's component -- non-zero, or +0: a stable static has no -0 component
    // (staticStable) -- gives the same bits whichever sign it had. So o = u + x exactly, and the two rotations of
    // that side are not computed: 60 of the ~150 instructions per contact point.
    const bool id1 = s1 && q1.w == 1.f && q1.x == 0.f && q1.y == 0.f && q1.z == 0.f;
    const bool id2 = s2 && q2.w == 1.f && q2.x == 0.f && q2.y == 0.f && q2.z == 0.f;
    for (int i = 0; i < 4; i++) {
        sc->lambda[i] = 0.f;
        if (i >= m.n) continue;
        // selected, not indexed: an indexed read would put the whole manifold in local memory
        const V3 c1 = i == 0 ? m.pts[0] : i == 1 ? m.pts[1] : i == 2 ? m.pts[2] : m.pts[3];
        const float depth = i == 0 ? m.depths[0] : i == 1 ? m.depths[1] : i == 2 ? m.depths[2] : m.depths[3];
        V3 c2 = c1 - m.normal * depth;
        const V3 u1 = c1 - x1, u2 = c2 - x2;
        V3 r1 = zero, r2 = zero, r1w = u1, r2w = u2;      // for an identity static only r*w + x is used below
        if (!id1) { r1 = rotate(qinv(q1), u1); r1w = rotate(q1, r1); }
        if (!id2) { r2 = rotate(qinv(q2), u2); r2w = rotate(q2, r2); }
        // relative pre-solve velocity; a static body's pre-solve velocity is zero
        V3 va = s1 ? zero : pv1 + cross(po1, r1w);
        V3 vb = s2 ? zero : pv2 + cross(po2, r2w);
        sc->vn_bar[i] = dot(m.normal, va - vb);
        V3 o1 = s1 ? r1w + x1 : r1;
        V3 o2 = s2 ? r2w + x2 : r2;
        sc->r1[i][0] = o1.x; sc->r1[i][1] = o1.y; sc->r1[i][2] = o1.z;
        sc->r2[i][0] = o2.x; sc->r2[i][1] = o2.y; sc->r2[i][2] = o2.z;
    }
}

MDR_HD void probeContact(const StepParams &p, int sub, int pi, const Manifold *m, int ref, int alt)
{
    if (pi >= p.probe.max_contacts) return;
    ContactDev o;
    o.ref = ref; o.ref_pad = 0; o.alt = alt; o.alt_pad = 0;
#pragma unroll
    for (int i = 0; i < 4; i++) {
        const bool on = m && i < m->n;
        o.pts[i][0] = on ? m->pts[i].x : 0.f; o.pts[i][1] = on ? m->pts[i].y : 0.f;
        o.pts[i][2] = on ? m->pts[i].z : 0.f; o.pts[i][3] = on ? m->depths[i] : 0.f;
        o.lambda_n[i] = 0.f;
    }
    o.num_points = m ? m->n : 0;
    o.normal[0] = m ? m->normal.x : 0.f; o.normal[1] = m ? m->normal.y : 0.f; o.normal[2] = m ? m->normal.z : 0.f;
    p.probe.contacts[(size_t)sub * p.probe.max_contacts + pi] = o;
}

// A hull-hull pair that a warp found separated by a FACE axis is, in a resting or slowly moving scene, separated
// by the same face in the next substep and the next step. The pair slot remembers that face (WorldView::hint), and
// the pair's own thread re-evaluates just that one axis -- the same world plane, the same vertex scan, operand
// for operand what the full query computes for that face (queryFaceDirections, narrowphase.cpp:352-378;
// faceSeparationLane in phys_sat_warp.cuh) -- before asking for a warp. A positive value is the value the
// reference's scan would meet at that face (or it has stopped at an earlier face, positive as well): doSAT
// reports no contact either way (:678-758), so returning "no contact" here is exact, not an approximation.
// A stale hint (another pair now sits in the slot) is just an axis that may fail to separate: the warp runs.
constexpr uint32_t kHintFaceB = 64u;     // hint byte = 1 + face (< 64) + kHintFaceB * (a face of the pair's second hull)
MDR_HD bool satHintSeparates(const WorldView &v, const StepParams &p, int a, int b, uint32_t hint)
{
    const uint32_t code = hint - 1u;
    const int f = (int)(code & (kHintFaceB - 1u));
    const bool face_of_b = (code & kHintFaceB) != 0u;
    const BodyXf xf = bodyXf(v, p, face_of_b ? b : a), xo = bodyXf(v, p, face_of_b ? a : b);
    if (f >= MDR_LDG(&xf.o->num_faces)) return false;
    const Plane pl = worldPlane(xf, p.a.planes + 4 * MDR_LDG(&xf.o->face_off), f);
    const float *src_v = p.a.verts + 3 * MDR_LDG(&xo.o->vert_off);
    const int nv = MDR_LDG(&xo.o->num_verts);
    float m = FLT_MAX;
    for (int k = 0; k < nv; k++) {
        const float d = dot(pl.n, worldVert(xo, src_v, k));
        if (d < m) m = d;
    }
    return m - pl.d > 0.0f;
}

// runNarrowphase (CPU branch) + narrowphaseDispatch + generateContacts
// (narrowphase.cpp:1515-1728,1182-1364,1366-1513) for pair slot `pi`. A slot
// without contact gets a zero header.
MDR_HD void narrowphasePair(WorldView &v, const StepParams &p, int pi, bool probing, int sub)
{
    SolverContact *sc = contactSlot(v, pi);
    sc->hdr = 0;
    uint32_t pr = v.pairs[pi];
    if (pr & PAIR_SAT) { pr &= ~PAIR_SAT; v.pairs[pi] = pr; }     // last substep's mark
    int a = pr & 1023, b = (pr >> 10) & 1023;
    // order by primitive type bits: hull (2) before plane (4)
    if (!(v.meta[a] & META_HULL) && (v.meta[b] & META_HULL)) { int t = a; a = b; b = t; }
    const bool a_hull = (v.meta[a] & META_HULL) != 0, b_hull = (v.meta[b] & META_HULL) != 0;
    Manifold m;
    int ref = -1, alt = -1;
    if (overlaps(ldBox(v.box, a), ldBox(v.box, b))) {
        if (a_hull && b_hull) {
#if defined(__CUDA_ARCH__)
            // the face that separated this slot's hulls last time, if any, then a whole warp (phys_sat_warp.cuh)
            const uint32_t hint = v.hint[pi];
            if (hint && !probing && satHintSeparates(v, p, a, b, hint)) return;
            v.pairs[pi] = pr | PAIR_SAT;
            return;
#else
            int r = hullHullContact(m, bodyXf(v, p, a), bodyXf(v, p, b), p);
            if (r == 1) { ref = a; alt = b; }
            else if (r == 2) { ref = b; alt = a; }
#endif
        } else if (a_hull) {
            Plane plane;
            plane.n = rotate(ld4(v.rot, b), mk3(0.f, 0.f, 1.f));
            plane.d = dot(plane.n, ld3(v.pos, b));
            if (hullPlaneContact(m, bodyXf(v, p, a), p, plane)) { ref = b; alt = a; }
        }
        // Other primitive combinations assert in the reference
        // (narrowphase.cpp:1197-1225,1268-1313): no contact.
    }
    const bool hit = ref >= 0 && m.n > 0;
    if (hit) {
        emitContact(v, p, sc, m, ref, alt);
#if defined(__CUDA_ARCH__)
        if (v.level_any) {
            const int lv = pairLevel(pr);
            atomicOr(&v.level_any[sub & 1], 1u << (lv < 31 ? lv : 31));
        }
#endif
    }
    if (probing) probeContact(p, sub, pi, hit ? &m : nullptr, ref, alt);
}

// Per body, per substep: world AABB of the object box, the narrowphase's cheap
// reject (narrowphase.cpp:1583-1596).
MDR_HD void worldBoxes(const BlockCtx &c, WorldView &v, const StepParams &p, int sub)
{
    // same row -> thread mapping as substepBodies: no barrier between the two
    const int rows = sub > 0 ? v.misc[3] : v.n;
    for (int k = c.tid(); k < rows; k += c.nthreads()) {
        const int i = v.row_perm[k];
        const ObjDev *o = p.a.objs + (v.meta[i] & 0xffffu);
        stBox(v.box, i, applyTRS(objAABB(o), ld3(v.pos, i), ld4(v.rot, i), ld3(v.scale, i)));
    }
}

}  // namespace mdr
