// Host-side construction of the shared collision-object tables: half-edge
// meshes from polygon lists (reference: HalfEdgeMesh::construct,
// src/physics/geometry.cpp:52-194) and the flat arrays the device reads
// (reference: PhysicsLoader / ObjectManager, src/physics/physics_assets.cpp:
// 163-396, include/madrona/physics.hpp:275-290). Plain C++, init time only.
#pragma once
#include <cmath>
#include <cstdint>
#include <cstring>
#include <map>
#include <string>
#include <utility>
#include <vector>

#include "phys_types.cuh"
#include "../../include/madrona_b200.h"

namespace mdr {

struct HostAssets {
    std::vector<ObjDev> objs;
    std::vector<float> verts;       // xyz
    std::vector<float> planes;      // nx ny nz d
    std::vector<HEdgeDev> hedges;
    std::vector<uint32_t> edges;
    std::vector<uint32_t> polys;
    int max_verts = 1, max_faces = 1;
    std::string error;
};

// Quirks of the reference construction that parity depends on: a face's entry
// half-edge is the LAST one created for it (so face vertices are walked
// starting from the polygon's last vertex), and an undirected edge is recorded,
// as the EARLIER twin's index, at the moment its second half-edge appears -- the
// edge list order is what breaks ties in the edge-edge SAT query.
inline bool buildHalfEdgeMesh(HostAssets &out, ObjDev &o, const float *verts, int num_verts,
                              const uint32_t *words, int num_words)
{
    o.vert_off = (int32_t)(out.verts.size() / 3);
    o.num_verts = num_verts;
    o.face_off = (int32_t)out.polys.size();
    o.hedge_off = (int32_t)out.hedges.size();
    o.edge_off = (int32_t)out.edges.size();
    out.verts.insert(out.verts.end(), verts, verts + 3 * num_verts);

    std::vector<HEdgeDev> hedges;
    std::map<std::pair<uint32_t, uint32_t>, uint32_t> by_ends;
    int k = 0;
    uint32_t face = 0;
    while (k < num_words) {
        const uint32_t n = words[k];
        if (n < 3 || k + 1 + (int)n > num_words) { out.error = "malformed polygon list"; return false; }
        const uint32_t *poly = words + k + 1;
        const uint32_t first = (uint32_t)hedges.size();
        for (uint32_t v = 0; v < n; v++) {
            const uint32_t a = poly[v], b = poly[(v + 1) % n];
            if (a >= (uint32_t)num_verts || by_ends.count({ a, b })) {
                out.error = "polygon list is not a closed 2-manifold";
                return false;
            }
            const uint32_t idx = (uint32_t)hedges.size();
            HEdgeDev he;
            he.next = v + 1 < n ? idx + 1 : first;
            he.twin = 0; he.root = a; he.poly = face;
            hedges.push_back(he);
            auto tw = by_ends.find({ b, a });
            if (tw != by_ends.end()) {
                hedges[idx].twin = tw->second;
                hedges[tw->second].twin = idx;
                out.edges.push_back(tw->second);
            }
            by_ends[{ a, b }] = idx;
        }
        const uint32_t entry = first + n - 1;
        out.polys.push_back(entry);

        // Face plane from the first three vertices met from the entry half-edge
        // (geometry.cpp:160-190), fp32, unfused.
        V3 pts[3];
        uint32_t he = entry;
        for (int i = 0; i < 3; i++) {
            const float *pv = verts + 3 * hedges[he].root;
            pts[i] = mk3(pv[0], pv[1], pv[2]);
            he = hedges[he].next;
        }
        V3 nrm = normalize(cross(pts[1] - pts[0], pts[2] - pts[0]));
        float d = dot(nrm, pts[0]);
        out.planes.push_back(nrm.x); out.planes.push_back(nrm.y);
        out.planes.push_back(nrm.z); out.planes.push_back(d);
        k += (int)n + 1;
        face++;
    }
    o.num_faces = (int32_t)face;
    o.face_pairs = (face % 2u) == 0u && face > 0u;
    for (uint32_t f = 0; f + 1 < face && o.face_pairs; f += 2) {
        const float *a = &out.planes[4 * ((size_t)o.face_off + f)], *b = a + 4;
        // numerically opposite (a zero may have either sign: it cannot change a comparison downstream)
        if (!(a[0] == -b[0] && a[1] == -b[1] && a[2] == -b[2])) o.face_pairs = 0;
    }
    o.num_hedges = (int32_t)hedges.size();
    o.num_edges = (int32_t)out.edges.size() - o.edge_off;
    if (2 * o.num_edges != o.num_hedges) { out.error = "open mesh: an edge has no twin"; return false; }
    if (num_verts > MDR_MAX_HULL_VERTS || (int)face > MDR_MAX_HULL_FACES) {
        out.error = "hull exceeds the compiled vertex / face capacity of the narrowphase";
        return false;
    }
    out.hedges.insert(out.hedges.end(), hedges.begin(), hedges.end());
    if (num_verts > out.max_verts) out.max_verts = num_verts;
    if ((int)face > out.max_faces) out.max_faces = (int)face;
    return true;
}

inline bool buildAssets(HostAssets &out, const mdr_object *objs, int num_objs,
                        const float *verts, const uint32_t *poly_words)
{
    out.objs.resize(num_objs);
    for (int i = 0; i < num_objs; i++) {
        const mdr_object &s = objs[i];
        ObjDev &o = out.objs[i];
        memset(&o, 0, sizeof(o));
        o.type = s.type;
        for (int k = 0; k < 3; k++) {
            o.inv_inertia[k] = s.inv_inertia[k];
            o.inertia[k] = s.inv_inertia[k] == 0 ? 0.0f : 1.0f / s.inv_inertia[k];
            o.aabb[k] = s.aabb_min[k];
            o.aabb[3 + k] = s.aabb_max[k];
        }
        o.inv_mass = s.inv_mass; o.mu_s = s.mu_s; o.mu_d = s.mu_d;
        if (s.type == MDR_PRIM_HULL) {
            if (!buildHalfEdgeMesh(out, o, verts + 3 * s.vert_offset, s.num_verts,
                                   poly_words + s.poly_offset, s.num_poly_words))
                return false;
        } else if (s.type != MDR_PRIM_PLANE) {
            out.error = "unsupported primitive type (the reference asserts on spheres)";
            return false;
        }
    }
    // Never hand the device a null table.
    if (out.verts.empty()) out.verts.assign(3, 0.f);
    if (out.planes.empty()) out.planes.assign(4, 0.f);
    if (out.hedges.empty()) out.hedges.resize(1);
    if (out.edges.empty()) out.edges.resize(1);
    if (out.polys.empty()) out.polys.resize(1);
    return true;
}

// Step constants (reference: SolverData ctor src/physics/physics.cpp:15-32 and
// the BVH expansion factors :1023-1027).
inline void fillStepConstants(StepParams &p, const mdr_phys_config &cfg)
{
    p.h = cfg.delta_t / (float)cfg.num_substeps;
    p.g[0] = cfg.gravity[0]; p.g[1] = cfg.gravity[1]; p.g[2] = cfg.gravity[2];
    p.restitution_threshold = 2.f * len(mk3(p.g[0], p.g[1], p.g[2])) * p.h;
    p.vel_expand = 2.f * cfg.delta_t;
    p.accel_expand = 100.f * cfg.delta_t * cfg.delta_t;
    p.substeps = cfg.num_substeps;
}

}  // namespace mdr
