// TEST INFRASTRUCTURE ONLY. Drives the UNMODIFIED reference asset path
//   PhysicsLoader::loadHullFromDisk   src/physics/physics_assets.cpp:205-252
//   imp::ImportedObject::importObject src/common/importer.cpp:35-437 (OBJ reader + meshoptimizer vertex remap,
//                                     external/meshoptimizer/src/indexgenerator.cpp:234-287, vendored)
//   HalfEdgeMesh::construct           src/physics/geometry.cpp:52-194
// compiled where the sources lie (oracle/Makefile, target assets_ref) and hands the loaded hull back as flat
// tables. Used by tests/test_assets.py and tests/golden/make_asset_golden.py; never by the package.
#include <madrona/physics_assets.hpp>

#include <cstdint>
#include <cstring>

using namespace madrona;
using namespace madrona::phys;

extern "C" {

// counts = { half-edges, edges, polygons, vertices }; arrays must hold `cap` entries of each kind.
// half_edges: [4*n] next, twin, root vertex, polygon. Returns 0, or -1 when a table exceeds cap.
int32_t orc_load_hull_obj(const char *obj_path, int32_t cap, float *verts, uint32_t *half_edges, uint32_t *edges,
                          uint32_t *polys, float *planes, float *aabb6, int32_t *counts)
{
    PhysicsLoader loader(PhysicsLoader::StorageType::CPU, 1);
    PhysicsLoader::LoadedHull hull = loader.loadHullFromDisk(obj_path);
    const geometry::HalfEdgeMesh &m = hull.collisionMesh;
    counts[0] = (int32_t)m.mHalfEdgeCount; counts[1] = (int32_t)m.mEdgeCount;
    counts[2] = (int32_t)m.mPolygonCount; counts[3] = (int32_t)m.mVertexCount;
    for (int k = 0; k < 4; k++) if (counts[k] > cap) return -1;
    for (uint32_t i = 0; i < m.mHalfEdgeCount; i++) {
        half_edges[4*i] = m.mHalfEdges[i].next;
        half_edges[4*i+1] = m.mHalfEdges[i].twin;
        half_edges[4*i+2] = m.mHalfEdges[i].rootVertex;
        half_edges[4*i+3] = m.mHalfEdges[i].polygon;
    }
    memcpy(edges, m.mEdges, 4 * m.mEdgeCount);
    memcpy(polys, m.mPolygons, 4 * m.mPolygonCount);
    memcpy(planes, m.mFacePlanes, 16 * m.mPolygonCount);
    memcpy(verts, m.mVertices, 12 * m.mVertexCount);
    memcpy(aabb6, &hull.aabb, 24);
    return 0;
}

}
