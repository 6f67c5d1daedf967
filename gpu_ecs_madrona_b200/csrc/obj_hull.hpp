// Collision hull from a Wavefront OBJ: the asset half of SURVEY §8f rank 4. Host code, init time only.
//
// Stands in for PhysicsLoader::loadHullFromDisk (src/physics/physics_assets.cpp:205-252) and the importer under
// it (imp::ImportedObject::importObject / loadOBJ, src/common/importer.cpp:35-437): the result is the vertex
// array, the polygon list (FastPolygonList words, include/madrona/physics.hpp:12-35) and the object-space AABB
// that mdr_phys_create takes for a hull object; buildHalfEdgeMesh (phys_assets.hpp) makes the half-edge tables
// from them exactly as HalfEdgeMesh::construct does.
//
// What parity depends on (pinned against the compiled reference, tests/test_assets.py):
//  * corners are first un-indexed, then merged again by BYTE equality of (position [, normal] [, uv]) and
//    numbered in order of first appearance in the face list (the importer's meshoptimizer remap,
//    external/meshoptimizer/src/indexgenerator.cpp:234-287): vertex order is the order the faces mention them,
//    not the order of the `v` lines, unused `v` lines vanish, -0.0 and 0.0 stay apart;
//  * only ' ' separates numbers on `v` lines, ' ' and '\r' separate corners on `f` lines; a corner is
//    pos[/[uv][/normal]], 1-based, no negative (relative) indices;
//  * decimal -> float is correctly rounded (the reference: fast_float; here: std::from_chars), no leading '+';
//  * the AABB starts at vertex 0 and grows with `<` / `else if >` per axis (math.hpp:1022-1041).
// Where the reference aborts (FATAL, assert) or runs into undefined behaviour this loader returns an error
// string instead: malformed numbers, index 0 or out of range, corners with and without normals / uvs mixed,
// an `o` line between the `v` lines and the `f` lines of a mesh, not exactly one mesh in the file. One
// deliberate difference: a mesh whose faces are ALL triangles makes the reference read a null faceCounts array
// (importer.cpp:337-352 hands out nullptr, physics_assets.cpp:221-235 dereferences it); here every face of such
// a mesh has three corners, which is what the importer's own convention means.
#pragma once
#include <charconv>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <string_view>
#include <vector>

namespace mdr {

struct ObjHull {
    std::vector<float> verts;        // xyz, merged, in order of first mention
    std::vector<uint32_t> words;     // [n, i0 .. i(n-1)]*
    int num_faces = 0;
    float aabb[6] = { 0, 0, 0, 0, 0, 0 };
    std::string error;
};

namespace objdetail {

inline const char *skipBlanks(const char *p, const char *end)
{
    while (p < end && *p == ' ') p++;
    return p;
}

// One decimal number at p; nullptr when there is none.
inline const char *readFloat(const char *p, const char *end, float &v)
{
    auto r = std::from_chars(p, end, v, std::chars_format::general);
    if (r.ptr == p) return nullptr;
    if (r.ec == std::errc::result_out_of_range) {
        // from_chars leaves v untouched: overflow -> +-inf, underflow -> the rounded subnormal / zero
        std::string tmp(p, r.ptr);
        v = strtof(tmp.c_str(), nullptr);
    }
    return r.ptr;
}

inline bool readVec(std::string_view s, int n, float *out)
{
    const char *p = s.data(), *end = s.data() + s.size();
    for (int k = 0; k < n; k++) {
        p = skipBlanks(p, end);
        p = readFloat(p, end, out[k]);
        if (!p) return false;
    }
    return true;
}

inline const char *readU32(const char *p, const char *end, uint32_t &v)
{
    auto r = std::from_chars(p, end, v, 10);
    if (r.ptr == p || r.ec != std::errc()) return nullptr;
    return r.ptr;
}

struct Corner { uint32_t pos, normal, uv; };

// pos[/[uv][/normal]]
inline const char *readCorner(const char *p, const char *end, Corner &c)
{
    c = Corner { 0, 0, 0 };
    p = readU32(p, end, c.pos);
    if (!p) return nullptr;
    if (p == end || *p != '/') return p;
    p++;
    if (p < end && *p == '/') {
        c.uv = 0;
    } else {
        p = readU32(p, end, c.uv);
        if (!p) return nullptr;
    }
    if (p == end || *p != '/') return p;
    p++;
    return readU32(p, end, c.normal);
}

}  // namespace objdetail

inline bool loadObjHull(const char *text, size_t len, ObjHull &out)
{
    using namespace objdetail;
    std::vector<float> positions, normals, uvs;
    std::vector<Corner> corners;
    std::vector<uint32_t> face_counts;
    int meshes = 0;
    auto fail = [&](const char *why) { out.error = why; return false; };

    auto commit = [&]() -> bool {
        if (corners.empty()) {
            if (!positions.empty() || !normals.empty() || !uvs.empty())
                return fail("an `o` line (or the end of the file) follows vertex data that no face uses");
            return true;
        }
        if (meshes == 1) return fail("more than one mesh in the file (the reference asserts exactly one)");
        const size_t np = positions.size() / 3, nn = normals.size() / 3, nt = uvs.size() / 2;
        const bool with_n = corners[0].normal > 0, with_t = corners[0].uv > 0;
        // key of a corner: the bytes of what it refers to
        std::vector<float> keys;
        const int kw = 3 + (with_n ? 3 : 0) + (with_t ? 2 : 0);
        keys.reserve(corners.size() * kw);
        for (const Corner &c : corners) {
            if (c.pos == 0 || c.pos > np) return fail("face refers to a vertex that does not exist");
            if ((c.normal > 0) != with_n || (c.uv > 0) != with_t)
                return fail("corners with and without normals / texture coordinates are mixed");
            if (c.normal > nn || c.uv > nt) return fail("face refers to a normal / texture coordinate that does not exist");
            const float *p = &positions[3 * (c.pos - 1)];
            keys.insert(keys.end(), p, p + 3);
            if (with_n) { const float *q = &normals[3 * (c.normal - 1)]; keys.insert(keys.end(), q, q + 3); }
            if (with_t) { const float *q = &uvs[2 * (c.uv - 1)]; keys.insert(keys.end(), q, q + 2); }
        }
        // merge by byte equality, numbered by first appearance (open addressing over corner indices)
        const size_t n = corners.size();
        size_t buckets = 1;
        while (buckets < n + n / 4 + 1) buckets *= 2;
        std::vector<uint32_t> table(buckets, ~0u), remap(n);
        uint32_t next = 0;
        for (size_t i = 0; i < n; i++) {
            const unsigned char *kb = (const unsigned char *)&keys[i * kw];
            uint64_t h = 1469598103934665603ull;
            for (int b = 0; b < kw * 4; b++) h = (h ^ kb[b]) * 1099511628211ull;
            size_t slot = (size_t)h & (buckets - 1);
            while (true) {
                const uint32_t e = table[slot];
                if (e == ~0u) {
                    table[slot] = (uint32_t)i;
                    remap[i] = next++;
                    out.verts.insert(out.verts.end(), &keys[i * kw], &keys[i * kw] + 3);
                    break;
                }
                if (memcmp(&keys[(size_t)e * kw], kb, (size_t)kw * 4) == 0) { remap[i] = remap[e]; break; }
                slot = (slot + 1) & (buckets - 1);
            }
        }
        size_t k = 0;
        for (uint32_t fc : face_counts) {
            out.words.push_back(fc);
            for (uint32_t j = 0; j < fc; j++) out.words.push_back(remap[k++]);
        }
        out.num_faces = (int)face_counts.size();
        positions.clear(); normals.clear(); uvs.clear(); corners.clear(); face_counts.clear();
        meshes++;
        return true;
    };

    const char *p = text, *end = text + len;
    while (p < end) {
        const char *nl = (const char *)memchr(p, '\n', (size_t)(end - p));
        std::string_view line(p, (size_t)((nl ? nl : end) - p));
        p = nl ? nl + 1 : end;
        if (line.empty()) continue;
        const char c0 = line[0], c1 = line.size() > 1 ? line[1] : '\0';
        if (c0 == '#' || c0 == 's') continue;
        if (c0 == 'o') {
            if (!commit()) return false;
        } else if (c0 == 'v') {
            float v[3];
            if (c1 == ' ') {
                if (!readVec(line.substr(1), 3, v)) return fail("malformed `v` line");
                positions.insert(positions.end(), v, v + 3);
            } else if (c1 == 'n') {
                if (!readVec(line.substr(2), 3, v)) return fail("malformed `vn` line");
                normals.insert(normals.end(), v, v + 3);
            } else if (c1 == 't') {
                if (!readVec(line.substr(2), 2, v)) return fail("malformed `vt` line");
                uvs.insert(uvs.end(), v, v + 2);
            }
        } else if (c0 == 'f') {
            const char *q = line.data() + 1, *qe = line.data() + line.size();
            uint32_t count = 0;
            while (true) {
                while (q < qe && (*q == ' ' || *q == '\r')) q++;
                if (q == qe) break;
                Corner c;
                q = readCorner(q, qe, c);
                if (!q) return fail("malformed `f` line");
                corners.push_back(c);
                count++;
            }
            if (count == 0) return fail("`f` line without corners");
            face_counts.push_back(count);
        }
    }
    if (!commit()) return false;
    if (meshes != 1) return fail("no mesh in the file");

    const float *v = out.verts.data();
    for (int k = 0; k < 3; k++) out.aabb[k] = out.aabb[3 + k] = v[k];
    for (size_t i = 1; i < out.verts.size() / 3; i++)
        for (int k = 0; k < 3; k++) {
            const float x = v[3 * i + k];
            if (x < out.aabb[k]) out.aabb[k] = x;
            else if (x > out.aabb[3 + k]) out.aabb[3 + k] = x;
        }
    return true;
}

}  // namespace mdr
