// Observation feeder: what the reference's RenderingSystem nodes compute for a
// renderer (src/mw/render/interop.cpp:10-96), written to plain device buffers
// instead of a Vulkan TLAS / Metal instance buffer (graphics APIs: unavailable
// on this box). instanceTransformSetup -> one row-major 3x4 object-to-world
// matrix per body (Mat3x4::fromTRS, include/madrona/math.hpp:929-962, laid out
// as interop.cpp:27-41 does) + its object id; updateViewData -> PackedViewData
// {rotation, camera position, yScale} per view (interop.hpp:28-31,
// interop.cpp:69-79). The reference appends instances with an atomic counter
// (order across worlds arbitrary); here the order is world-major, row order.
#pragma once
#include "phys_world.cuh"

namespace mdr {

// out[12] = matrix[3][4] row-major, the translation in column 3
MDR_HD void instanceTransform(V3 t, Q4 r, V3 s, float *out)
{
    const float x2 = r.x * r.x, y2 = r.y * r.y, z2 = r.z * r.z;
    const float xz = r.x * r.z, xy = r.x * r.y, yz = r.y * r.z;
    const float wx = r.w * r.x, wy = r.w * r.y, wz = r.w * r.z;
    const V3 ds = mk3(2.f * s.x, 2.f * s.y, 2.f * s.z);
    const V3 c0 = mk3(s.x - ds.x * (y2 + z2), ds.x * (xy + wz), ds.x * (xz - wy));
    const V3 c1 = mk3(ds.y * (xy - wz), s.y - ds.y * (x2 + z2), ds.y * (yz + wx));
    const V3 c2 = mk3(ds.z * (xz + wy), ds.z * (yz - wx), s.z - ds.z * (x2 + y2));
    out[0] = c0.x; out[1] = c1.x; out[2] = c2.x; out[3] = t.x;
    out[4] = c0.y; out[5] = c1.y; out[6] = c2.y; out[7] = t.y;
    out[8] = c0.z; out[9] = c1.z; out[10] = c2.z; out[11] = t.z;
}

// out[8] = rotation w,x,y,z, camera position xyz, yScale
MDR_HD void packView(V3 pos, Q4 rot, V3 camera_offset, V3 world_offset, float y_scale, float *out)
{
    const V3 cam = pos + camera_offset + world_offset;
    out[0] = rot.w; out[1] = rot.x; out[2] = rot.y; out[3] = rot.z;
    out[4] = cam.x; out[5] = cam.y; out[6] = cam.z; out[7] = y_scale;
}

struct FeederJob {
    float world_offset[3];
    float *transforms;         // [W][stride][12]
    int32_t *object_ids;       // [W][stride]; -1 past the world's rows
    int32_t views_per_world;
    const int32_t *view_rows;  // [W][V] row the camera rides on
    const float *camera_offset;// [W][V][3]
    const float *y_scale;      // [W][V]
    float *views;              // [W][V][8]
};

MDR_HD void feedInstance(const PhysTables &t, const FeederJob &job, int w, int i)
{
    const size_t r = (size_t)w * t.stride + i;
    if (i < t.row_count[w]) {
        const V3 off = mk3(job.world_offset[0], job.world_offset[1], job.world_offset[2]);
        instanceTransform(ld3(t.pos, r) + off, ld4(t.rot, r), ld3(t.scale, r), job.transforms + 12 * r);
        job.object_ids[r] = t.obj[r];
    } else {
        for (int k = 0; k < 12; k++) job.transforms[12 * r + k] = 0.f;
        job.object_ids[r] = -1;
    }
}

MDR_HD void feedView(const PhysTables &t, const FeederJob &job, int w, int k)
{
    const size_t vi = (size_t)w * job.views_per_world + k;
    const size_t r = (size_t)w * t.stride + job.view_rows[vi];
    packView(ld3(t.pos, r), ld4(t.rot, r), ld3(job.camera_offset, vi),
             mk3(job.world_offset[0], job.world_offset[1], job.world_offset[2]), job.y_scale[vi], job.views + 8 * vi);
}

#if defined(__CUDACC__)
__global__ void physFeedInstancesKernel(PhysTables t, FeederJob job)
{
    const size_t total = (size_t)t.num_worlds * t.stride;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x)
        feedInstance(t, job, (int)(g / t.stride), (int)(g % t.stride));
}

__global__ void physFeedViewsKernel(PhysTables t, FeederJob job)
{
    const size_t total = (size_t)t.num_worlds * job.views_per_world;
    for (size_t g = (size_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (size_t)gridDim.x * blockDim.x)
        feedView(t, job, (int)(g / job.views_per_world), (int)(g % job.views_per_world));
}
#endif

}  // namespace mdr
