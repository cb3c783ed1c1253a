// scene.cuh — device data layout of the scene path (north_star stages n3-n7; the reference has no
// counterpart: its scene is the literals of RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:8,12-13).
// Shared by scene.cu (C-ABI glue), scene_build.cu (GPU LBVH build) and kernels_scene.cu (wavefront
// path tracer).  Not part of the public ABI.
#pragma once
#include <cstddef>
#include <cstdint>
#include <cuda_runtime.h>

#include "../../include/ptb200.h"
#include "row_shard.h"

namespace ptb {

// 64-byte two-child node (fetched as two 256-bit loads):
//   a = (c0.lo.x, c0.hi.x, c0.lo.y, c0.hi.y)
//   b = (c1.lo.x, c1.hi.x, c1.lo.y, c1.hi.y)
//   c = (c0.lo.z, c0.hi.z, c1.lo.z, c1.hi.z)
//   d = (ref0, ref1, 0, 0)
// A child reference r >= 0 is the index of another node; r < 0 is a leaf: with q = ~r,
// first = q & 0x07FFFFFF (position in the leaf-ordered primitive arrays) and count = (q >> 27) + 1.
struct __align__(16) BvhNode {
    float4 a, b, c;
    int4 d;
};
static_assert(sizeof(BvhNode) == 64, "node must be 64 bytes");
// kernels_scene.cu load_node() reads a node as two 32 B-aligned 256-bit loads (a|b, c|d): node arrays come from cudaMalloc
// (256 B aligned) and are indexed in whole nodes
static_assert(offsetof(BvhNode, c) == 32, "second half of a node must start at byte 32");

constexpr uint32_t LEAF_FIRST_MASK = 0x07FFFFFFu;
constexpr int LEAF_COUNT_SHIFT = 27;
constexpr uint32_t MAX_PRIMS_PER_CLASS = 1u << 27;
constexpr uint32_t MAX_LEAF = 8;
constexpr int EMPTY_REF = (int)0x80000000;   // "no child" (single-leaf hierarchies)

__host__ __device__ inline int make_leaf_ref(uint32_t first, uint32_t count)
{
    return ~(int)(first | ((count - 1u) << LEAF_COUNT_SHIFT));
}

struct DeviceBvh {
    BvhNode *nodes = nullptr;     // n_nodes entries; node 0 is the root when root_ref == 0
    uint32_t *order = nullptr;    // n_prims: class-local primitive index at each leaf-order position
    uint32_t n_nodes = 0, n_prims = 0, n_leaves = 0;
    int root_ref = 0;             // node 0, or a leaf reference when everything collapsed into one leaf
};

// Launch-uniform parameters of the scene kernels (kernel argument -> constant bank).
struct SceneK {
    uint32_t width, height;
    uint32_t row_offset, row_stride, row_block;
    float div_x, div_y;
    float cam_o[3], cam_u[3], cam_v[3], cam_w[3];
    float sky_h[3], sky_z[3];
    float t_min;
    uint32_t shade_mode, max_depth, seed;
    uint32_t n_spheres, n_tris, n_materials;
    // spheres in leaf order: centre+radius, and (original id, material)
    const float4 *sph;
    const uint2 *sph_meta;
    const BvhNode *sph_nodes;
    int sph_root;
    // triangles in leaf order, 3 float4 each: (v0 | original id), (e1 | material), (e2 | 0)
    const float4 *tri;
    const BvhNode *tri_nodes;
    int tri_root;
    const ptb_material *materials;
    uint32_t n_sph_nodes, n_tri_nodes;  // node counts (bounds-checked build)
    // next-event estimation towards a directional light (ptb_scene_set_sun); sun_on = 0: none
    uint32_t sun_on;
    float sun_dir[3], sun_rad[3];
};

// Wavefront path state, structure of arrays, ping-pong (in/out) per bounce:
//   s0 = (o.x, o.y, o.z, shard-local pixel index | sample of the wave << 26), s1 = (d.x, d.y, d.z, thr.r), s2 = (thr.g, thr.b)
struct PathQueue {
    float4 *s0 = nullptr;
    float4 *s1 = nullptr;
    float2 *s2 = nullptr;
};

struct SceneStatsDev {
    unsigned long long rays, nodes_visited, prims_tested, scatters, paths, max_stack;
};

struct SceneState {
    ptb_scene_desc desc{};          // host copy (pointers nulled)
    ptb_scene_info info{};
    DeviceBvh sph_bvh, tri_bvh;
    float4 *sph = nullptr;
    uint2 *sph_meta = nullptr;
    float4 *tri = nullptr;
    ptb_material *materials = nullptr;
    PathQueue q[2];
    uint32_t q_cap = 0;             // paths per queue = capacity of one wave
    float4 *slices = nullptr;       // batched waves: q_cap radiance slots, [sample of the wave][pixel of the shard]
    uint32_t wave_samples = 0;      // samples per wave: 0 = as many as fit the queues (ptb_scene_set_wave_samples)
    uint32_t *counts = nullptr;     // max_depth + 2 queue counters
    uint32_t counts_len = 0;
    float4 *frame_accum = nullptr;  // scratch accumulation for ptb_compute_frame
    SceneStatsDev *stats = nullptr;
    uint32_t *error_flag = nullptr; // traversal stack overflow etc.
    bool count_stats = false;
    uint32_t last_n0 = 0;           // camera rays of the last wave
    bool sun_on = false;            // ptb_scene_set_sun
    float sun_dir[3] = {0.f, 1.f, 0.f}, sun_rad[3] = {0.f, 0.f, 0.f};
};

// scene_build.cu: builds one hierarchy from per-primitive boxes (lo/hi as float4, w unused).
cudaError_t build_lbvh(const float4 *prim_lo, const float4 *prim_hi, uint32_t n, uint32_t max_leaf, float pad,
                       DeviceBvh *out, cudaStream_t s);
void free_bvh(DeviceBvh *b);

// kernels_scene.cu
cudaError_t launch_sphere_bounds(const float4 *sph, uint32_t n, float4 *lo, float4 *hi, cudaStream_t s);
cudaError_t launch_tri_bounds(const float *verts9, uint32_t n, float4 *lo, float4 *hi, cudaStream_t s);
cudaError_t launch_pack_spheres(const float4 *sph_in, const uint32_t *mat_in, const uint32_t *order, uint32_t n,
                                float4 *sph_out, uint2 *meta_out, cudaStream_t s);
cudaError_t launch_pack_tris(const float *verts9, const uint32_t *mat_in, const uint32_t *order, uint32_t n,
                             uint32_t id_base, float4 *tri_out, cudaStream_t s);
// one wave: samples [sample, sample + batch) of every pixel of the row shard, all bounces; adds into accum
// (batch > 1 needs st->slices to hold batch slices and batch * shard pixels <= st->q_cap)
cudaError_t launch_scene_wave(const SceneK &k, SceneState *st, float4 *accum, int32_t *hit_ids, uint32_t sample, uint32_t batch,
                              bool jitter, cudaStream_t s, uint64_t *launches);
cudaError_t launch_frame_finish(const float4 *accum, float4 *f32_out, uint32_t *rgba8_out, uint32_t n_pixels,
                                cudaStream_t s);
cudaError_t launch_trace_rays(const SceneK &k, SceneState *st, uint64_t n, const float *o, const float *d, float t_min,
                              float t_max, int mode, int32_t *out_prim, float *out_t, cudaStream_t s);

}  // namespace ptb
