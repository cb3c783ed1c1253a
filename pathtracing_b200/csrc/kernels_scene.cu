// kernels_scene.cu — sm_100a kernels of the scene path (north_star stages n3, n5, n6, n7).
// The reference's only tracing arithmetic is RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:6-39;
// its camera ray (:7-10), sphere quadratic (:16-27) and normal / miss colours (:28-36) are
// generalised here, everything else (triangles, hierarchy traversal, materials, bounces, queues)
// has no reference counterpart.  The arithmetic spec is frozen in oracle/pt_scene_oracle.h; this
// file restates it independently.
//
// THIS TRANSLATION UNIT IS COMPILED WITH -fmad=false (csrc/Makefile): every `*`, `+`, `-`, `/`,
// sqrtf below is one IEEE binary32 round-to-nearest operation, never contracted, which is what
// makes the results bit-identical to the CPU oracle.  FMAs appear only where written as fmaf().
//
// Wavefront organisation: one wave = `batch` samples of every pixel of this context's row shard
// (scene.cu packs as many as fit the path queues), camera rays generated in 8x4 pixel tiles.
// Per bounce ONE fused kernel (closest hit over both hierarchies + scatter), run by a persistent
// grid whose warps pull 32-path chunks from an atomic cursor.  Paths that continue are repacked
// into the next queue with warp ballot + prefix popcount + one atomicAdd per warp, so every
// bounce starts from a dense array (structure of arrays, 40 B per path); from a certain depth on
// the few paths left finish inside one launch (tail mode).  With batch = 1 a pixel owns at most
// one live path per wave and terminated paths add their radiance to the float4 accumulation
// buffer with a plain read-modify-write; with batch > 1 every path stores its radiance into its
// own slot of slices[sample of the wave][pixel of the shard] and reduce_slices_kernel adds the
// slices in sample order — the same additions in the same order, so the results do not depend
// on the wave size.
#include <cfloat>
#include <cstdlib>

#include "ptb_math.cuh"
#include "scene.cuh"

namespace ptb {
namespace {

constexpr int STACK_N = 64;
#ifndef PTB_BOUNCE_MIN_BLOCKS
// resident 256-thread blocks per SM the bounce kernel is compiled for (register cap).  Measured with the lean traversal
// stack (profiles/r01_notes.md): 3 / 4 / 5 / 6 blocks -> c4 1,064 / 1,051 / 1,109 / 1,083, c3 3,781 / 3,732 / 3,739 / 3,671
// Msamples/s; 5 = 48 registers with ~100 B of spills, 40 warps per SM to hide the L2 latency of the node fetches.
#define PTB_BOUNCE_MIN_BLOCKS 5
#endif
constexpr unsigned FULL = 0xFFFFFFFFu;

// compute-sanitizer is closed on the GPU pool this is developed on (profiles/r02_sanitizer.txt), so the indices the scene
// kernels compute themselves are checked by a build of their own: -DPTB_BOUNDS_CHECK (experiments/build_variant.sh) turns
// every PTB_CHK into a test that raises a bit of the scene's error word, which the next ptb_wait / read reports.  The
// product build compiles them away.
#ifdef PTB_BOUNDS_CHECK
#define PTB_CHK(flagp, cond, bit)                                                                                      \
    do {                                                                                                               \
        if (!(cond)) atomicOr((flagp), (bit));                                                                         \
    } while (0)
#else
#define PTB_CHK(flagp, cond, bit)                                                                                      \
    do {                                                                                                               \
    } while (0)
#endif
constexpr uint32_t CHK_NODE = 2u, CHK_PRIM = 4u, CHK_QUEUE_OUT = 8u, CHK_SLOT = 16u, CHK_LEAFQ = 32u, CHK_PIXEL = 64u, CHK_IN = 128u;

struct Counters {
    uint32_t nodes = 0, prims = 0, max_stack = 0;
};

struct RayT {
    float ox, oy, oz, dx, dy, dz;
    float ix, iy, iz;      // ~1/d      (culling only: never reaches a result)
    float px, py, pz;      // -o * (1/d) (culling only; FMA form of the slab test)
    float a, a2, a4;       // d.d, 2*(d.d), 4*(d.d)  (Test.hlsl:17, hoisted per ray)
};

struct Best {
    float t;
    float t_cull;  // t * (1 + 2^-19): what node culling compares against
    int id;        // unified original primitive id (spheres 0..ns-1, triangles ns..)
    int pos;       // leaf-order position inside its class
    int cls;       // 0 sphere, 1 triangle
};

__device__ __forceinline__ float dot3f(float ax, float ay, float az, float bx, float by, float bz)
{
    float s = ax * bx;
    s = s + ay * by;
    s = s + az * bz;
    return s;
}

__device__ __forceinline__ float rcp_approx(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}

// One 64 B node in two 256-bit loads (sm_100 LDG.256) instead of four 128-bit ones.  The lanes of a warp read different
// nodes, and the L1 data pipe spends one wavefront per lane and load instruction whatever its width: with 7-14 lanes active
// the walk was as close to the limit of that pipe (76 % of peak on the 1 M-triangle scene) as to its issue slots.
__device__ __forceinline__ void load_node(const BvhNode *__restrict__ n, float4 &a, float4 &b, float4 &c, int2 &d)
{
#ifndef PTB_NODE_LDG128
    float d2, d3;
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w) : "l"(n));
    asm("ld.global.nc.v8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
        : "=f"(c.x), "=f"(c.y), "=f"(c.z), "=f"(c.w), "=r"(d.x), "=r"(d.y), "=f"(d2), "=f"(d3) : "l"(&n->c));
#else
    a = __ldg(&n->a); b = __ldg(&n->b); c = __ldg(&n->c);
    d = __ldg(reinterpret_cast<const int2 *>(&n->d));
#endif
}

__device__ __forceinline__ void make_ray(RayT &r, float ox, float oy, float oz, float dx, float dy, float dz)
{
    r.ox = ox; r.oy = oy; r.oz = oz;
    r.dx = dx; r.dy = dy; r.dz = dz;
    // 1 ulp; absorbed by the culling slack.  A direction component of exactly 0 (or a denormal, flushed) gives +-inf, and an
    // infinite reciprocal poisons the FMA slab form: plane * inf + (-o * inf) is inf - inf = NaN on one side of the origin
    // and -inf on the other, so a box the ray lies INSIDE came out as the empty interval and whole subtrees were culled
    // (found by the 2-GPU scene test: 1 path in 3*10^5 takes a scattered direction with d.x == 0).  Clamped to +-2^96 the
    // products stay finite and the slab of a parallel ray is [-huge, +huge] inside the box, same-signed outside.
    constexpr float RCP_MAX = 7.9228163e28f;
    r.ix = fminf(fmaxf(rcp_approx(dx), -RCP_MAX), RCP_MAX);
    r.iy = fminf(fmaxf(rcp_approx(dy), -RCP_MAX), RCP_MAX);
    r.iz = fminf(fmaxf(rcp_approx(dz), -RCP_MAX), RCP_MAX);
    r.px = -(ox * r.ix); r.py = -(oy * r.iy); r.pz = -(oz * r.iz);
    r.a = dot3f(dx, dy, dz, dx, dy, dz);
    r.a2 = 2.0f * r.a;
    r.a4 = 4.0f * r.a;
}

// Test.hlsl:16-27 generalised; far root when the near one is outside (t_min, t_max).
__device__ __forceinline__ bool hit_sphere(const float4 cr, const RayT &r, float t_min, float t_max, float &t_out)
{
    const float ocx = r.ox - cr.x, ocy = r.oy - cr.y, ocz = r.oz - cr.z;      // :16
    const float b = 2.0f * dot3f(ocx, ocy, ocz, r.dx, r.dy, r.dz);           // :18
    const float cc = dot3f(ocx, ocy, ocz, ocx, ocy, ocz) - cr.w * cr.w;      // :19
    const float disc = b * b - r.a4 * cc;                                    // :21
    if (!(disc > 0.0f)) return false;                                        // :24
    const float s = sqrtf(disc);
    float t = ((-b) - s) / r.a2;                                             // :27
    if (!(t > t_min && t < t_max)) {
        t = ((-b) + s) / r.a2;
        if (!(t > t_min && t < t_max)) return false;
    }
    t_out = t;
    return true;
}

// Moeller-Trumbore, two-sided.  v0, e1 = v1-v0, e2 = v2-v0 (single exact-rounded subtractions,
// done once at upload).
__device__ __forceinline__ bool hit_triangle(const float4 v0, const float4 e1, const float4 e2, const RayT &r, float t_min,
                                             float t_max, float &t_out)
{
    const float px = r.dy * e2.z - r.dz * e2.y;
    const float py = r.dz * e2.x - r.dx * e2.z;
    const float pz = r.dx * e2.y - r.dy * e2.x;
    const float det = dot3f(e1.x, e1.y, e1.z, px, py, pz);
    if (!(det != 0.0f)) return false;
    const float inv = 1.0f / det;
    const float tx = r.ox - v0.x, ty = r.oy - v0.y, tz = r.oz - v0.z;
    const float u = dot3f(tx, ty, tz, px, py, pz) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return false;
    const float qx = ty * e1.z - tz * e1.y;
    const float qy = tz * e1.x - tx * e1.z;
    const float qz = tx * e1.y - ty * e1.x;
    const float w = dot3f(r.dx, r.dy, r.dz, qx, qy, qz) * inv;
    if (!(w >= 0.0f && u + w <= 1.0f)) return false;
    const float t = dot3f(e2.x, e2.y, e2.z, qx, qy, qz) * inv;
    if (!(t > t_min && t < t_max)) return false;
    t_out = t;
    return true;
}

// Conservative slab test of one child box against [t_min, t_far].  FMA_SLAB: t = fma(plane, 1/d, -o/d)
// (3 FFMA2 per box; its cancellation error is at most half an ulp of the ray origin in space, covered
// by the 2^-18-of-scene-extent padding the boxes get at build time, which includes the camera);
// otherwise t = (plane - o) * (1/d) for rays whose origin may lie far outside the scene
// (ptb_trace_rays).  The interval is widened by 2^-19 of |t| (approximate reciprocal + 3 roundings),
// so a box is never rejected when a primitive inside it would be accepted.  fminf/fmaxf drop NaNs.
template <bool FMA_SLAB>
__device__ __forceinline__ bool slab(float lox, float hix, float loy, float hiy, float loz, float hiz, const RayT &r,
                                     float t_min, float t_far, float &t_near)
{
    float ax, bx, ay, by, az, bz;
    if (FMA_SLAB) {
        // both planes of an axis in one FFMA2 (two IEEE-rounded lanes per issue slot; the node's 16 B loads deliver the pair
        // in adjacent registers and the ray's constants are broadcast operands): 6 instructions instead of 12, same values
        const float2 tx = __ffma2_rn(make_float2(lox, hix), make_float2(r.ix, r.ix), make_float2(r.px, r.px));
        const float2 ty = __ffma2_rn(make_float2(loy, hiy), make_float2(r.iy, r.iy), make_float2(r.py, r.py));
        const float2 tz = __ffma2_rn(make_float2(loz, hiz), make_float2(r.iz, r.iz), make_float2(r.pz, r.pz));
        ax = tx.x; bx = tx.y; ay = ty.x; by = ty.y; az = tz.x; bz = tz.y;
    } else {
        ax = (lox - r.ox) * r.ix; bx = (hix - r.ox) * r.ix;
        ay = (loy - r.oy) * r.iy; by = (hiy - r.oy) * r.iy;
        az = (loz - r.oz) * r.iz; bz = (hiz - r.oz) * r.iz;
    }
    t_near = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_min));
    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_far));
    return t_near <= fmaf(fabsf(tf), 1.9073486e-6f, tf);     // widened away from zero whatever the sign of tf
}

template <int CLS, bool ANY>
__device__ __forceinline__ bool test_leaf(const SceneK &k, int leaf, const RayT &r, float t_min, float t_max, Best &best,
                                          uint32_t &prims, uint32_t *error_flag)
{
    (void)error_flag;
    const uint32_t q = (uint32_t)~leaf;
    const uint32_t first = q & LEAF_FIRST_MASK, count = (q >> LEAF_COUNT_SHIFT) + 1u;
    prims += count;
    PTB_CHK(error_flag, first + count <= (CLS == 0 ? k.n_spheres : k.n_tris), CHK_PRIM);
    for (uint32_t i = first; i < first + count; ++i) {
        float t;
        int id;
        if (CLS == 0) {
            const float4 cr = __ldg(&k.sph[i]);
            if (!hit_sphere(cr, r, t_min, t_max, t)) continue;
            if (ANY) return true;
            if (!(t <= best.t)) continue;
            id = (int)__ldg(&k.sph_meta[i]).x;
        } else {
            const float4 v0 = __ldg(&k.tri[3 * (size_t)i]), e1 = __ldg(&k.tri[3 * (size_t)i + 1]),
                         e2 = __ldg(&k.tri[3 * (size_t)i + 2]);
            if (!hit_triangle(v0, e1, e2, r, t_min, t_max, t)) continue;
            if (ANY) return true;
            if (!(t <= best.t)) continue;
            id = __float_as_int(v0.w);
        }
        if (t < best.t || id < best.id || best.id < 0) {     // equal t: the smaller id wins (order independent)
            best.t = t;
            best.t_cull = fmaf(fabsf(t), 1.9073486e-6f, t);
            best.id = id;
            best.pos = (int)i;
            best.cls = CLS;
        }
    }
    return false;
}

// Stack-based traversal of one hierarchy, "while-while": the inner loop descends through nodes (test both child boxes,
// go to the nearer, push the farther) until the lane holds a leaf or has finished, then the leaf's primitives are tested.
// Lanes walk independently and the CALLER reconverges the warp (__syncwarp) before it shades: without that, independent
// thread scheduling let every lane run the shading that follows on its own, 2.6-3.0 active threads per instruction
// (profiles/r01_notes.md).  Two things an earlier version had were measured out again once waves carried millions of
// paths (profiles/r01_notes.md): lockstep rounds (the warp reconverging before every round of primitive tests; c3 3,516 vs
// 3,734 Msamples/s without) and entry distances stored next to the stacked references so that a pop could skip entries
// the closest hit had ruled out since (2-4 % fewer node visits, but twice the stack's L1 footprint — 2 x 128 B lines per
// warp and stack level; c4 929 -> 1,050 without).  A stale entry is culled one level later by the box tests of its children.
// Lanes without a ray pass active = false.  CLS 0 = spheres, 1 = triangles.  ANY: stop at the first hit.
#ifdef PTB_CHUNK_SMEM_STACK
__device__ __noinline__ int *short_stack_base()
{
    __shared__ int s[PTB_CHUNK_SMEM_STACK * 256];
    return s;
}
#endif

template <int CLS, bool ANY, bool STATS, bool FMA_SLAB, bool BRANCH_FREE = false>
__device__ __forceinline__ bool traverse(const SceneK &k, const RayT &r, float t_min, float t_max, bool active, Best &best,
                                         Counters &cn, uint32_t *error_flag)
{
    const BvhNode *__restrict__ nodes = CLS == 0 ? k.sph_nodes : k.tri_nodes;
    const int root = CLS == 0 ? k.sph_root : k.tri_root;
    if (root == EMPTY_REF) return false;               // launch-uniform
    int cur = active ? root : EMPTY_REF;               // node (>= 0), leaf (< 0) or EMPTY_REF = finished
    int stack[STACK_N];
    uint32_t prims = 0;
    bool found = false;
    if (BRANCH_FREE) {
        // Branch-free node step, for walks whose lanes disagree about the outcome of most steps (scattered rays in a
        // hierarchy of overlapping boxes: the sphere class behind the camera rays).  stack[0] is a sentinel — popping it
        // ends the walk, so a pop needs no emptiness test —, a push is one predicated store, a pop one predicated load,
        // and the three outcomes of a step (both children hit / one / none) are selects instead of three branches that a
        // divergent warp runs one after the other: 18 instructions behind the box decision for everybody, against 4 / 8 / 9
        // per outcome PRESENT in the warp.  Measured (profiles/r02_notes.md section 12): c3 +11 % with every launch on the
        // chunk schedule, but the coherent camera rays and the well-separated boxes of the terrain mostly see one outcome
        // per step and lose (c4 -17 %), hence the template switch.  The stack position is clamped to the last slot
        // (memory-safe); a walk that reached it is reported behind the loop, one level early.
        stack[0] = EMPTY_REF;
        int sp = 1, max_sp = 1;                        // sp: next free slot
        for (;;) {
            while (cur >= 0) {
                PTB_CHK(error_flag, (uint32_t)cur < (CLS == 0 ? k.n_sph_nodes : k.n_tri_nodes), CHK_NODE);
                float4 na, nb, nc;
                int2 nd;
                load_node(&nodes[cur], na, nb, nc, nd);
                if (STATS) ++cn.nodes;
                const float t_far = ANY ? t_max : best.t_cull;
                float tn0, tn1;
                const bool h0 = slab<FMA_SLAB>(na.x, na.y, na.z, na.w, nc.x, nc.y, r, t_min, t_far, tn0);
                const bool h1 = slab<FMA_SLAB>(nb.x, nb.y, nb.z, nb.w, nc.z, nc.w, r, t_min, t_far, tn1) && nd.y != EMPTY_REF;
                const bool both = h0 && h1, none = !(h0 || h1);
                const bool second_first = both ? tn1 < tn0 : !h0;      // descend into child 1 (both hit: the nearer one)
                int nxt = second_first ? nd.y : nd.x;
                const int other = second_first ? nd.x : nd.y;
                if (both) stack[sp] = other;
                if (both) ++sp;
                sp = min(sp, STACK_N - 1);
                max_sp = max(max_sp, sp);
                if (none) --sp;
                if (none) nxt = stack[sp];
                cur = nxt;
            }
            if (cur == EMPTY_REF) break;
            if (test_leaf<CLS, ANY>(k, cur, r, t_min, t_max, best, prims, error_flag)) {
                found = true;
                break;
            }
            --sp;
            cur = stack[sp];
        }
        if (max_sp >= STACK_N - 1) atomicOr(error_flag, 1u);
        if (STATS) {
            cn.max_stack = max(cn.max_stack, (uint32_t)(max_sp - 1));
            cn.prims += prims;
        }
        return ANY ? found : best.id >= 0;
    }
    int sp = 0;

    // A/B build (-DPTB_CHUNK_SMEM_STACK=N, experiments/build_variant.sh): the first N stack levels in shared memory, one
    // 4-byte column per thread ([level][thread]: conflict-free whatever the lanes' depths), deeper levels in local memory
    // as before.  Measured in round 2 (profiles/r02_notes.md section 5): no gain, so the product build keeps the local stack.
#ifdef PTB_CHUNK_SMEM_STACK
    int *const s_short = short_stack_base();            // one array for both hierarchies (they are walked one after the other)
#define PTB_STACK_AT(i) (*((i) < PTB_CHUNK_SMEM_STACK ? &s_short[(i) * 256 + threadIdx.x] : &stack[(i)]))
#else
#define PTB_STACK_AT(i) stack[(i)]
#endif
#define PTB_POP() (cur = sp > 0 ? PTB_STACK_AT(sp - 1) : EMPTY_REF, sp = sp > 0 ? sp - 1 : 0)

    for (;;) {
        while (cur >= 0) {
            PTB_CHK(error_flag, (uint32_t)cur < (CLS == 0 ? k.n_sph_nodes : k.n_tri_nodes), CHK_NODE);
            float4 na, nb, nc;
            int2 nd;
            load_node(&nodes[cur], na, nb, nc, nd);
            if (STATS) ++cn.nodes;
            const float t_far = ANY ? t_max : best.t_cull;
            float tn0, tn1;
            const bool h0 = slab<FMA_SLAB>(na.x, na.y, na.z, na.w, nc.x, nc.y, r, t_min, t_far, tn0);
            const bool h1 = slab<FMA_SLAB>(nb.x, nb.y, nb.z, nb.w, nc.z, nc.w, r, t_min, t_far, tn1) && nd.y != EMPTY_REF;
            if (h0 && h1) {
                const bool swap = tn1 < tn0;
                cur = swap ? nd.y : nd.x;
                if (sp < STACK_N) {
                    PTB_STACK_AT(sp) = swap ? nd.x : nd.y;
                    ++sp;
                    if (STATS) cn.max_stack = max(cn.max_stack, (uint32_t)sp);
                } else {
                    atomicOr(error_flag, 1u);
                }
            } else if (h0) {
                cur = nd.x;
            } else if (h1) {
                cur = nd.y;
            } else {
                PTB_POP();
            }
        }
        if (cur == EMPTY_REF) break;
        if (test_leaf<CLS, ANY>(k, cur, r, t_min, t_max, best, prims, error_flag)) {
            found = true;
            break;
        }
        PTB_POP();
    }
#undef PTB_POP
#undef PTB_STACK_AT
    if (STATS) cn.prims += prims;
    return ANY ? found : best.id >= 0;
}

// closest hit over both hierarchies (active = false for lanes without a ray)
// SECONDARY: the rays are scattered ones (not camera rays): the sphere class then walks with the branch-free node step
template <bool STATS, bool FMA_SLAB, bool SECONDARY = false>
__device__ __forceinline__ bool closest_hit(const SceneK &k, const RayT &r, float t_min, float t_max, bool active, Best &best,
                                            Counters &cn, uint32_t *error_flag)
{
    best.t = t_max;
    best.t_cull = t_max;
    best.id = -1;
    best.pos = 0;
    best.cls = 0;
    traverse<0, false, STATS, FMA_SLAB, SECONDARY>(k, r, t_min, t_max, active, best, cn, error_flag);
    traverse<1, false, STATS, FMA_SLAB>(k, r, t_min, t_max, active, best, cn, error_flag);
    return best.id >= 0;
}

// ---- sampling spec (oracle/pt_scene_oracle.c: pto_sincos2pi, pto_unit_vector, pto_scatter) ------
__device__ __forceinline__ void sincos2pi(float u, float &sn, float &cs)
{
    const float x = u * 4.0f;
    const int q = (int)x;
    const float f = x - (float)q;
    const float a = (f - 0.5f) * 1.57079637f;
    const float a2 = a * a;
    float ps = 2.75573192e-6f;
    ps = ps * a2 + -1.98412698e-4f;
    ps = ps * a2 + 8.33333333e-3f;
    ps = ps * a2 + -1.66666667e-1f;
    ps = ps * a2 + 1.0f;
    const float s = a * ps;
    float pc = 2.48015873e-5f;
    pc = pc * a2 + -1.38888889e-3f;
    pc = pc * a2 + 4.16666667e-2f;
    pc = pc * a2 + -0.5f;
    pc = pc * a2 + 1.0f;
    const float c = pc;
    const float sb = (s + c) * 0.707106769f;
    const float cb = (c - s) * 0.707106769f;
    switch (q & 3) {
    case 0: sn = sb; cs = cb; break;
    case 1: sn = cb; cs = -sb; break;
    case 2: sn = -sb; cs = -cb; break;
    default: sn = -cb; cs = sb; break;
    }
}

__device__ __forceinline__ void unit_vector(float u1, float u2, float &vx, float &vy, float &vz)
{
    const float z = 1.0f - 2.0f * u1;
    const float r2 = 1.0f - z * z;
    const float rr = sqrtf(fmaxf(r2, 0.0f));
    float sn, cs;
    sincos2pi(u2, sn, cs);
    vx = rr * cs;
    vy = rr * sn;
    vz = z;
}

__device__ __forceinline__ void normalize3f(float &x, float &y, float &z)
{
    const float len = sqrtf(dot3f(x, y, z, x, y, z));
    x = x / len;
    y = y / len;
    z = z / len;
}

// returns false when the path is absorbed; d is replaced by the scattered direction
__device__ __forceinline__ bool scatter(const ptb_material &m, float &dx, float &dy, float &dz, float nox, float noy, float noz,
                                        float u1, float u2, float u3)
{
    const float dn_out = dot3f(dx, dy, dz, nox, noy, noz);
    const bool front = dn_out < 0.0f;
    const float nx = front ? nox : -nox, ny = front ? noy : -noy, nz = front ? noz : -noz;
    float rx, ry, rz;
    if (m.type == PTB_MAT_LAMBERT) {
        float vx, vy, vz;
        unit_vector(u1, u2, vx, vy, vz);
        rx = nx + vx; ry = ny + vy; rz = nz + vz;
        if (fabsf(rx) < 1e-8f && fabsf(ry) < 1e-8f && fabsf(rz) < 1e-8f) { rx = nx; ry = ny; rz = nz; }
    } else if (m.type == PTB_MAT_METAL) {
        float vx, vy, vz;
        unit_vector(u1, u2, vx, vy, vz);
        const float dn = dot3f(dx, dy, dz, nx, ny, nz);
        const float kk = 2.0f * dn;
        float fx = dx - kk * nx, fy = dy - kk * ny, fz = dz - kk * nz;
        normalize3f(fx, fy, fz);
        rx = fx + m.fuzz * vx; ry = fy + m.fuzz * vy; rz = fz + m.fuzz * vz;
        if (!(dot3f(rx, ry, rz, nx, ny, nz) > 0.0f)) return false;
    } else {
        const float ri = front ? 1.0f / m.ior : m.ior;
        const float dn = dot3f(dx, dy, dz, nx, ny, nz);
        const float cs = fminf(-dn, 1.0f);
        const float sn = sqrtf(fmaxf(1.0f - cs * cs, 0.0f));
        const bool cannot = ri * sn > 1.0f;
        float r0 = (1.0f - ri) / (1.0f + ri);
        r0 = r0 * r0;
        const float mm = 1.0f - cs;
        const float m2 = mm * mm;
        const float m5 = (m2 * m2) * mm;
        const float refl = r0 + (1.0f - r0) * m5;
        if (cannot || refl > u3) {
            const float kk = 2.0f * dn;
            rx = dx - kk * nx; ry = dy - kk * ny; rz = dz - kk * nz;
        } else {
            const float px = ri * (dx + cs * nx), py = ri * (dy + cs * ny), pz = ri * (dz + cs * nz);
            const float kk = -sqrtf(fabsf(1.0f - dot3f(px, py, pz, px, py, pz)));
            rx = px + kk * nx; ry = py + kk * ny; rz = pz + kk * nz;
        }
    }
    normalize3f(rx, ry, rz);
    dx = rx; dy = ry; dz = rz;
    return true;
}

// outward geometric normal at the hit (Test.hlsl:28-29 for spheres) and the material id
__device__ __forceinline__ void hit_frame(const SceneK &k, const Best &b, float Px, float Py, float Pz, float &nx, float &ny,
                                          float &nz, uint32_t &mat)
{
    if (b.cls == 0) {
        const float4 cr = __ldg(&k.sph[b.pos]);
        nx = Px - cr.x; ny = Py - cr.y; nz = Pz - cr.z;
        mat = __ldg(&k.sph_meta[b.pos]).y;
    } else {
        const float4 e1 = __ldg(&k.tri[3 * (size_t)b.pos + 1]), e2 = __ldg(&k.tri[3 * (size_t)b.pos + 2]);
        nx = e1.y * e2.z - e1.z * e2.y;
        ny = e1.z * e2.x - e1.x * e2.z;
        nz = e1.x * e2.y - e1.y * e2.x;
        mat = (uint32_t)__float_as_int(e1.w);
    }
    normalize3f(nx, ny, nz);
}

// ---- the per-bounce kernel ----------------------------------------------------------------------
struct WaveArgs {
    PathQueue in, out;
    const uint32_t *n_in;     // paths in `in` (depth >= 1)
    uint32_t *n_out;          // compaction counter of `out`
    uint32_t *cursor;         // work cursor of this launch
    uint32_t n0;              // depth 0: 32 x (number of 8x4 pixel tiles covering this context's row shard)
    uint32_t n0_pixels;       // depth 0: pixels in this context's row shard
    uint32_t tiles_x, rows;   // depth 0: tiles per tile row, rows in the shard
    uint32_t depth, sample;   // sample index of the wave's first sample
    // samples of every pixel traced by this wave (>= 1).  A path's id word is its pixel's index within the row shard
    // (shard row * width + x); batch > 1: | local sample << PIX_BITS, and its radiance goes to
    // slices[local sample][shard index] (one plain store per path, summed in sample order by reduce_slices_kernel —
    // the same additions in the same order as one wave per sample)
    uint32_t batch, tiles;
    float4 *slices;
    int jitter;
    int tail;                 // 1: finish every path of this launch in-kernel (bounce loop, no more queues)
    uint32_t q_cap;           // capacity of the path queues and of the slices buffer (bounds-checked build)
    uint32_t refill;          // bounce_kernel: walkers pause for a refill when fewer than this many lanes of the warp walk
    float4 *accum;
    int32_t *hit_ids;
    SceneStatsDev *stats;
    uint32_t *error_flag;
};

constexpr uint32_t PIX_BITS = 26u, PIX_MASK = (1u << PIX_BITS) - 1u;

// Test.hlsl:7 for pixel (x, y), sample `s_local` of the wave (jittered inside the pixel when the wave asks for it)
__device__ __forceinline__ void camera_uv(const SceneK &k, const WaveArgs &w, uint32_t x, uint32_t y, uint32_t s_local, float &uvx,
                                          float &uvy)
{
    float fx = (float)x, fy = (float)y;
    if (w.jitter) {
        const U3 v = pcg3d(y * k.width + x, w.sample + s_local, k.seed);
        fx = fx + bits_to_unit(v.x);
        fy = fy + bits_to_unit(v.y);
    }
    uvx = (fx / k.div_x) * 2.0f - 1.0f;
    uvy = (fy / k.div_y) * 2.0f - 1.0f;
}

// Sun term of a hit that is about to scatter (spec: oracle/pt_scene_oracle.h "sun"): for a Lambert material lit from its
// front side, c = (throughput * albedo) * (radiance * n.l) with the throughput BEFORE this scatter; returns whether a shadow
// ray has to confirm it.  d = incoming direction, n_out = outward normal.
template <bool SUN>
__device__ __forceinline__ bool sun_term(const SceneK &k, const ptb_material &m, float dx, float dy, float dz, float nox, float noy,
                                         float noz, float tr, float tg, float tb, float &cr, float &cg, float &cb)
{
    if (!SUN || m.type != PTB_MAT_LAMBERT) return false;
    const bool front = dot3f(dx, dy, dz, nox, noy, noz) < 0.0f;
    const float nx = front ? nox : -nox, ny = front ? noy : -noy, nz = front ? noz : -noz;
    const float c = dot3f(nx, ny, nz, k.sun_dir[0], k.sun_dir[1], k.sun_dir[2]);
    if (!(c > 0.0f)) return false;
    cr = (tr * m.albedo[0]) * (k.sun_rad[0] * c);
    cg = (tg * m.albedo[1]) * (k.sun_rad[1] * c);
    cb = (tb * m.albedo[2]) * (k.sun_rad[2] * c);
    return true;
}

// Radiance (cr, cg, cb) of one path segment into the path's slot / its pixel's accumulator.  Without the sun a path adds
// radiance once, when it ends; with it (slices always in use) the slot is the path's running sum L = L + x from +0.
template <bool DEPTH0, bool SUN>
__device__ __forceinline__ void add_radiance(const SceneK &k, const WaveArgs &w, uint32_t s_local, uint32_t loc, uint32_t pix,
                                             float cr, float cg, float cb, bool add)
{
    if (w.slices) {                                 // launch-uniform: batched wave
        PTB_CHK(w.error_flag, (size_t)s_local * w.n0_pixels + loc < w.q_cap && loc < w.n0_pixels, CHK_SLOT);
        float4 *slot = &w.slices[(size_t)s_local * w.n0_pixels + loc];
        if (SUN) {
            if (DEPTH0) {
                __stcs(slot, make_float4(0.0f + cr, 0.0f + cg, 0.0f + cb, 1.0f));
            } else if (add) {
                float4 v = *slot;
                v.x = v.x + cr; v.y = v.y + cg; v.z = v.z + cb;
                *slot = v;
            }
        } else if (DEPTH0 || add) {
            // every path stores its slot at depth 0 (zero radiance while it is still alive) and once more when it ends
            __stcs(slot, make_float4(cr, cg, cb, 1.0f));
        }
    } else if (DEPTH0) {
        PTB_CHK(w.error_flag, pix < k.width * k.height, CHK_PIXEL);
        float4 acc = w.accum[pix];
        acc.x = acc.x + cr; acc.y = acc.y + cg; acc.z = acc.z + cb; acc.w = acc.w + 1.0f;
        w.accum[pix] = acc;
    } else if (add) {
        PTB_CHK(w.error_flag, pix < k.width * k.height, CHK_PIXEL);
        float4 acc = w.accum[pix];
        acc.x = acc.x + cr; acc.y = acc.y + cg; acc.z = acc.z + cb;
        w.accum[pix] = acc;
    }
}

template <bool DEPTH0, bool STATS, bool SUN>
__global__ void __launch_bounds__(256, PTB_BOUNCE_MIN_BLOCKS) bounce_kernel_chunk(const SceneK k, const WaveArgs w)
{
    const unsigned lane = threadIdx.x & 31u;
    const uint32_t n_in = DEPTH0 ? w.n0 : *w.n_in;
    Counters cn;
    uint32_t n_rays = 0, n_scatter = 0;
    for (;;) {
        uint32_t base = 0;
        if (lane == 0) base = atomicAdd(w.cursor, 32u);
        base = __shfl_sync(FULL, base, 0);
        if (base >= n_in) break;
        const uint32_t i = base + lane;
        bool alive = i < n_in;
        uint32_t px = 0, py_shard = 0, s_local = 0;
        if (DEPTH0) {
            // a warp takes one 8x4 pixel tile: primary rays of a warp stay close in BOTH image axes (32x1 strips were
            // measurably less coherent on the 4K terrain), and the compacted queues inherit the locality
            uint32_t tile = i >> 5;
            if (w.batch > 1u) {                         // launch-uniform
                s_local = tile / w.tiles;
                tile -= s_local * w.tiles;
            }
            px = (tile % w.tiles_x) * 8u + (lane & 7u);
            py_shard = (tile / w.tiles_x) * 4u + (lane >> 3);
            alive = alive && px < k.width && py_shard < w.rows;
        }
        uint32_t pix = 0, loc = 0;
        float ox = 0, oy = 0, oz = 0, dx = 0, dy = 0, dz = 1, tr = 1, tg = 1, tb = 1, uvx = 0, uvy = 0;
        if (alive) {
            if (DEPTH0) {
                const uint32_t x = px, yy = py_shard;
                const uint32_t y = shard_row(k.row_offset, k.row_stride, k.row_block, yy);
                pix = y * k.width + x;
                loc = yy * k.width + x;
                float fx = (float)x, fy = (float)y;
                if (w.jitter) {
                    const U3 v = pcg3d(pix, w.sample + s_local, k.seed);
                    fx = fx + bits_to_unit(v.x);
                    fy = fy + bits_to_unit(v.y);
                }
                uvx = (fx / k.div_x) * 2.0f - 1.0f;                      // Test.hlsl:7
                uvy = (fy / k.div_y) * 2.0f - 1.0f;
                dx = (uvx * k.cam_u[0] + uvy * k.cam_v[0]) + k.cam_w[0];   // Test.hlsl:10 with a camera basis
                dy = (uvx * k.cam_u[1] + uvy * k.cam_v[1]) + k.cam_w[1];
                dz = (uvx * k.cam_u[2] + uvy * k.cam_v[2]) + k.cam_w[2];
                normalize3f(dx, dy, dz);
                ox = k.cam_o[0]; oy = k.cam_o[1]; oz = k.cam_o[2];       // Test.hlsl:8
            } else {
                // streaming (evict-first) access: the path state is touched once per bounce and must not
                // push the hierarchy out of L2
                PTB_CHK(w.error_flag, i < w.q_cap, CHK_IN);
                const float4 s0 = __ldcs(&w.in.s0[i]), s1 = __ldcs(&w.in.s1[i]);
                const float2 s2 = __ldcs(&w.in.s2[i]);
                ox = s0.x; oy = s0.y; oz = s0.z;
                loc = __float_as_uint(s0.w);
                if (w.batch > 1u) {                     // launch-uniform
                    s_local = loc >> PIX_BITS;
                    loc &= PIX_MASK;
                }
                const uint32_t yy = loc / k.width;
                pix = shard_row(k.row_offset, k.row_stride, k.row_block, yy) * k.width + (loc - yy * k.width);
                dx = s1.x; dy = s1.y; dz = s1.z; tr = s1.w; tg = s2.x; tb = s2.y;
            }
        }
        uint32_t depth = w.depth;
      bounce_again:
        float cr = 0.0f, cg = 0.0f, cb = 0.0f;     // radiance this path adds now
        bool add = false;
        // closest hit: every lane takes part (lanes past the end of the queue pass alive = false), the warp reconverges behind it
        RayT r;
        make_ray(r, ox, oy, oz, dx, dy, dz);
        Best best;
        const bool hit = closest_hit<STATS, true, !DEPTH0>(k, r, k.t_min, __int_as_float(0x7F800000) /* +inf */, alive, best, cn, w.error_flag);
        __syncwarp();
        const bool shaded = alive;
        bool need_shadow = false;                  // sun: this hit's direct light is added if the shadow ray finds nothing
        if (alive) {
            if (STATS) ++n_rays;
            if (DEPTH0 && w.hit_ids) w.hit_ids[pix] = hit ? best.id : -1;
            float Px = 0, Py = 0, Pz = 0, nx = 0, ny = 0, nz = 0;
            uint32_t mat = 0;
            if (hit) {
                Px = ox + best.t * dx; Py = oy + best.t * dy; Pz = oz + best.t * dz;      // Test.hlsl:28
                hit_frame(k, best, Px, Py, Pz, nx, ny, nz, mat);
            }
            if (k.shade_mode == PTB_SHADE_NORMALS) {
                if (hit) { cr = nx * 0.5f + 0.5f; cg = ny * 0.5f + 0.5f; cb = nz * 0.5f + 0.5f; }   // Test.hlsl:31
                else { cr = uvx; cg = uvy; cb = 0.0f; }                                              // Test.hlsl:36
                add = true;
                alive = false;
            } else if (!hit) {
                const float a = (dy + 1.0f) * 0.5f;
                cr = tr * ((1.0f - a) * k.sky_h[0] + a * k.sky_z[0]);
                cg = tg * ((1.0f - a) * k.sky_h[1] + a * k.sky_z[1]);
                cb = tb * ((1.0f - a) * k.sky_h[2] + a * k.sky_z[2]);
                add = true;
                alive = false;
            } else if (depth == k.max_depth) {
                alive = false;                      // bounce budget spent: contributes nothing
            } else {
                const ptb_material m = k.materials[mat < k.n_materials ? mat : 0u];
                const U3 v = pcg3d(pix, w.sample + s_local, k.seed + (depth + 1u) * 0x9E3779B9u);
                if (STATS) ++n_scatter;
                need_shadow = sun_term<SUN>(k, m, dx, dy, dz, nx, ny, nz, tr, tg, tb, cr, cg, cb);
                if (scatter(m, dx, dy, dz, nx, ny, nz, bits_to_unit(v.x), bits_to_unit(v.y), bits_to_unit(v.z))) {
                    tr = tr * m.albedo[0]; tg = tg * m.albedo[1]; tb = tb * m.albedo[2];
                    ox = Px; oy = Py; oz = Pz;
                } else {
                    alive = false;
                }
            }
        }
        if (SUN && __any_sync(FULL, need_shadow)) {       // shadow rays: any-hit walk, the whole warp takes part
            RayT sr;
            make_ray(sr, ox, oy, oz, k.sun_dir[0], k.sun_dir[1], k.sun_dir[2]);
            Best sb;
            sb.t = __int_as_float(0x7F800000); sb.t_cull = sb.t; sb.id = -1; sb.pos = 0; sb.cls = 0;
            bool occluded = traverse<0, true, STATS, true>(k, sr, k.t_min, sb.t, need_shadow, sb, cn, w.error_flag);
            occluded = traverse<1, true, STATS, true>(k, sr, k.t_min, sb.t, need_shadow && !occluded, sb, cn, w.error_flag) || occluded;
            __syncwarp();
            if (STATS && need_shadow) ++n_rays;
            if (need_shadow && !occluded) add = true;
            else if (need_shadow) cr = cg = cb = 0.0f;
        }
        if (shaded) add_radiance<DEPTH0, SUN>(k, w, s_local, loc, pix, cr, cg, cb, add);
        const unsigned live = __ballot_sync(FULL, alive);
        if (!DEPTH0 && w.tail) {
            // tail mode (launch-uniform): the few paths still alive at this depth finish here, bounce after bounce, with
            // no further launches or queues — deep queues are too small to fill the GPU, so their per-launch latency
            // dominated (profiles/r01_notes.md).  Lanes whose path has ended idle until the warp's last path ends.
            if (live) {
                ++depth;
                goto bounce_again;
            }
            continue;
        }
        // queue compaction: warp ballot + prefix popcount, one atomic per warp
        if (live) {
            uint32_t obase = 0;
            if (lane == 0) obase = atomicAdd(w.n_out, (uint32_t)__popc(live));
            obase = __shfl_sync(FULL, obase, 0);
            if (alive) {
                const uint32_t o = obase + (uint32_t)__popc(live & ((1u << lane) - 1u));
                PTB_CHK(w.error_flag, o < w.q_cap, CHK_QUEUE_OUT);
                __stcs(&w.out.s0[o], make_float4(ox, oy, oz, __uint_as_float(loc | (s_local << PIX_BITS))));
                __stcs(&w.out.s1[o], make_float4(dx, dy, dz, tr));
                __stcs(&w.out.s2[o], make_float2(tg, tb));
            }
        }
    }
    if (STATS) {
        unsigned long long v[4] = {cn.nodes, cn.prims, n_rays, n_scatter};
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int o = 16; o; o >>= 1) v[j] += __shfl_xor_sync(FULL, v[j], o);
        uint32_t ms = cn.max_stack;
#pragma unroll
        for (int o = 16; o; o >>= 1) ms = max(ms, __shfl_xor_sync(FULL, ms, o));
        if (lane == 0) {
            atomicAdd(&w.stats->nodes_visited, v[0]);
            atomicAdd(&w.stats->prims_tested, v[1]);
            atomicAdd(&w.stats->rays, v[2]);
            atomicAdd(&w.stats->scatters, v[3]);
            atomicMax(&w.stats->max_stack, (unsigned long long)ms);
            if (DEPTH0 && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&w.stats->paths, (unsigned long long)w.n0_pixels * w.batch);
        }
    }
}

// ---- the per-bounce kernel: lockstep node steps, per-warp leaf queue, per-lane refill ----------------------------------
// Same work as bounce_kernel_chunk above, other schedule.  There every lane walks its hierarchy on its own: at any moment some
// lanes of a warp test boxes, some test primitives, some have finished, and behind the first bounce the kernel ran at 6-12
// active threads per instruction — boxes at 16, primitives at 6.5 (profiles/r02_notes.md).  Here
//   * a walk is a sequence of NODE STEPS taken by the whole warp together: every walking lane fetches its node, tests the two
//     child boxes, descends / pushes / pops.  A child that is a leaf is not visited: it goes, as a (lane, class, leaf)
//     word, into a queue the warp keeps in shared memory (slots by ballot + popcount: the warp is converged there);
//   * whenever the queue holds 32 leaves the warp tests 32 leaves at once, one per lane, whatever path they belong to:
//     the lane reads the owner's ray from shared memory and folds its leaf's best hit into the owner's closest-hit word with
//     one shared-memory atomicMin on (t bits << 32 | primitive id) — t > t_min >= 0, so the integer order is the order of
//     (t, id) and the tie rule "equal t: the smaller id wins" comes for free.  The owner meanwhile walks on with the culling
//     distance it knew (a stale distance only visits more nodes; what is found is the minimum over a superset);
//   * every LANE owns a path: when fewer than w.refill lanes are still walking, the queue is flushed, the lanes whose walk has
//     ended shade and scatter their paths together, take new paths from the cursor (one atomicAdd per warp for all its free
//     lanes) and the warp walks on with all lanes busy (Aila & Laine 2009, dynamic fetch).  In tail mode a path that
//     scatters stays in its lane and starts its next walk at once: a persistent path tracer until the cursor runs dry.
// Which lane traces a path, in which order primitives are tested and the queues fill changes nothing in any result: every
// path's arithmetic is its own, its radiance goes to its own slot (or its pixel's only live path), the closest hit is a
// minimum.
#ifndef PTB_REFILL_DEFAULT
#define PTB_REFILL_DEFAULT 8
#endif
#ifndef PTB_QUEUE_MIN_BLOCKS
#define PTB_QUEUE_MIN_BLOCKS 4
#endif
constexpr unsigned long long NO_HIT = 0x7F800000FFFFFFFFull;   // t = +inf, id = -1

// Shared memory of bounce_kernel, addressed through 32-bit shared-window addresses and explicit ld/st.shared: left to itself
// the compiler re-derives every address from %tid and the CTA's shared window inside the walk loop (S2R, 6 instructions
// per access) rather than keep a base in a register.
struct QueueShared {
    float ray[8][256];                         // o.xyz, d.xyz, 2 d.d, 4 d.d of the path in each lane
    unsigned long long best[256];              // its closest hit so far: t bits << 32 | id
    uint32_t pos[256];                         // ... leaf-order position | class << 31
    unsigned long long queue[8][32u + 64u];    // per warp: lane << 59 | class << 58 | ~(leaf reference)
};
__device__ __forceinline__ uint32_t lds32(uint32_t a)
{
    uint32_t v;
    asm volatile("ld.shared.b32 %0, [%1];" : "=r"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ float ldsf(uint32_t a) { return __uint_as_float(lds32(a)); }
__device__ __forceinline__ void sts32(uint32_t a, uint32_t v) { asm volatile("st.shared.b32 [%0], %1;" ::"r"(a), "r"(v) : "memory"); }
__device__ __forceinline__ unsigned long long lds64(uint32_t a)
{
    unsigned long long v;
    asm volatile("ld.shared.b64 %0, [%1];" : "=l"(v) : "r"(a) : "memory");
    return v;
}
__device__ __forceinline__ void sts64(uint32_t a, unsigned long long v) { asm volatile("st.shared.b64 [%0], %1;" ::"r"(a), "l"(v) : "memory"); }
__device__ __forceinline__ unsigned long long pack64(uint32_t hi, uint32_t lo)
{
    unsigned long long v;
    asm("mov.b64 %0, {%1, %2};" : "=l"(v) : "r"(lo), "r"(hi));
    return v;
}
__device__ __forceinline__ void atom_min64(uint32_t a, unsigned long long v)
{
    unsigned long long old;
    asm volatile("atom.shared.min.u64 %0, [%1], %2;" : "=l"(old) : "r"(a), "l"(v) : "memory");
}

// What a lane keeps of its ray while it walks: the slab form only.  Origin, direction and the two sphere constants live in
// shared memory (s_ray), where the lane that tests one of this path's leaves reads them; throughput, pixel and depth are
// needed again only when the path is shaded and stay in local memory.
struct WalkRay {
    float ix, iy, iz, px, py, pz;
};

// the FMA form of slab<> on the six numbers a walking lane keeps in registers
__device__ __forceinline__ bool slab_walk(float lox, float hix, float loy, float hiy, float loz, float hiz, const WalkRay &r,
                                          float t_min, float t_far, float &t_near)
{
    const float2 tx = __ffma2_rn(make_float2(lox, hix), make_float2(r.ix, r.ix), make_float2(r.px, r.px));   // see slab<>
    const float2 ty = __ffma2_rn(make_float2(loy, hiy), make_float2(r.iy, r.iy), make_float2(r.py, r.py));
    const float2 tz = __ffma2_rn(make_float2(loz, hiz), make_float2(r.iz, r.iz), make_float2(r.pz, r.pz));
    const float ax = tx.x, bx = tx.y, ay = ty.x, by = ty.y, az = tz.x, bz = tz.y;
    t_near = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), t_min));
    const float tf = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fminf(fmaxf(az, bz), t_far));
    return t_near <= fmaf(fabsf(tf), 1.9073486e-6f, tf);     // widened away from zero whatever the sign of tf
}

template <bool DEPTH0, bool STATS, bool SUN>
__global__ void __launch_bounds__(256, PTB_QUEUE_MIN_BLOCKS) bounce_kernel(const SceneK k, const WaveArgs w)
{
    __shared__ QueueShared S;
    uint32_t tid = threadIdx.x;
    uint32_t sa = (uint32_t)__cvta_generic_to_shared(&S);
    asm volatile("" : "+r"(tid), "+r"(sa));            // keep both in registers (see QueueShared)
    const unsigned lane = tid & 31u;
    const unsigned lt = (1u << lane) - 1u;
    const uint32_t sa_ray = sa + (uint32_t)offsetof(QueueShared, ray), sa_best = sa + (uint32_t)offsetof(QueueShared, best);
    const uint32_t sa_pos = sa + (uint32_t)offsetof(QueueShared, pos);
    const uint32_t sa_queue = sa + (uint32_t)offsetof(QueueShared, queue) + (tid >> 5) * (uint32_t)sizeof(S.queue[0]);
    const uint32_t wbase = tid & ~31u;
    const uint32_t n_in = DEPTH0 ? w.n0 : *w.n_in;
    const int first_cls = k.sph_root != EMPTY_REF ? 0 : 1;
    const int first_root = k.sph_root != EMPTY_REF ? k.sph_root : k.tri_root;      // EMPTY_REF: a scene without primitives
    const int tri_after = first_cls == 0 ? k.tri_root : EMPTY_REF;                 // where a walk goes when the spheres are done
    const float INF = __int_as_float(0x7F800000);
    Counters cn;
    uint32_t n_rays = 0, n_scatter = 0;
    int stack[STACK_N];                                // local memory, as in the chunk schedule (a shared-memory short stack
                                                       // was measured: no faster, and its 12 KB per block come out of L1)
    // throughput r, g, b, id word, depth; while a shadow ray is out: the scattered direction [5..7] and the sun term [8..10]
    uint32_t cold_[11] = {0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u, 0u};
    volatile uint32_t *const cold = cold_;
    bool has = false;                  // this lane holds a path (walking, or walked and not shaded yet)
    bool shadow = false;               // ... and what it walks is the path's shadow ray towards the sun (any hit ends it)
    bool exhausted = false;            // warp-uniform: the cursor has run past the queue
    int cur = EMPTY_REF, sp = 0, cls = 0;
    WalkRay r = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    float t_cull = INF;                // culling distance: the closest hit known at the last queue flush, widened
    uint32_t qn = 0;                   // warp-uniform: leaves waiting in the queue

    // a new ray for the path in this lane: slab form to registers, the rest to shared memory, walk state reset
    auto start_walk = [&](float ox, float oy, float oz, float dx, float dy, float dz) {
        RayT f;
        make_ray(f, ox, oy, oz, dx, dy, dz);
        r.ix = f.ix; r.iy = f.iy; r.iz = f.iz; r.px = f.px; r.py = f.py; r.pz = f.pz;
        const uint32_t ra = sa_ray + tid * 4u;
        sts32(ra, __float_as_uint(ox)); sts32(ra + 1024u, __float_as_uint(oy)); sts32(ra + 2048u, __float_as_uint(oz));
        sts32(ra + 3072u, __float_as_uint(dx)); sts32(ra + 4096u, __float_as_uint(dy)); sts32(ra + 5120u, __float_as_uint(dz));
        sts32(ra + 6144u, __float_as_uint(f.a2)); sts32(ra + 7168u, __float_as_uint(f.a4));
        sts64(sa_best + tid * 8u, NO_HIT);
        t_cull = INF;
        cur = first_root; cls = first_cls; sp = 0;
    };

    // tests the primitives of the top `nb` (<= 32) leaves of the queue, one leaf per lane; every lane of the warp calls it
    auto test_batch = [&](uint32_t nb) {
        __syncwarp();                                        // queue words and rays other lanes have just written
        const bool act = lane < nb;
        const unsigned long long e = act ? lds64(sa_queue + (qn - nb + lane) * 8u) : pack64(lane << 27, 0u);
        const uint32_t ehi = (uint32_t)(e >> 32);
        const unsigned own = wbase + (ehi >> 27);
        const uint32_t ra = sa_ray + own * 4u;
        RayT er;
        er.ox = ldsf(ra); er.oy = ldsf(ra + 1024u); er.oz = ldsf(ra + 2048u);
        er.dx = ldsf(ra + 3072u); er.dy = ldsf(ra + 4096u); er.dz = ldsf(ra + 5120u);
        er.a2 = ldsf(ra + 6144u); er.a4 = ldsf(ra + 7168u);
        const uint32_t c = (ehi >> 26) & 1u, q = (uint32_t)e;
        const uint32_t first = q & LEAF_FIRST_MASK, count = act ? (q >> LEAF_COUNT_SHIFT) + 1u : 0u;
        unsigned long long key = NO_HIT;
        uint32_t pos = 0;
        PTB_CHK(w.error_flag, first + count <= (c == 0u ? k.n_spheres : k.n_tris) && own < 256u, CHK_PRIM);
        for (uint32_t i = first; i < first + count; ++i) {
            float t;
            int id;
            if (c == 0u) {
                const float4 cr = __ldg(&k.sph[i]);
                if (!hit_sphere(cr, er, k.t_min, INF, t)) continue;
                id = (int)__ldg(&k.sph_meta[i]).x;
            } else {
                const float4 v0 = __ldg(&k.tri[3 * (size_t)i]), e1 = __ldg(&k.tri[3 * (size_t)i + 1]),
                             e2 = __ldg(&k.tri[3 * (size_t)i + 2]);
                if (!hit_triangle(v0, e1, e2, er, k.t_min, INF, t)) continue;
                id = __float_as_int(v0.w);
            }
            const unsigned long long kk = ((unsigned long long)__float_as_uint(t) << 32) | (uint32_t)id;
            if (kk < key) {
                key = kk;
                pos = i;
            }
        }
        const bool hit = key < lds64(sa_best + own * 8u);
        if (hit) atom_min64(sa_best + own * 8u, key);
        __syncwarp();
        if (hit && lds64(sa_best + own * 8u) == key) sts32(sa_pos + own * 4u, pos | (c << 31));
        qn -= nb;
    };

    for (;;) {
        // ---- refill: every lane without a path takes the next one of the queue ----
        const unsigned idle = __ballot_sync(FULL, !has);
        if (idle != 0u && !exhausted) {                      // warp-uniform
            const uint32_t n = (uint32_t)__popc(idle);
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(w.cursor, n);
            base = __shfl_sync(FULL, base, 0);
            exhausted = base >= n_in || n_in - base <= n;
            const uint32_t i = base + (uint32_t)__popc(idle & lt);
            if (!has && i < n_in) {
                if (DEPTH0) {
                    // path i = lane (i & 31) of 8x4 pixel tile (i >> 5): primary rays fetched together stay close in BOTH
                    // image axes (32x1 strips were measurably less coherent on the 4K terrain), and the compacted queues
                    // inherit the locality
                    uint32_t tile = i >> 5, s_local = 0;
                    if (w.batch > 1u) {                      // launch-uniform
                        s_local = tile / w.tiles;
                        tile -= s_local * w.tiles;
                    }
                    const uint32_t x = (tile % w.tiles_x) * 8u + (i & 7u), yy = (tile / w.tiles_x) * 4u + ((i >> 3) & 3u);
                    if (x < k.width && yy < w.rows) {
                        const uint32_t y = shard_row(k.row_offset, k.row_stride, k.row_block, yy);
                        float uvx, uvy;
                        camera_uv(k, w, x, y, s_local, uvx, uvy);
                        float dx = (uvx * k.cam_u[0] + uvy * k.cam_v[0]) + k.cam_w[0];   // Test.hlsl:10 with a camera basis
                        float dy = (uvx * k.cam_u[1] + uvy * k.cam_v[1]) + k.cam_w[1];
                        float dz = (uvx * k.cam_u[2] + uvy * k.cam_v[2]) + k.cam_w[2];
                        normalize3f(dx, dy, dz);
                        start_walk(k.cam_o[0], k.cam_o[1], k.cam_o[2], dx, dy, dz);      // Test.hlsl:8
                        cold[0] = cold[1] = cold[2] = __float_as_uint(1.0f);
                        cold[3] = (yy * k.width + x) | (s_local << PIX_BITS);
                        has = true;
                    }
                } else {
                    // streaming (evict-first) access: the path state is touched once per bounce and must not
                    // push the hierarchy out of L2
                    PTB_CHK(w.error_flag, i < w.q_cap, CHK_IN);
                const float4 s0 = __ldcs(&w.in.s0[i]), s1 = __ldcs(&w.in.s1[i]);
                    const float2 s2 = __ldcs(&w.in.s2[i]);
                    start_walk(s0.x, s0.y, s0.z, s1.x, s1.y, s1.z);
                    cold[0] = __float_as_uint(s1.w); cold[1] = __float_as_uint(s2.x); cold[2] = __float_as_uint(s2.y);
                    cold[3] = __float_as_uint(s0.w);
                    has = true;
                }
                if (has) cold[4] = w.depth;
            }
        }
        const unsigned walkers0 = __ballot_sync(FULL, cur != EMPTY_REF);
        if (__ballot_sync(FULL, has) == 0u) {
            if (exhausted) break;
            continue;                                        // depth 0: a fetch of tile padding only
        }

        // ---- walk, in node steps of the whole warp, until too few lanes are still walking ----
        // pausing pays only if the paused warp can do something else: shade-and-continue (tail) or fetch (queue not dry)
        const int pause_below = (exhausted && !w.tail) ? 0 : min((int)w.refill, __popc(walkers0));
        for (;;) {
            // a lane walks while cur != EMPTY_REF.  c0 / c1: what this step found — cur itself when it is a leaf (a hierarchy
            // that is one leaf), else the children whose boxes the ray hits, the nearer one first
            int c0 = cur, c1 = EMPTY_REF;
            const uint32_t tag = (lane << 27) | ((uint32_t)cls << 26);
            if (cur >= 0) {
                const BvhNode *__restrict__ nodes = cls == 0 ? k.sph_nodes : k.tri_nodes;
                PTB_CHK(w.error_flag, (uint32_t)cur < (cls == 0 ? k.n_sph_nodes : k.n_tri_nodes), CHK_NODE);
                float4 na, nb, nc;
                int2 nd;
                load_node(&nodes[cur], na, nb, nc, nd);
                if (STATS) ++cn.nodes;
                float tn0, tn1;
                const bool h0 = slab_walk(na.x, na.y, na.z, na.w, nc.x, nc.y, r, k.t_min, t_cull, tn0);
                const bool h1 = slab_walk(nb.x, nb.y, nb.z, nb.w, nc.z, nc.w, r, k.t_min, t_cull, tn1);
                const bool swap = h0 && h1 && tn1 < tn0;
                const int r0 = h0 ? nd.x : EMPTY_REF, r1 = h1 ? nd.y : EMPTY_REF;   // nd.y == EMPTY_REF: no second child
                c0 = swap ? r1 : r0;
                c1 = swap ? r0 : r1;
            }
            // children that are leaves go to the queue, nodes to the walk.  As unsigned numbers node indices lie below
            // EMPTY_REF (0x80000000) and leaf references above it.
            const uint32_t u0 = (uint32_t)c0, u1 = (uint32_t)c1;
            const bool l0 = u0 > 0x80000000u, l1 = u1 > 0x80000000u;
            const int n0 = (int)min(u0, 0x80000000u), n1 = (int)min(u1, 0x80000000u);     // the node, or EMPTY_REF
            if (cur != EMPTY_REF) {
                if ((n0 | n1) >= 0) {                        // two nodes: the farther one waits on the stack
                    if (sp < STACK_N) {
                        stack[sp++] = n1;
                        if (STATS) cn.max_stack = max(cn.max_stack, (uint32_t)sp);
                    } else {
                        atomicOr(w.error_flag, 1u);
                    }
                }
                cur = n0 >= 0 ? n0 : n1;
                if (cur == EMPTY_REF) {
                    if (sp > 0) {
                        cur = stack[--sp];
                    } else if (cls == 0) {                   // spheres done: the triangle hierarchy next (or EMPTY_REF: finished)
                        cls = 1;
                        cur = tri_after;
                    }
                }
            }
            // the warp is converged here: queue slots by ballot, no atomics
            const unsigned m0 = __ballot_sync(FULL, l0), m1 = __ballot_sync(FULL, l1);
            const uint32_t at0 = qn + __popc(m0 & lt), at1 = qn + __popc(m0) + __popc(m1 & lt);
            PTB_CHK(w.error_flag, (!l0 || at0 < 96u) && (!l1 || at1 < 96u) && (cur == EMPTY_REF || sp <= STACK_N), CHK_LEAFQ);
            if (l0) sts64(sa_queue + at0 * 8u, pack64(tag, ~u0));
            if (l1) sts64(sa_queue + at1 * 8u, pack64(tag, ~u1));
            qn += __popc(m0) + __popc(m1);
            if (STATS) cn.prims += (l0 ? (~u0 >> LEAF_COUNT_SHIFT) + 1u : 0u) + (l1 ? (~u1 >> LEAF_COUNT_SHIFT) + 1u : 0u);
            if (qn >= 32u) {                                 // warp-uniform
                do {
                    test_batch(32u);
                } while (qn >= 32u);
                __syncwarp();
                const float tb_ = ldsf(sa_best + tid * 8u + 4u);
                t_cull = fmaf(fabsf(tb_), 1.9073486e-6f, tb_);
                if (SUN && shadow && tb_ < INF) cur = EMPTY_REF;     // any hit: the sun is occluded, nothing more to learn
            }
            const unsigned wk = __ballot_sync(FULL, cur != EMPTY_REF);
            if (wk == 0u || __popc(wk) < pause_below) break;
        }
        if (qn != 0u) test_batch(qn);                        // flush: the paths about to be shaded need every result
        __syncwarp();
        if (SUN && shadow && cur != EMPTY_REF && ldsf(sa_best + tid * 8u + 4u) < INF) cur = EMPTY_REF;

        // ---- shade and scatter the paths whose walk has ended ----
        bool cont = false;                                   // goes to the next bounce's queue
        float Px = 0, Py = 0, Pz = 0, dx = 0, dy = 0, dz = 0, tr = 0, tg = 0, tb = 0;
        uint32_t idw = 0;
        if (SUN && has && cur == EMPTY_REF && shadow) {
            // the shadow ray of a Lambert hit has finished: add the sun term if nothing was hit, then the path goes on
            // with the scattered ray that has been waiting (origin = the shadow ray's)
            shadow = false;
            idw = cold[3];
            const uint32_t s_local = w.batch > 1u ? idw >> PIX_BITS : 0u;
            const uint32_t loc = w.batch > 1u ? idw & PIX_MASK : idw;
            if (STATS) ++n_rays;
            if (lds64(sa_best + tid * 8u) == NO_HIT)
                add_radiance<false, SUN>(k, w, s_local, loc, 0u, __uint_as_float(cold[8]), __uint_as_float(cold[9]), __uint_as_float(cold[10]), true);
            const uint32_t ra = sa_ray + tid * 4u;
            Px = ldsf(ra); Py = ldsf(ra + 1024u); Pz = ldsf(ra + 2048u);
            dx = __uint_as_float(cold[5]); dy = __uint_as_float(cold[6]); dz = __uint_as_float(cold[7]);
            tr = __uint_as_float(cold[0]); tg = __uint_as_float(cold[1]); tb = __uint_as_float(cold[2]);
            if (!DEPTH0 && w.tail) {
                start_walk(Px, Py, Pz, dx, dy, dz);
                cold[4] = cold[4] + 1u;
            } else {
                cont = true;
                has = false;
            }
        } else if (has && cur == EMPTY_REF) {
            Best best;
            const unsigned long long bk = lds64(sa_best + tid * 8u);
            const uint32_t bp = lds32(sa_pos + tid * 4u);
            best.t = __uint_as_float((uint32_t)(bk >> 32));
            best.t_cull = best.t;
            best.id = (int)(uint32_t)bk;
            best.pos = (int)(bp & 0x7FFFFFFFu);
            best.cls = (int)(bp >> 31);
            const bool hit = best.id >= 0;
            tr = __uint_as_float(cold[0]); tg = __uint_as_float(cold[1]); tb = __uint_as_float(cold[2]);
            idw = cold[3];
            const uint32_t depth = cold[4];
            const uint32_t s_local = w.batch > 1u ? idw >> PIX_BITS : 0u;
            const uint32_t loc = w.batch > 1u ? idw & PIX_MASK : idw;
            const uint32_t yy = loc / k.width, x = loc - yy * k.width;
            const uint32_t y = shard_row(k.row_offset, k.row_stride, k.row_block, yy);
            const uint32_t pix = y * k.width + x;
            float cr = 0.0f, cg = 0.0f, cb = 0.0f;           // radiance this path adds now
            float sr = 0.0f, sg = 0.0f, sb = 0.0f;           // ... and may add once its shadow ray has come back empty
            bool add = false, alive = true, need_shadow = false;
            if (STATS) ++n_rays;
            if (DEPTH0 && w.hit_ids) w.hit_ids[pix] = hit ? best.id : -1;
            float nx = 0, ny = 0, nz = 0;
            uint32_t mat = 0;
            const uint32_t ra = sa_ray + tid * 4u;
            dx = ldsf(ra + 3072u); dy = ldsf(ra + 4096u); dz = ldsf(ra + 5120u);
            if (hit) {
                Px = ldsf(ra) + best.t * dx; Py = ldsf(ra + 1024u) + best.t * dy; Pz = ldsf(ra + 2048u) + best.t * dz;   // Test.hlsl:28
                hit_frame(k, best, Px, Py, Pz, nx, ny, nz, mat);
            }
            if (k.shade_mode == PTB_SHADE_NORMALS) {
                if (hit) {
                    cr = nx * 0.5f + 0.5f; cg = ny * 0.5f + 0.5f; cb = nz * 0.5f + 0.5f;         // Test.hlsl:31
                } else {
                    camera_uv(k, w, x, y, s_local, cr, cg);                                      // Test.hlsl:36
                    cb = 0.0f;
                }
                add = true;
                alive = false;
            } else if (!hit) {
                const float a = (dy + 1.0f) * 0.5f;
                cr = tr * ((1.0f - a) * k.sky_h[0] + a * k.sky_z[0]);
                cg = tg * ((1.0f - a) * k.sky_h[1] + a * k.sky_z[1]);
                cb = tb * ((1.0f - a) * k.sky_h[2] + a * k.sky_z[2]);
                add = true;
                alive = false;
            } else if (depth == k.max_depth) {
                alive = false;                               // bounce budget spent: contributes nothing
            } else {
                const ptb_material m = k.materials[mat < k.n_materials ? mat : 0u];
                const U3 v = pcg3d(pix, w.sample + s_local, k.seed + (depth + 1u) * 0x9E3779B9u);
                if (STATS) ++n_scatter;
                need_shadow = sun_term<SUN>(k, m, dx, dy, dz, nx, ny, nz, tr, tg, tb, sr, sg, sb);
                if (scatter(m, dx, dy, dz, nx, ny, nz, bits_to_unit(v.x), bits_to_unit(v.y), bits_to_unit(v.z))) {
                    tr = tr * m.albedo[0]; tg = tg * m.albedo[1]; tb = tb * m.albedo[2];
                } else {
                    alive = false;
                }
            }
            add_radiance<DEPTH0, SUN>(k, w, s_local, loc, pix, cr, cg, cb, add);
            if (!alive) {
                has = false;
            } else if (need_shadow) {                        // the shadow ray first; the scattered ray waits in local memory
                cold[0] = __float_as_uint(tr); cold[1] = __float_as_uint(tg); cold[2] = __float_as_uint(tb);
                cold[5] = __float_as_uint(dx); cold[6] = __float_as_uint(dy); cold[7] = __float_as_uint(dz);
                cold[8] = __float_as_uint(sr); cold[9] = __float_as_uint(sg); cold[10] = __float_as_uint(sb);
                start_walk(Px, Py, Pz, k.sun_dir[0], k.sun_dir[1], k.sun_dir[2]);
                shadow = true;
            } else if (!DEPTH0 && w.tail) {                  // launch-uniform: the path stays in this lane
                start_walk(Px, Py, Pz, dx, dy, dz);
                cold[0] = __float_as_uint(tr); cold[1] = __float_as_uint(tg); cold[2] = __float_as_uint(tb);
                cold[4] = depth + 1u;
            } else {
                cont = true;
                has = false;
            }
        }
        // queue compaction: warp ballot + prefix popcount, one atomic per warp
        const unsigned outm = __ballot_sync(FULL, cont);
        if (outm != 0u) {
            uint32_t obase = 0;
            if (lane == 0) obase = atomicAdd(w.n_out, (uint32_t)__popc(outm));
            obase = __shfl_sync(FULL, obase, 0);
            if (cont) {
                const uint32_t o = obase + (uint32_t)__popc(outm & lt);
                PTB_CHK(w.error_flag, o < w.q_cap, CHK_QUEUE_OUT);
                __stcs(&w.out.s0[o], make_float4(Px, Py, Pz, __uint_as_float(idw)));
                __stcs(&w.out.s1[o], make_float4(dx, dy, dz, tr));
                __stcs(&w.out.s2[o], make_float2(tg, tb));
            }
        }
    }
    if (STATS) {
        unsigned long long v[4] = {cn.nodes, cn.prims, n_rays, n_scatter};
#pragma unroll
        for (int j = 0; j < 4; ++j)
#pragma unroll
            for (int o = 16; o; o >>= 1) v[j] += __shfl_xor_sync(FULL, v[j], o);
        uint32_t ms = cn.max_stack;
#pragma unroll
        for (int o = 16; o; o >>= 1) ms = max(ms, __shfl_xor_sync(FULL, ms, o));
        if (lane == 0) {
            atomicAdd(&w.stats->nodes_visited, v[0]);
            atomicAdd(&w.stats->prims_tested, v[1]);
            atomicAdd(&w.stats->rays, v[2]);
            atomicAdd(&w.stats->scatters, v[3]);
            atomicMax(&w.stats->max_stack, (unsigned long long)ms);
            if (DEPTH0 && blockIdx.x == 0 && threadIdx.x == 0) atomicAdd(&w.stats->paths, (unsigned long long)w.n0_pixels * w.batch);
        }
    }
}

// ---- upload helpers -------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) sphere_bounds_kernel(const float4 *__restrict__ sph, uint32_t n, float4 *lo, float4 *hi)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 c = sph[i];
    const float r = fabsf(c.w);
    lo[i] = make_float4(c.x - r, c.y - r, c.z - r, 0.0f);
    hi[i] = make_float4(c.x + r, c.y + r, c.z + r, 0.0f);
}

__global__ void __launch_bounds__(256) tri_bounds_kernel(const float *__restrict__ v, uint32_t n, float4 *lo, float4 *hi)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float *p = v + 9 * (size_t)i;
    lo[i] = make_float4(fminf(p[0], fminf(p[3], p[6])), fminf(p[1], fminf(p[4], p[7])), fminf(p[2], fminf(p[5], p[8])), 0.0f);
    hi[i] = make_float4(fmaxf(p[0], fmaxf(p[3], p[6])), fmaxf(p[1], fmaxf(p[4], p[7])), fmaxf(p[2], fmaxf(p[5], p[8])), 0.0f);
}

__global__ void __launch_bounds__(256) pack_spheres_kernel(const float4 *__restrict__ in, const uint32_t *__restrict__ mat,
                                                           const uint32_t *__restrict__ order, uint32_t n, float4 *out, uint2 *meta)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t src = order[i];
    out[i] = in[src];
    meta[i] = make_uint2(src, mat[src]);
}

__global__ void __launch_bounds__(256) pack_tris_kernel(const float *__restrict__ v, const uint32_t *__restrict__ mat,
                                                        const uint32_t *__restrict__ order, uint32_t n, uint32_t id_base, float4 *out)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t src = order[i];
    const float *p = v + 9 * (size_t)src;
    out[3 * (size_t)i + 0] = make_float4(p[0], p[1], p[2], __int_as_float((int)(id_base + src)));
    out[3 * (size_t)i + 1] = make_float4(p[3] - p[0], p[4] - p[1], p[5] - p[2], __int_as_float((int)mat[src]));
    out[3 * (size_t)i + 2] = make_float4(p[6] - p[0], p[7] - p[1], p[8] - p[2], 0.0f);
}

// Batched waves: accum[pixel] += slices[0][pixel], then slices[1][pixel], ... in sample order — the additions one wave
// per sample performs (depth 0 adds (0,0,0,1) or the radiance, a later bounce adds the radiance: x + 0 first, then + c).
__global__ void __launch_bounds__(256) reduce_slices_kernel(const SceneK k, float4 *accum, const float4 *__restrict__ slices,
                                                            uint32_t batch, uint32_t n_shard)
{
    for (uint32_t j = blockIdx.x * blockDim.x + threadIdx.x; j < n_shard; j += gridDim.x * blockDim.x) {
        const uint32_t yy = j / k.width, x = j - yy * k.width;
        const uint32_t pix = shard_row(k.row_offset, k.row_stride, k.row_block, yy) * k.width + x;
        float4 acc = accum[pix];
        for (uint32_t s = 0; s < batch; ++s) {
            const float4 c = __ldcs(&slices[(size_t)s * n_shard + j]);
            acc.x = (acc.x + 0.0f) + c.x; acc.y = (acc.y + 0.0f) + c.y; acc.z = (acc.z + 0.0f) + c.z;
            acc.w = acc.w + 1.0f;
        }
        accum[pix] = acc;
    }
}

__global__ void __launch_bounds__(256) frame_finish_kernel(const float4 *__restrict__ accum, float4 *f32_out, uint32_t *rgba8_out,
                                                           uint32_t n)
{
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        float4 c = accum[i];
        c.w = 1.0f;
        if (f32_out) f32_out[i] = c;
        rgba8_out[i] = pack_rgba8(c.x, c.y, c.z, c.w);
    }
}

template <bool STATS>
__global__ void __launch_bounds__(256) trace_rays_kernel(const SceneK k, uint64_t n, const float *__restrict__ o,
                                                         const float *__restrict__ d, float t_min, float t_max, int mode,
                                                         int32_t *out_prim, float *out_t, uint32_t *error_flag)
{
    const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < n;                          // no early return: lanes without a ray just skip the walk
    const uint64_t j = active ? i : 0;
    RayT r;
    make_ray(r, o[3 * j], o[3 * j + 1], o[3 * j + 2], d[3 * j], d[3 * j + 1], d[3 * j + 2]);
    Counters cn;
    Best best;
    if (mode == 0) {                                    // launch-uniform
        const bool hit = closest_hit<STATS, false>(k, r, t_min, t_max, active, best, cn, error_flag);
        if (active) {
            out_prim[i] = hit ? best.id : -1;
            if (out_t) out_t[i] = hit ? best.t : 0.0f;
        }
    } else {
        best.t = t_max; best.t_cull = t_max; best.id = -1; best.pos = 0; best.cls = 0;
        bool hit = traverse<0, true, STATS, false>(k, r, t_min, t_max, active, best, cn, error_flag);
        hit = traverse<1, true, STATS, false>(k, r, t_min, t_max, active && !hit, best, cn, error_flag) || hit;
        if (active) out_prim[i] = hit ? 1 : 0;
    }
}

int g_wave_blocks[2][2][2][2] = {};

template <bool DEPTH0, bool STATS, bool SUN, bool CHUNK>
int wave_blocks()
{
    int &cached = g_wave_blocks[DEPTH0][STATS][SUN][CHUNK];
    if (cached == 0) {
        int per_sm = 0, dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        const cudaError_t e = CHUNK ? cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bounce_kernel_chunk<DEPTH0, STATS, SUN>, 256, 0)
                                    : cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, bounce_kernel<DEPTH0, STATS, SUN>, 256, 0);
        if (e != cudaSuccess || per_sm < 1) per_sm = 1;
        cached = per_sm * sms;
    }
    return cached;
}

template <bool DEPTH0, bool STATS, bool SUN>
void launch_bounce_t(const SceneK &k, const WaveArgs &w, bool chunk, cudaStream_t s)
{
    if (chunk) bounce_kernel_chunk<DEPTH0, STATS, SUN><<<wave_blocks<DEPTH0, STATS, SUN, true>(), 256, 0, s>>>(k, w);
    else bounce_kernel<DEPTH0, STATS, SUN><<<wave_blocks<DEPTH0, STATS, SUN, false>(), 256, 0, s>>>(k, w);
}

template <bool DEPTH0, bool STATS>
void launch_bounce(const SceneK &k, const WaveArgs &w, bool chunk, cudaStream_t s)
{
    if (k.sun_on) launch_bounce_t<DEPTH0, STATS, true>(k, w, chunk, s);
    else launch_bounce_t<DEPTH0, STATS, false>(k, w, chunk, s);
}

}  // namespace

cudaError_t launch_sphere_bounds(const float4 *sph, uint32_t n, float4 *lo, float4 *hi, cudaStream_t s)
{
    if (n) sphere_bounds_kernel<<<(n + 255u) / 256u, 256, 0, s>>>(sph, n, lo, hi);
    return cudaGetLastError();
}

cudaError_t launch_tri_bounds(const float *verts9, uint32_t n, float4 *lo, float4 *hi, cudaStream_t s)
{
    if (n) tri_bounds_kernel<<<(n + 255u) / 256u, 256, 0, s>>>(verts9, n, lo, hi);
    return cudaGetLastError();
}

cudaError_t launch_pack_spheres(const float4 *sph_in, const uint32_t *mat_in, const uint32_t *order, uint32_t n, float4 *sph_out,
                                uint2 *meta_out, cudaStream_t s)
{
    if (n) pack_spheres_kernel<<<(n + 255u) / 256u, 256, 0, s>>>(sph_in, mat_in, order, n, sph_out, meta_out);
    return cudaGetLastError();
}

cudaError_t launch_pack_tris(const float *verts9, const uint32_t *mat_in, const uint32_t *order, uint32_t n, uint32_t id_base,
                             float4 *tri_out, cudaStream_t s)
{
    if (n) pack_tris_kernel<<<(n + 255u) / 256u, 256, 0, s>>>(verts9, mat_in, order, n, id_base, tri_out);
    return cudaGetLastError();
}

cudaError_t launch_scene_wave(const SceneK &k, SceneState *st, float4 *accum, int32_t *hit_ids, uint32_t sample, uint32_t batch,
                              bool jitter, cudaStream_t s, uint64_t *launches)
{
    const uint32_t rows = shard_row_count(k.height, k.row_offset, k.row_stride, k.row_block);
    const uint32_t n0_pixels = rows * k.width;
    if (n0_pixels == 0) return cudaSuccess;
    const uint32_t tiles_x = (k.width + 7u) / 8u, tiles_y = (rows + 3u) / 4u;
    const uint64_t n0_64 = (uint64_t)tiles_x * tiles_y * 32u * batch;
    if (n0_64 > 0xFFF00000ull) return cudaErrorInvalidValue;      // scene.cu sizes waves so that this cannot happen (the
                                                                  // margin: every warp runs the cursor past the end once)
    const uint32_t n0 = (uint32_t)n0_64;
    const uint32_t depths = k.shade_mode == PTB_SHADE_NORMALS ? 1u : k.max_depth + 1u;
    // per-path radiance slots: batched waves, and every wave when the sun is on (a path then adds radiance several times)
    const bool use_slices = batch > 1u || k.sun_on;
    if (use_slices && !st->slices) return cudaErrorInvalidValue;
    // bounces >= tail_depth run inside ONE launch (0 disables): queues that deep are too small to fill the GPU.  With the
    // queue schedule in the tail launch the tail starts at bounce 3 (c4: 3, 4, 5, 6 -> 1,433 / 1,422 / 1,415 / 1,412
    // Msamples/s); with the chunk schedule there, see below.  PTB200_TAIL_DEPTH overrides for A/B runs.
    const char *e_tail = std::getenv("PTB200_TAIL_DEPTH");
    const int tail_env = e_tail ? std::atoi(e_tail) : -1;
    // A/B switches for profiling runs: PTB200_BOUNCE=chunk selects the 32-paths-per-warp schedule, PTB200_REFILL the number
    // of walking lanes below which a warp pauses to shade and refetch (0: never).
    const char *e_bounce = std::getenv("PTB200_BOUNCE"), *e_refill = std::getenv("PTB200_REFILL");
    // PTB200_BOUNCE: "chunk" = every launch with the chunk schedule, "queue" = bounces >= 1 with the queue schedule,
    // "tail" = only the tail launch with the queue schedule (there a path that scatters stays in its lane; on the big
    // early bounces its node step costs twice the instructions of the chunk schedule's for 1.5-2x the active lanes;
    // profiles/r02_notes.md)
    // Unset: by scene.  With triangles, the tail launch runs the queue schedule (c4 1,433 against 1,407-1,413 Msamples/s with
    // chunks everywhere).  A scene of spheres only stays on the chunk schedule throughout: since the sphere class walks its
    // scattered rays with the branch-free node step (traverse<>), chunks beat the queue-schedule tail at every wave size
    // (c3, 1 / 2 / 4 / 8 / 32 samples per wave: 3,594 / 4,048 / 4,364 / 4,565 / 4,686 against 3,281 / 3,763 / 4,159 / 4,377 / 4,511).
    const bool spheres_only = k.n_tris == 0u;
    const bool q_all = e_bounce && e_bounce[0] == 'q';
    const bool q_tail = e_bounce ? (q_all || e_bounce[0] == 't') : !spheres_only;
    const bool chunk_d0 = !std::getenv("PTB200_D0_QUEUE");
    const int refill_i = e_refill ? std::atoi(e_refill) : PTB_REFILL_DEFAULT;
    const uint32_t refill = (uint32_t)(refill_i < 0 ? 0 : refill_i > 32 ? 32 : refill_i);
    // chunk-schedule tail: from bounce 6, or 8 for waves of >= 16 Mi camera rays, whose queues stay large to the end
    // (c3, tail from 4 / 6 / never: 1 sample per wave 3,598 / 3,594 / 3,471, 8 per wave 4,523 (6) / 4,565 (never))
    const uint32_t tail_depth = tail_env >= 0 ? (uint32_t)tail_env : (q_tail ? 3u : ((uint64_t)n0_pixels * batch >= (16u << 20) ? 8u : 6u));
    // counters: [0, depths] queue sizes, [depths+1, 2*depths+1] work cursors
    cudaError_t e = cudaMemsetAsync(st->counts, 0, sizeof(uint32_t) * st->counts_len, s);
    if (e != cudaSuccess) return e;
    for (uint32_t d = 0; d < depths; ++d) {
        WaveArgs w{};
        w.in = st->q[(d + 1u) & 1u];
        w.out = st->q[d & 1u];
        w.n_in = st->counts + d;
        w.n_out = st->counts + d + 1u;
        w.cursor = st->counts + (k.max_depth + 2u) + d;
        w.n0 = n0;
        w.n0_pixels = n0_pixels;
        w.tiles_x = tiles_x;
        w.rows = rows;
        w.depth = d;
        w.sample = sample;
        w.batch = batch;
        w.tiles = tiles_x * tiles_y;
        w.slices = use_slices ? st->slices : nullptr;
        w.jitter = jitter ? 1 : 0;
        w.tail = (tail_depth > 0 && d >= tail_depth && d + 1u < depths) ? 1 : 0;
        w.accum = accum;
        w.hit_ids = hit_ids;
        w.stats = st->stats;
        w.error_flag = st->error_flag;
        w.refill = refill;
        w.q_cap = st->q_cap;
        if (d == 0) {
            // camera rays are coherent: the chunk schedule already runs them at 28 of 32 lanes, without the queue's overheads
            if (st->count_stats) launch_bounce<true, true>(k, w, chunk_d0, s);
            else launch_bounce<true, false>(k, w, chunk_d0, s);
        } else {
            const bool chunk = w.tail ? !q_tail : !q_all;
            if (st->count_stats) launch_bounce<false, true>(k, w, chunk, s);
            else launch_bounce<false, false>(k, w, chunk, s);
        }
        if (launches) ++*launches;
        if (w.tail) break;
    }
    if (use_slices) {
        const unsigned blocks = min((n0_pixels + 255u) / 256u, 148u * 8u);
        reduce_slices_kernel<<<blocks, 256, 0, s>>>(k, accum, st->slices, batch, n0_pixels);
        if (launches) ++*launches;
    }
    return cudaGetLastError();
}

cudaError_t launch_frame_finish(const float4 *accum, float4 *f32_out, uint32_t *rgba8_out, uint32_t n_pixels, cudaStream_t s)
{
    const unsigned blocks = min((n_pixels + 255u) / 256u, 148u * 8u);
    frame_finish_kernel<<<blocks, 256, 0, s>>>(accum, f32_out, rgba8_out, n_pixels);
    return cudaGetLastError();
}

cudaError_t launch_trace_rays(const SceneK &k, SceneState *st, uint64_t n, const float *o, const float *d, float t_min, float t_max,
                              int mode, int32_t *out_prim, float *out_t, cudaStream_t s)
{
    if (n == 0) return cudaSuccess;
    const unsigned blocks = (unsigned)((n + 255u) / 256u);
    if (st->count_stats) trace_rays_kernel<true><<<blocks, 256, 0, s>>>(k, n, o, d, t_min, t_max, mode, out_prim, out_t, st->error_flag);
    else trace_rays_kernel<false><<<blocks, 256, 0, s>>>(k, n, o, d, t_min, t_max, mode, out_prim, out_t, st->error_flag);
    return cudaGetLastError();
}

}  // namespace ptb
