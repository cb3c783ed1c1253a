// scene_build.cu — GPU construction of the flattened two-child BVH (north_star stage n4; the
// reference has no hierarchy of any kind — whole-tree grep, SURVEY.md §0 — so this is new surface).
//
// Pipeline, all on the context's stream, no host round trips except the final node count:
//   1. centroid bounds of the primitive boxes            (warp-shuffle reduce + ordered-int atomics)
//   2. 63-bit Morton code of every centroid (21 bits/axis)
//   3. cub::DeviceRadixSort::SortPairs(code, primitive index)
//   4. Karras 2012 hierarchy: one thread per internal node finds its key range and split
//   5. bottom-up refit with per-node arrival counters: boxes, surface-area-heuristic cost, and the
//      SAH leaf-collapse decision (subtree of <= max_leaf primitives becomes one leaf when testing
//      all of them is cheaper than descending)
//   6. survivors = internal nodes not inside a collapsed subtree; exclusive scan -> compact ids
//   7. emit 64 B nodes (scene.cuh layout) with child boxes padded outwards by `pad`
// Culling in the traversal kernels is conservative (padded boxes, padded slab interval), so the
// closest hit a ray finds does not depend on the shape of this tree (oracle/pt_scene_oracle.h).
#include <cfloat>
#include <cstdlib>
#include <cub/cub.cuh>

#include "scene.cuh"

namespace ptb {
namespace {

#ifndef PTB_ROTATE_DEFAULT
#define PTB_ROTATE_DEFAULT 4
#endif
#ifndef PTB_MORTON_CUBIC_DEFAULT
#define PTB_MORTON_CUBIC_DEFAULT true
#endif
constexpr float SAH_C_PRIM = 1.0f;       // one primitive test (+ its 16/48 B fetch)

__device__ __forceinline__ uint32_t f2ord(float f)
{
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t u)
{
    return __uint_as_float((u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u);
}

__global__ void init_bounds_kernel(uint32_t *cb)
{
    if (threadIdx.x < 3) cb[threadIdx.x] = 0xFFFFFFFFu;        // min
    else if (threadIdx.x < 6) cb[threadIdx.x] = 0u;            // max
}

__global__ void __launch_bounds__(256) centroid_bounds_kernel(const float4 *__restrict__ lo, const float4 *__restrict__ hi,
                                                              uint32_t n, uint32_t *cb)
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const float4 a = lo[i], b = hi[i];
        const float c[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            mn[k] = fminf(mn[k], c[k]);
            mx[k] = fmaxf(mx[k], c[k]);
        }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o; o >>= 1) {
            mn[k] = fminf(mn[k], __shfl_xor_sync(0xFFFFFFFFu, mn[k], o));
            mx[k] = fmaxf(mx[k], __shfl_xor_sync(0xFFFFFFFFu, mx[k], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            atomicMin(&cb[k], f2ord(mn[k]));
            atomicMax(&cb[3 + k], f2ord(mx[k]));
        }
    }
}

__device__ __forceinline__ uint64_t spread21(uint32_t v)
{
    uint64_t x = v & 0x1FFFFFu;
    x = (x | x << 32) & 0x1F00000000FFFFull;
    x = (x | x << 16) & 0x1F0000FF0000FFull;
    x = (x | x << 8) & 0x100F00F00F00F00Full;
    x = (x | x << 4) & 0x10C30C30C30C30C3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}

// cubic != 0: one cell size on all three axes (every axis is scaled by the LARGEST centroid extent), so that the code's
// leading bits split space where it is actually wide: a terrain's height then only enters at the bit level where cells are
// as small as its relief, instead of taking every third split from the top (which cuts a height field into interleaved,
// overlapping halves).  cubic == 0: each axis scaled by its own extent.
__global__ void __launch_bounds__(256) morton_kernel(const float4 *__restrict__ lo, const float4 *__restrict__ hi, uint32_t n,
                                                     const uint32_t *__restrict__ cb, int cubic, uint64_t *keys, uint32_t *vals)
{
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const float4 a = lo[i], b = hi[i];
    const float c[3] = {0.5f * (a.x + b.x), 0.5f * (a.y + b.y), 0.5f * (a.z + b.z)};
    const float ext_max = fmaxf(fmaxf(ord2f(cb[3]) - ord2f(cb[0]), ord2f(cb[4]) - ord2f(cb[1])), ord2f(cb[5]) - ord2f(cb[2]));
    uint32_t q[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const float mn = ord2f(cb[k]), mx = ord2f(cb[3 + k]);
        const float ext = cubic ? ext_max : mx - mn;
        float t = ext > 0.0f ? (c[k] - mn) / ext : 0.0f;
        t = fminf(fmaxf(t * 2097152.0f, 0.0f), 2097151.0f);
        q[k] = (uint32_t)t;
    }
    keys[i] = (spread21(q[0]) << 2) | (spread21(q[1]) << 1) | spread21(q[2]);
    vals[i] = i;
}

// Karras: length of the common prefix of keys i and j (index bits break ties); -1 outside [0,n).
__device__ __forceinline__ int delta(const uint64_t *__restrict__ keys, int n, int i, uint64_t ki, int j)
{
    if (j < 0 || j >= n) return -1;
    const uint64_t kj = keys[j];
    if (ki == kj) return 64 + __clz((uint32_t)i ^ (uint32_t)j);
    return __clzll((long long)(ki ^ kj));
}

struct BuildTmp {
    int *left, *right;        // children of internal node i: >= 0 internal, < 0 ~leaf position
    int *parent_int;          // parent of internal node (root: -1)
    int *parent_leaf;         // parent of leaf position
    uint32_t *range_lo, *range_hi;
    float4 *nlo, *nhi;        // internal node boxes (unpadded)
    float *ncost;
    uint32_t *arrive;         // arrival counters
    uint8_t *marked;          // collapse decision
    uint32_t *flags, *compact;
};

__global__ void __launch_bounds__(256) karras_kernel(const uint64_t *__restrict__ keys, int n, BuildTmp t)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const uint64_t ki = keys[i];
    const int d = delta(keys, n, i, ki, i + 1) - delta(keys, n, i, ki, i - 1) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, ki, i - d);
    int lmax = 2;
    while (delta(keys, n, i, ki, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int s = lmax >> 1; s >= 1; s >>= 1)
        if (delta(keys, n, i, ki, i + (l + s) * d) > dmin) l += s;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, ki, j);
    int s = 0;
    int step = l;
    do {
        step = (step + 1) >> 1;
        if (delta(keys, n, i, ki, i + (s + step) * d) > dnode) s += step;
    } while (step > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int lc = (lo == gamma) ? ~gamma : gamma;
    const int rc = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    t.left[i] = lc;
    t.right[i] = rc;
    t.range_lo[i] = (uint32_t)lo;
    t.range_hi[i] = (uint32_t)hi;
    if (lc >= 0) t.parent_int[lc] = i; else t.parent_leaf[~lc] = i;
    if (rc >= 0) t.parent_int[rc] = i; else t.parent_leaf[~rc] = i;
    if (i == 0) t.parent_int[0] = -1;
}

__device__ __forceinline__ float box_area(float4 lo, float4 hi)
{
    const float dx = hi.x - lo.x, dy = hi.y - lo.y, dz = hi.z - lo.z;
    return 2.0f * (dx * dy + dy * dz + dz * dx);
}

__device__ __forceinline__ float4 ldcg4(const float4 *p) { return __ldcg(p); }

__device__ __forceinline__ float union_area(float4 alo, float4 ahi, float4 blo, float4 bhi)
{
    return box_area(make_float4(fminf(alo.x, blo.x), fminf(alo.y, blo.y), fminf(alo.z, blo.z), 0.0f),
                    make_float4(fmaxf(ahi.x, bhi.x), fmaxf(ahi.y, bhi.y), fmaxf(ahi.z, bhi.z), 0.0f));
}

// box and cost of child reference c (internal node or one primitive)
__device__ __forceinline__ void child_box(int c, const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi,
                                          const uint32_t *__restrict__ vals, const BuildTmp &t, float4 &lo, float4 &hi, float &cost)
{
    if (c >= 0) { lo = ldcg4(&t.nlo[c]); hi = ldcg4(&t.nhi[c]); cost = __ldcg(&t.ncost[c]); }
    else { const uint32_t v = vals[~c]; lo = prim_lo[v]; hi = prim_hi[v]; cost = SAH_C_PRIM * box_area(lo, hi); }
}

// Tree rotation (Kensler 2008) at node `cur`, tried when its subtree is complete: swapping one child with one grandchild on
// the other side leaves cur's box alone and changes only the box of the child that is rebuilt; the swap with the largest
// drop in surface area is taken.  Only children holding more than 2 * max_leaf primitives are rebuilt, so the subtrees that
// may collapse into leaves (which need their contiguous primitive ranges) are moved as wholes, never restructured.
__device__ __forceinline__ void try_rotation(int cur, const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi,
                                             const uint32_t *__restrict__ vals, uint32_t min_prims, float c_trav, BuildTmp &t)
{
    int side_best = -1, gc_best = 0;
    float gain_best = 0.0f;
    for (int side = 0; side < 2; ++side) {                 // side: the child that is rebuilt (0 = left, 1 = right)
        const int b = side == 0 ? t.left[cur] : t.right[cur], a = side == 0 ? t.right[cur] : t.left[cur];
        if (b < 0 || t.range_hi[b] - t.range_lo[b] + 1u <= min_prims || *(volatile uint8_t *)&t.marked[b]) continue;
        float4 alo, ahi, glo[2], ghi[2];
        float acost, gcost[2];
        child_box(a, prim_lo, prim_hi, vals, t, alo, ahi, acost);
        const int g[2] = {__ldcg(&t.left[b]), __ldcg(&t.right[b])};      // b's own rotation, by another thread, may have changed them
        child_box(g[0], prim_lo, prim_hi, vals, t, glo[0], ghi[0], gcost[0]);
        child_box(g[1], prim_lo, prim_hi, vals, t, glo[1], ghi[1], gcost[1]);
        const float area_b = box_area(ldcg4(&t.nlo[b]), ldcg4(&t.nhi[b]));
        for (int j = 0; j < 2; ++j) {                      // grandchild g[j] moves up, a takes its place next to g[1 - j]
            const float gain = area_b - union_area(alo, ahi, glo[1 - j], ghi[1 - j]);
            if (gain > gain_best) { gain_best = gain; side_best = side; gc_best = j; }
        }
    }
    // grandchild <-> grandchild: cur = ((LL, LR), (RL, RR)) -> ((RL, LR), (LL, RR)) or ((RR, LR), (RL, LL)); both children are rebuilt
    {
        const int L = t.left[cur], R = t.right[cur];
        if (L >= 0 && R >= 0 && t.range_hi[L] - t.range_lo[L] + 1u > min_prims && t.range_hi[R] - t.range_lo[R] + 1u > min_prims &&
            !*(volatile uint8_t *)&t.marked[L] && !*(volatile uint8_t *)&t.marked[R]) {
            const int g[4] = {__ldcg(&t.left[L]), __ldcg(&t.right[L]), __ldcg(&t.left[R]), __ldcg(&t.right[R])};
            float4 glo[4], ghi[4];
            float gcost[4];
            for (int j = 0; j < 4; ++j) child_box(g[j], prim_lo, prim_hi, vals, t, glo[j], ghi[j], gcost[j]);
            const float area_now = box_area(ldcg4(&t.nlo[L]), ldcg4(&t.nhi[L])) + box_area(ldcg4(&t.nlo[R]), ldcg4(&t.nhi[R]));
            // partner of LL (g[0]) among R's children: RL (g[2]) or RR (g[3]) moves to L, LL moves to R
            int pick = -1;
            float gain_gg = gain_best;
            for (int j = 2; j < 4; ++j) {
                const int other = j == 2 ? 3 : 2;
                const float area_new = union_area(glo[j], ghi[j], glo[1], ghi[1]) + union_area(glo[0], ghi[0], glo[other], ghi[other]);
                if (area_now - area_new > gain_gg) { gain_gg = area_now - area_new; pick = j; }
            }
            if (pick >= 0) {
                const int other = pick == 2 ? 3 : 2;
                // L' = (g[pick], LR), R' = (LL, g[other])
                t.left[L] = g[pick]; t.right[L] = g[1];
                t.left[R] = g[0];    t.right[R] = g[other];
                if (g[pick] >= 0) t.parent_int[g[pick]] = L; else t.parent_leaf[~g[pick]] = L;
                if (g[0] >= 0) t.parent_int[g[0]] = R; else t.parent_leaf[~g[0]] = R;
                const float4 llo = make_float4(fminf(glo[pick].x, glo[1].x), fminf(glo[pick].y, glo[1].y), fminf(glo[pick].z, glo[1].z), 0.0f);
                const float4 lhi = make_float4(fmaxf(ghi[pick].x, ghi[1].x), fmaxf(ghi[pick].y, ghi[1].y), fmaxf(ghi[pick].z, ghi[1].z), 0.0f);
                const float4 rlo = make_float4(fminf(glo[0].x, glo[other].x), fminf(glo[0].y, glo[other].y), fminf(glo[0].z, glo[other].z), 0.0f);
                const float4 rhi = make_float4(fmaxf(ghi[0].x, ghi[other].x), fmaxf(ghi[0].y, ghi[other].y), fmaxf(ghi[0].z, ghi[other].z), 0.0f);
                t.nlo[L] = llo; t.nhi[L] = lhi; t.ncost[L] = c_trav * box_area(llo, lhi) + gcost[pick] + gcost[1];
                t.nlo[R] = rlo; t.nhi[R] = rhi; t.ncost[R] = c_trav * box_area(rlo, rhi) + gcost[0] + gcost[other];
                __threadfence();
                return;
            }
        }
    }
    if (side_best < 0 || !(gain_best > 0.0f)) return;
    const int b = side_best == 0 ? t.left[cur] : t.right[cur], a = side_best == 0 ? t.right[cur] : t.left[cur];
    const int bl = __ldcg(&t.left[b]), br = __ldcg(&t.right[b]);
    const int up = gc_best == 0 ? bl : br, stay = gc_best == 0 ? br : bl;
    // cur = (up, b'), b' = (a, stay)
    t.left[b] = a;
    t.right[b] = stay;
    if (a >= 0) t.parent_int[a] = b; else t.parent_leaf[~a] = b;
    if (up >= 0) t.parent_int[up] = cur; else t.parent_leaf[~up] = cur;
    if (side_best == 0) t.right[cur] = up; else t.left[cur] = up;
    float4 alo, ahi, slo, shi;
    float acost, scost;
    child_box(a, prim_lo, prim_hi, vals, t, alo, ahi, acost);
    child_box(stay, prim_lo, prim_hi, vals, t, slo, shi, scost);
    const float4 lo = make_float4(fminf(alo.x, slo.x), fminf(alo.y, slo.y), fminf(alo.z, slo.z), 0.0f);
    const float4 hi = make_float4(fmaxf(ahi.x, shi.x), fmaxf(ahi.y, shi.y), fmaxf(ahi.z, shi.z), 0.0f);
    t.nlo[b] = lo;
    t.nhi[b] = hi;
    t.ncost[b] = c_trav * box_area(lo, hi) + acost + scost;
    __threadfence();
}

__global__ void __launch_bounds__(256) refit_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi,
                                                    const uint32_t *__restrict__ vals, int n, uint32_t max_leaf,
                                                    float SAH_C_TRAVERSE, int rotate, BuildTmp t)
{
    const int p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    int cur = t.parent_leaf[p];
    while (cur >= 0) {
        __threadfence();
        if (atomicAdd(&t.arrive[cur], 1u) == 0u) return;      // the sibling subtree finishes this node
        __threadfence();
        if (rotate) try_rotation(cur, prim_lo, prim_hi, vals, max_leaf * (uint32_t)rotate, SAH_C_TRAVERSE, t);
        const int lc = t.left[cur], rc = t.right[cur];
        float4 llo, lhi, rlo, rhi;
        float lcost, rcost;
        if (lc >= 0) { llo = ldcg4(&t.nlo[lc]); lhi = ldcg4(&t.nhi[lc]); lcost = __ldcg(&t.ncost[lc]); }
        else { const uint32_t v = vals[~lc]; llo = prim_lo[v]; lhi = prim_hi[v]; lcost = SAH_C_PRIM * box_area(llo, lhi); }
        if (rc >= 0) { rlo = ldcg4(&t.nlo[rc]); rhi = ldcg4(&t.nhi[rc]); rcost = __ldcg(&t.ncost[rc]); }
        else { const uint32_t v = vals[~rc]; rlo = prim_lo[v]; rhi = prim_hi[v]; rcost = SAH_C_PRIM * box_area(rlo, rhi); }
        const float4 lo = make_float4(fminf(llo.x, rlo.x), fminf(llo.y, rlo.y), fminf(llo.z, rlo.z), 0.0f);
        const float4 hi = make_float4(fmaxf(lhi.x, rhi.x), fmaxf(lhi.y, rhi.y), fmaxf(lhi.z, rhi.z), 0.0f);
        const float area = box_area(lo, hi);
        const uint32_t cnt = t.range_hi[cur] - t.range_lo[cur] + 1u;
        const float cost_int = SAH_C_TRAVERSE * area + lcost + rcost;
        const float cost_leaf = SAH_C_PRIM * area * (float)cnt;
        const bool collapse = cnt <= max_leaf && cost_leaf <= cost_int;
        t.nlo[cur] = lo;
        t.nhi[cur] = hi;
        t.ncost[cur] = collapse ? cost_leaf : cost_int;
        t.marked[cur] = collapse ? 1 : 0;
        cur = t.parent_int[cur];
    }
}

__global__ void __launch_bounds__(256) survive_kernel(int n_internal, uint32_t max_leaf, BuildTmp t)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_internal) return;
    bool dead = t.marked[i] != 0;
    int a = t.parent_int[i];
    while (!dead && a >= 0 && (t.range_hi[a] - t.range_lo[a] + 1u) <= max_leaf) {
        dead = t.marked[a] != 0;
        a = t.parent_int[a];
    }
    t.flags[i] = dead ? 0u : 1u;
}

__device__ __forceinline__ void child_of(int c, const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi,
                                         const uint32_t *__restrict__ vals, const BuildTmp &t, float pad, float4 &lo, float4 &hi,
                                         int &ref, uint32_t &leaves)
{
    if (c >= 0) {
        lo = t.nlo[c];
        hi = t.nhi[c];
        if (t.flags[c]) ref = (int)t.compact[c];
        else { ref = make_leaf_ref(t.range_lo[c], t.range_hi[c] - t.range_lo[c] + 1u); ++leaves; }
    } else {
        const uint32_t v = vals[~c];
        lo = prim_lo[v];
        hi = prim_hi[v];
        ref = make_leaf_ref((uint32_t)~c, 1u);
        ++leaves;
    }
    lo.x -= pad; lo.y -= pad; lo.z -= pad;
    hi.x += pad; hi.y += pad; hi.z += pad;
}

__global__ void __launch_bounds__(256) emit_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi,
                                                   const uint32_t *__restrict__ vals, int n_internal, float pad, BuildTmp t,
                                                   BvhNode *nodes, uint32_t *n_leaves)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_internal || !t.flags[i]) return;
    float4 l0, h0, l1, h1;
    int r0, r1;
    uint32_t leaves = 0;
    child_of(t.left[i], prim_lo, prim_hi, vals, t, pad, l0, h0, r0, leaves);
    child_of(t.right[i], prim_lo, prim_hi, vals, t, pad, l1, h1, r1, leaves);
    BvhNode nd;
    nd.a = make_float4(l0.x, h0.x, l0.y, h0.y);
    nd.b = make_float4(l1.x, h1.x, l1.y, h1.y);
    nd.c = make_float4(l0.z, h0.z, l1.z, h1.z);
    nd.d = make_int4(r0, r1, 0, 0);
    nodes[t.compact[i]] = nd;
    if (leaves) atomicAdd(n_leaves, leaves);
}

// Everything collapsed into one leaf (or a single primitive): node 0 holds that leaf as child 0
// and an empty child 1.
__global__ void single_leaf_kernel(const float4 *__restrict__ prim_lo, const float4 *__restrict__ prim_hi, uint32_t n, float pad,
                                   BvhNode *nodes)
{
    float4 lo = prim_lo[0], hi = prim_hi[0];
    for (uint32_t i = 1; i < n; ++i) {
        const float4 a = prim_lo[i], b = prim_hi[i];
        lo = make_float4(fminf(lo.x, a.x), fminf(lo.y, a.y), fminf(lo.z, a.z), 0.0f);
        hi = make_float4(fmaxf(hi.x, b.x), fmaxf(hi.y, b.y), fmaxf(hi.z, b.z), 0.0f);
    }
    BvhNode nd;
    nd.a = make_float4(lo.x - pad, hi.x + pad, lo.y - pad, hi.y + pad);
    nd.b = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    nd.c = make_float4(lo.z - pad, hi.z + pad, 0.0f, 0.0f);
    nd.d = make_int4(make_leaf_ref(0u, n), EMPTY_REF, 0, 0);
    nodes[0] = nd;
}

// Cost of one node visit (two box tests + one 64 B fetch) in units of one primitive test.  The primitive tests are
// exactly rounded (IEEE division and square root), a node visit is 12 FMAs and a handful of min/max: measured
// (profiles/r01_notes.md) 0.5 / 1 / 1.5 / 2 / 3 -> c3 3,831 / 3,739 / 3,689 / 3,606 / 3,342, c4 1,110 / 1,107 / 1,070 /
// 1,007 / 864 Msamples/s.  PTB200_SAH_CTRAV overrides.
float sah_c_traverse()
{
    static const float v = [] {
        const char *e = std::getenv("PTB200_SAH_CTRAV");
        return e ? (float)std::atof(e) : 0.5f;
    }();
    return v;
}

// PTB200_ROTATE=0|1: tree rotations in the refit pass (A/B; profiles/r02_notes.md)
bool morton_cubic()         // PTB200_MORTON=axis|cube (A/B)
{
    const char *e = std::getenv("PTB200_MORTON");
    return e ? e[0] == 'c' : PTB_MORTON_CUBIC_DEFAULT;
}

int tree_rotations()        // number of refit passes that try rotations (0: none)
{
    const char *e = std::getenv("PTB200_ROTATE");
    const int v = e ? std::atoi(e) : PTB_ROTATE_DEFAULT;
    return v < 0 ? 0 : v > 16 ? 16 : v;
}
uint32_t rotate_min_factor()   // a child is rebuilt when it holds more than this many times max_leaf primitives
{
    const char *e = std::getenv("PTB200_ROTATE_MIN");
    const int v = e ? std::atoi(e) : 1;
    return (uint32_t)(v < 1 ? 1 : v);
}

template <typename T>
cudaError_t dmalloc(T **p, size_t count)
{
    return cudaMalloc((void **)p, (count ? count : 1) * sizeof(T));
}

}  // namespace

void free_bvh(DeviceBvh *b)
{
    cudaFree(b->nodes);
    cudaFree(b->order);
    *b = DeviceBvh{};
}

#define BK(call)                          \
    do {                                  \
        err = (call);                     \
        if (err != cudaSuccess) goto out; \
    } while (0)

cudaError_t build_lbvh(const float4 *prim_lo, const float4 *prim_hi, uint32_t n, uint32_t max_leaf, float pad, DeviceBvh *out,
                       cudaStream_t s)
{
    *out = DeviceBvh{};
    out->root_ref = EMPTY_REF;
    if (n == 0) return cudaSuccess;
    if (max_leaf < 1) max_leaf = 1;
    if (max_leaf > MAX_LEAF) max_leaf = MAX_LEAF;

    cudaError_t err = cudaSuccess;
    const int ni = (int)n - 1;
    uint32_t *cb = nullptr, *vals_a = nullptr, *vals_b = nullptr, *n_leaves = nullptr;
    uint64_t *keys_a = nullptr, *keys_b = nullptr;
    void *cub_tmp = nullptr;
    size_t sort_bytes = 0, scan_bytes = 0, cub_bytes = 0;
    BuildTmp t{};
    uint32_t h_last_flag = 0, h_last_compact = 0, h_leaves = 0;
    const unsigned blocks_n = (n + 255u) / 256u;
    const unsigned blocks_i = ni > 0 ? ((unsigned)ni + 255u) / 256u : 1u;

    BK(dmalloc(&cb, 6));
    BK(dmalloc(&keys_a, n));
    BK(dmalloc(&keys_b, n));
    BK(dmalloc(&vals_a, n));
    BK(dmalloc(&vals_b, n));
    BK(dmalloc(&n_leaves, 1));
    BK(dmalloc(&out->order, n));
    BK(cudaMemsetAsync(n_leaves, 0, sizeof(uint32_t), s));

    init_bounds_kernel<<<1, 32, 0, s>>>(cb);
    centroid_bounds_kernel<<<min(blocks_n, 148u * 8u), 256, 0, s>>>(prim_lo, prim_hi, n, cb);
    morton_kernel<<<blocks_n, 256, 0, s>>>(prim_lo, prim_hi, n, cb, morton_cubic() ? 1 : 0, keys_a, vals_a);
    BK(cudaGetLastError());

    BK(cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, keys_a, keys_b, vals_a, vals_b, (int)n, 0, 63, s));
    if (ni > 0) BK(cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (uint32_t *)nullptr, (uint32_t *)nullptr, ni, s));
    cub_bytes = sort_bytes > scan_bytes ? sort_bytes : scan_bytes;
    BK(cudaMalloc(&cub_tmp, cub_bytes ? cub_bytes : 1));
    BK(cub::DeviceRadixSort::SortPairs(cub_tmp, sort_bytes, keys_a, keys_b, vals_a, vals_b, (int)n, 0, 63, s));
    BK(cudaMemcpyAsync(out->order, vals_b, sizeof(uint32_t) * n, cudaMemcpyDeviceToDevice, s));

    if (ni == 0) {
        BK(dmalloc(&out->nodes, 1));
        single_leaf_kernel<<<1, 1, 0, s>>>(prim_lo, prim_hi, 1u, pad, out->nodes);
        BK(cudaGetLastError());
        out->n_nodes = 1;
        out->n_prims = n;
        out->n_leaves = 1;
        out->root_ref = 0;
        goto out;
    }

    BK(dmalloc(&t.left, ni));
    BK(dmalloc(&t.right, ni));
    BK(dmalloc(&t.parent_int, ni));
    BK(dmalloc(&t.parent_leaf, n));
    BK(dmalloc(&t.range_lo, ni));
    BK(dmalloc(&t.range_hi, ni));
    BK(dmalloc(&t.nlo, ni));
    BK(dmalloc(&t.nhi, ni));
    BK(dmalloc(&t.ncost, ni));
    BK(dmalloc(&t.arrive, ni));
    BK(dmalloc(&t.marked, ni));
    BK(dmalloc(&t.flags, ni));
    BK(dmalloc(&t.compact, ni));
    BK(cudaMemsetAsync(t.arrive, 0, sizeof(uint32_t) * ni, s));

    karras_kernel<<<blocks_i, 256, 0, s>>>(keys_b, (int)n, t);
    refit_kernel<<<blocks_n, 256, 0, s>>>(prim_lo, prim_hi, vals_b, (int)n, max_leaf, sah_c_traverse(), tree_rotations() > 0 ? (int)rotate_min_factor() : 0, t);
    for (int pass = 1; pass < tree_rotations(); ++pass) {      // further bottom-up passes over the rotated tree
        BK(cudaMemsetAsync(t.arrive, 0, sizeof(uint32_t) * ni, s));
        refit_kernel<<<blocks_n, 256, 0, s>>>(prim_lo, prim_hi, vals_b, (int)n, max_leaf, sah_c_traverse(), (int)rotate_min_factor(), t);
    }
    survive_kernel<<<blocks_i, 256, 0, s>>>(ni, max_leaf, t);
    BK(cudaGetLastError());
    BK(cub::DeviceScan::ExclusiveSum(cub_tmp, scan_bytes, t.flags, t.compact, ni, s));
    BK(cudaMemcpyAsync(&h_last_flag, t.flags + (ni - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    BK(cudaMemcpyAsync(&h_last_compact, t.compact + (ni - 1), sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    BK(cudaStreamSynchronize(s));
    out->n_nodes = h_last_flag + h_last_compact;
    out->n_prims = n;
    if (out->n_nodes == 0) {
        // the root itself collapsed: the whole class is one leaf
        BK(dmalloc(&out->nodes, 1));
        // boxes must follow leaf order only for the union, which is order independent
        single_leaf_kernel<<<1, 1, 0, s>>>(prim_lo, prim_hi, n, pad, out->nodes);
        BK(cudaGetLastError());
        out->n_nodes = 1;
        out->n_leaves = 1;
        out->root_ref = 0;
        goto out;
    }
    BK(dmalloc(&out->nodes, out->n_nodes));
    emit_kernel<<<blocks_i, 256, 0, s>>>(prim_lo, prim_hi, vals_b, ni, pad, t, out->nodes, n_leaves);
    BK(cudaGetLastError());
    BK(cudaMemcpyAsync(&h_leaves, n_leaves, sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
    BK(cudaStreamSynchronize(s));
    out->n_leaves = h_leaves;
    out->root_ref = 0;

out:
    if (err == cudaSuccess) err = cudaStreamSynchronize(s);
    cudaFree(cb); cudaFree(keys_a); cudaFree(keys_b); cudaFree(vals_a); cudaFree(vals_b); cudaFree(n_leaves);
    cudaFree(cub_tmp);
    cudaFree(t.left); cudaFree(t.right); cudaFree(t.parent_int); cudaFree(t.parent_leaf);
    cudaFree(t.range_lo); cudaFree(t.range_hi); cudaFree(t.nlo); cudaFree(t.nhi); cudaFree(t.ncost);
    cudaFree(t.arrive); cudaFree(t.marked); cudaFree(t.flags); cudaFree(t.compact);
    if (err != cudaSuccess) free_bvh(out);
    return err;
}

}  // namespace ptb
