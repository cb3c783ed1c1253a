// kernels_csmain.cu — sm_100a kernels for the reference's per-pixel tracing path.
//
//   csmain_frame_*      : one launch == one Renderer.ComputeFrame dispatch
//                         (RayTracing/Graphics/Renderer.cs:1006-1040) of CSMain
//                         (RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:3-40), 1 primary ray per
//                         pixel, stored as packed R8G8B8A8 UNORM (Renderer.cs:124) and optionally as
//                         the shader-side float4 and a hit-id plane.
//   csmain_accumulate   : north_star stage n1/n2 — spp jittered samples per pixel summed
//                         in-kernel into a float4 accumulation buffer.
//
// Arithmetic is the frozen "one rounding per op" evaluation of Test.hlsl, so results are
// bit-identical to the CPU oracle's.  Both kernels are FP32-issue bound (the IEEE divisions and
// square roots dominate), so the work here is cutting instructions WITHOUT changing a single
// result bit: MODE_FAST replaces each IEEE division / sqrt by a short FMA sequence that is
// proven (exhaustively, selftest.cu) or provable (Markstein) to round identically on the domain
// the launch-uniform checks in api.cu guarantee; MODE_GENERIC is the literal intrinsic form and
// is used whenever those checks fail.  No dense contraction anywhere => no tensor cores.
#include <algorithm>
#include <cstdlib>
#include <numeric>
#include "ptb_internal.h"
#include "ptb_math.cuh"

namespace ptb {

// MODE_SPIRV: the operation list of the reference's checked-in compiled module Assets/Shaders/Compiled/Ray/Test.spirv
// (reciprocal multiply for the uv division, disc = Fma(b, b, a * -3), -b - s = Fma(d.z, -2, -s), folded constants;
// PTB_CFG_SPIRV_ARITH, literal scene only) with literal IEEE intrinsics — bit-identical to the numpy evaluation of the
// module's words (tests/test_spirv_pin.py, tests/test_gpu_spirv.py).
enum Mode { MODE_GENERIC = 0, MODE_FAST = 1, MODE_FAST_ON_AXIS = 2, MODE_SPIRV = 3 };
template <int MODE> struct ModeTraits { static constexpr bool fast = MODE == MODE_FAST || MODE == MODE_FAST_ON_AXIS; };

// Launch-uniform subexpressions, evaluated once on the device with the same intrinsics.
__global__ void prepare_terms_kernel(FrameK *k)
{
    k->inv_div_x = __frcp_rn(k->div_x);
    k->inv_div_y = __frcp_rn(k->div_y);
    const V3 oc{fsub(k->cam[0], k->centre[0]), fsub(k->cam[1], k->centre[1]), fsub(k->cam[2], k->centre[2])};  // :16
    k->oc[0] = oc.x;
    k->oc[1] = oc.y;
    k->oc[2] = oc.z;
    k->c = fsub(dot3(oc, oc), fmul(k->radius, k->radius));   // :19
    k->c4 = fmul(4.0f, k->c);
    k->neg2ocz = -fmul(2.0f, oc.z);
    k->nzero = -0.0f;
}

cudaError_t prepare_terms(FrameK *k_host, cudaStream_t s)
{
    FrameK *d = nullptr;
    cudaError_t e = cudaMalloc(&d, sizeof(FrameK));
    if (e != cudaSuccess) return e;
    e = cudaMemcpyAsync(d, k_host, sizeof(FrameK), cudaMemcpyHostToDevice, s);
    if (e == cudaSuccess) {
        prepare_terms_kernel<<<1, 1, 0, s>>>(d);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(k_host, d, sizeof(FrameK), cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d);
    return e;
}

// Test.hlsl:7 for one axis: uv = (id / resolution) * 2 - 1.
template <int MODE>
__device__ __forceinline__ float screen_uv(float p, float div, float inv_div)
{
    if (MODE == MODE_GENERIC) return fsub(fmul(fdiv(p, div), 2.0f), 1.0f);
    if (MODE == MODE_SPIRV) return fsub(fmul(fmul(p, inv_div), 2.0f), 1.0f);   // Test.spirv [34]-[36]: id * fl(1/1080)
    // q*2 is exact, so one FMA rounds exactly like (q*2) - 1.
    return ffma(div_rn_uniform(p, div, inv_div), 2.0f, -1.0f);
}

// Test.hlsl:10-37 given uv.  Returns true on hit (disc > 0, strict); rgb = colour.xyz (alpha is 1).
template <int MODE>
__device__ __forceinline__ bool csmain_shade(const FrameK &k, float uvx, float uvy, float &r, float &g, float &bl)
{
    V3 d;
    float a, b, disc;
    if (MODE == MODE_SPIRV) {
        // Test.spirv, result ids in brackets; cam (0,0,1), centre 0, r 0.5 are folded into the module
        d = normalize3(V3{uvx, uvy, -1.0f});                                   // [40]
        a = dot3(d, d);                                                        // [41]
        b = fmul(2.0f, d.z);                                                   // [43]
        disc = ffma(b, b, fmul(a, -3.0f));                                     // [44] [45]
        if (disc > 0.0f) {                                                     // [46]
            const float s = fsqrt(disc);                                       // [50]
            const float t = fdiv(ffma(d.z, -2.0f, -s), fmul(2.0f, a));         // [51]-[54]
            const V3 n = normalize3(V3{fadd(0.0f, fmul(d.x, t)), fadd(0.0f, fmul(d.y, t)), fadd(1.0f, fmul(d.z, t))});   // [55]-[57]
            r = fadd(fmul(n.x, 0.5f), 0.5f);                                   // [58] [59]
            g = fadd(fmul(n.y, 0.5f), 0.5f);
            bl = fadd(fmul(n.z, 0.5f), 0.5f);
            return true;
        }
        r = uvx;                                                               // [64]
        g = uvy;
        bl = 0.0f;
        return false;
    } else if (MODE == MODE_GENERIC) {
        d = normalize3(V3{uvx, uvy, -1.0f});                                   // :10
        a = dot3(d, d);                                                        // :17
        b = fmul(2.0f, dot3(V3{k.oc[0], k.oc[1], k.oc[2]}, d));                // :18
        disc = fsub(fmul(b, b), fmul(fmul(4.0f, a), k.c));                     // :21
    } else {
        // :10 — dot(v,v) with v.z = -1: (-1)*(-1) == 1 exactly; len in [1, 2^12) by the launch check.
        const float len = sqrt_rn_fast(fadd(fadd(fmul(uvx, uvx), fmul(uvy, uvy)), 1.0f));
        const float yh = rcp_rn_fast(len);                                     // RN(1/len)
        d = V3{div_rn_by_rcp(uvx, len, yh), div_rn_by_rcp(uvy, len, yh), -yh};  // -1/len == -RN(1/len)
        a = fadd(fadd(fmul(d.x, d.x), fmul(d.y, d.y)), fmul(yh, yh));          // :17
        if (MODE == MODE_FAST_ON_AXIS)
            b = fmul(k.neg2ocz, yh);   // :18 with oc = (0,0,z): 2*(z*d.z) == RN((-2z)*yh), scalings exact
        else
            b = fmul(2.0f, dot3(V3{k.oc[0], k.oc[1], k.oc[2]}, d));
        disc = fsub(fmul(b, b), fmul(a, k.c4));                                // :21, (4a)c == a(4c) exactly scaled
    }
    if (disc > 0.0f) {                                                         // :24
        const float t = fdiv(fsub(-b, fsqrt(disc)), fmul(2.0f, a));            // :27
        const V3 pc{fsub(fadd(k.cam[0], fmul(t, d.x)), k.centre[0]),           // :28-29
                    fsub(fadd(k.cam[1], fmul(t, d.y)), k.centre[1]),
                    fsub(fadd(k.cam[2], fmul(t, d.z)), k.centre[2])};
        V3 n;
        const float dot2 = dot3(pc, pc);
        if (MODE != MODE_GENERIC && in_fast_range(dot2)) {
            const float len2 = sqrt_rn_fast(dot2);
            const float yh2 = rcp_rn_fast(len2);
            n = V3{div_rn_by_rcp(pc.x, len2, yh2), div_rn_by_rcp(pc.y, len2, yh2), div_rn_by_rcp(pc.z, len2, yh2)};
        } else {
            const float len2 = fsqrt(dot2);
            n = V3{fdiv(pc.x, len2), fdiv(pc.y, len2), fdiv(pc.z, len2)};
        }
        r = ffma(n.x, 0.5f, 0.5f);   // :31 — n*0.5 is exact, so the FMA rounds like (n*0.5)+0.5
        g = ffma(n.y, 0.5f, 0.5f);
        bl = ffma(n.z, 0.5f, 0.5f);
        return true;
    }
    r = uvx;                                                                   // :36
    g = uvy;
    bl = 0.0f;
    return false;
}

// Conservative classification of a whole pixel: true only if EVERY sample position in [x,x+1] x [y,y+1] certainly
// evaluates to disc <= 0 (a miss) in the exactly-rounded csmain_shade.  With exact arithmetic
//   disc = -(4/|raw|^2) * G(u,v),   G = c*(u^2+v^2+1) - (oc.x*u + oc.y*v - oc.z)^2,   raw = (u, v, -1);
// the rounded evaluation differs from it by at most ~5e-6*K (K = |c| + |oc|^2: d, a, b each carry a few ulp), i.e. the
// computed disc is certainly <= 0 wherever G >= 1.3e-6*K*|raw|^2.  G over the pixel is bounded from its value and
// gradient at the pixel centre plus the second-order term (|Hessian| <= 2K): the test below asks for 1e-4*K*|raw|max^2,
// ~80x the requirement, which also swallows the rounding of this (ordinary float) evaluation and of the per-sample
// uv.  Such pixels (85 % of a 16:9 frame of the default scene) skip normalisation, discriminant and hit branch for all
// of their samples: their colour is (uv.x, uv.y, 0) (Test.hlsl:36).  NaN anywhere -> false -> the full path.
// Checked by exhaustion per context (ptb_selftest_exact_math which = 3) in the GPU tests.
__device__ __forceinline__ bool certain_miss(const FrameK &k, float fx, float fy)
{
    if (!k.cm_on) return false;
    const float u = (fx + 0.5f) * k.cm_sx - 1.0f, v = (fy + 0.5f) * k.cm_sy - 1.0f;   // pixel centre in uv
    const float hx = fabsf(k.cm_sx) * 0.5005f, hy = fabsf(k.cm_sy) * 0.5005f;          // half extents (inflated)
    const float s = k.oc[0] * u + k.oc[1] * v - k.oc[2];
    const float G = k.c * (u * u + v * v + 1.0f) - s * s;
    const float gu = fabsf(2.0f * (k.c * u - s * k.oc[0])), gv = fabsf(2.0f * (k.c * v - s * k.oc[1]));
    const float au = fabsf(u) + hx, av = fabsf(v) + hy;
    const float r2max = au * au + av * av + 1.0f;
    const float lower = G - gu * hx - gv * hy - k.cm_K * (hx * hx + hy * hy);
    return lower >= 1e-4f * k.cm_K * r2max;
}

// Two samples per instruction (FFMA2/FMUL2/FADD2), fast modes only.  Lane i of every F2 is exactly
// the scalar csmain_shade<MODE> evaluation of sample i: same ops, same order, one rounding each.
// Negated lengths are carried (xn, lenN, denN) so the Markstein residuals fma(q, -y, v) need no
// separate negation.  h0/h1 = per-lane hit.
template <int MODE>
__device__ __forceinline__ void csmain_shade2(const FrameK &k, F2 uvx, F2 uvy, F2 &r, F2 &g, F2 &bl, bool &h0, bool &h1)
{
    const F2 z = splat(k.nzero);   // run-time -0.0: keeps ptxas from fusing mul+add (see fmul2a)
    // :10 — -(dot(v,v)) = RN(-(uvx^2 + uvy^2) - 1)
    const F2 xn = ffma2(fadd2(fmul2a(uvx, uvx, z), fmul2a(uvy, uvy, z)), splat(-1.0f), splat(-1.0f));
    const F2 lenN = neg_sqrt_rn_fast2(xn);
    const F2 yh = rcp_rn_fast2_neg(lenN);                        // RN(1/len); d.z = -yh
    const F2 dx = div_rn_by_rcp2(uvx, lenN, yh), dy = div_rn_by_rcp2(uvy, lenN, yh);
    const F2 a = fadd2(fadd2(fmul2a(dx, dx, z), fmul2a(dy, dy, z)), fmul2a(yh, yh, z));   // :17
    F2 b;                                                        // :18
    if (MODE == MODE_FAST_ON_AXIS)
        b = fmul2(splat(k.neg2ocz), yh);
    else
        b = fmul2(splat(2.0f), fadd2(fadd2(fmul2a(splat(k.oc[0]), dx, z), fmul2a(splat(k.oc[1]), dy, z)),
                                     fmul2a(splat(-k.oc[2]), yh, z)));
    const F2 disc = fadd2(fmul2a(b, b, z), fmul2a(a, splat(-k.c4), z));  // :21
    h0 = disc.x > 0.0f;                                          // :24
    h1 = disc.y > 0.0f;
    r = uvx;                                                     // :36
    g = uvy;
    bl = splat(0.0f);
    if (h0 || h1) {
        // :27 — sqrt(disc); a missing lane gets a dummy argument.  disc may be arbitrarily small next to the silhouette:
        // the packed exactly-rounded sequence inside its proven domain, IEEE sqrt per lane outside it
        const float d0 = h0 ? disc.x : 1.0f, d1 = h1 ? disc.y : 1.0f;
        F2 num;
        if (in_fast_range(d0) && in_fast_range(d1))
            num = fadd2(neg2(b), neg_sqrt_rn_fast2(make_float2(-d0, -d1)));     // (-b) + (-sqrt) == (-b) - sqrt, one rounding
        else
            num = make_float2(fsub(-b.x, fsqrt(d0)), fsub(-b.y, fsqrt(d1)));
        const F2 denN = fmul2(splat(-2.0f), a);                  // -(2a), 2a in (1,4)
        const F2 t = div_rn_by_rcp2(num, denN, rcp_rn_fast2_neg(denN));
        // :28-29 — pc = (cam + t*d) - centre, with t*d.z = RN(t*(-yh))
        const F2 pcx = fadd2(fadd2(splat(k.cam[0]), fmul2a(t, dx, z)), splat(-k.centre[0]));
        const F2 pcy = fadd2(fadd2(splat(k.cam[1]), fmul2a(t, dy, z)), splat(-k.centre[1]));
        const F2 pcz = fadd2(fadd2(splat(k.cam[2]), fmul2a(t, neg2(yh), z)), splat(-k.centre[2]));
        const F2 dot2 = fadd2(fadd2(fmul2a(pcx, pcx, z), fmul2a(pcy, pcy, z)), fmul2a(pcz, pcz, z));
        F2 nx, ny, nz;
        if (in_fast_range(dot2.x) && in_fast_range(dot2.y)) {
            const F2 len2N = neg_sqrt_rn_fast2(neg2(dot2));
            const F2 yh2 = rcp_rn_fast2_neg(len2N);
            nx = div_rn_by_rcp2(pcx, len2N, yh2);
            ny = div_rn_by_rcp2(pcy, len2N, yh2);
            nz = div_rn_by_rcp2(pcz, len2N, yh2);
        } else {   // outside the proven domain (or a dummy lane): literal IEEE per lane
            const float l0 = fsqrt(dot2.x), l1 = fsqrt(dot2.y);
            nx = make_float2(fdiv(pcx.x, l0), fdiv(pcx.y, l1));
            ny = make_float2(fdiv(pcy.x, l0), fdiv(pcy.y, l1));
            nz = make_float2(fdiv(pcz.x, l0), fdiv(pcz.y, l1));
        }
        const F2 rh = ffma2(nx, splat(0.5f), splat(0.5f));       // :31
        const F2 gh = ffma2(ny, splat(0.5f), splat(0.5f));
        const F2 bh = ffma2(nz, splat(0.5f), splat(0.5f));
        r = make_float2(h0 ? rh.x : r.x, h1 ? rh.y : r.y);
        g = make_float2(h0 ? gh.x : g.x, h1 ? gh.y : g.y);
        bl = make_float2(h0 ? bh.x : 0.0f, h1 ? bh.y : 0.0f);
    }
}

// Test.hlsl:7 for two samples of one axis.
__device__ __forceinline__ F2 screen_uv2(F2 p, float div, float inv_div)
{
    return ffma2(div_rn_uniform2(p, splat(-div), splat(inv_div)), splat(2.0f), splat(-1.0f));
}

// Conservative classification of the pixel SPAN [x0, x1] x {y} (integer pixel positions, no jitter: what a 1 spp frame
// evaluates): the same bound as certain_miss() with the span's centre and half extent.
__device__ __forceinline__ bool certain_miss_span(const FrameK &k, float fx0, float fx1, float fy)
{
    if (!k.cm_on) return false;
    const float u = (0.5f * (fx0 + fx1)) * k.cm_sx - 1.0f, v = fy * k.cm_sy - 1.0f;   // span centre in uv
    const float hx = fabsf(k.cm_sx) * (0.5f * (fx1 - fx0) + 0.0005f), hy = fabsf(k.cm_sy) * 0.0005f;
    const float s = k.oc[0] * u + k.oc[1] * v - k.oc[2];
    const float G = k.c * (u * u + v * v + 1.0f) - s * s;
    const float gu = fabsf(2.0f * (k.c * u - s * k.oc[0])), gv = fabsf(2.0f * (k.c * v - s * k.oc[1]));
    const float au = fabsf(u) + hx, av = fabsf(v) + hy;
    const float r2max = au * au + av * av + 1.0f;
    const float lower = G - gu * hx - gv * hy - k.cm_K * (hx * hx + hy * hy);
    return lower >= 1e-4f * k.cm_K * r2max;
}

// Four consecutive pixels of one row -> one STG.128 of packed RGBA8 (plus the optional float4 / hit-id planes).
// row_span[y] = [first, end) quad columns of row y that need the full path; every quad outside is a certain miss
// (row_span_kernel below evaluates certain_miss_span once per context).  nullptr: no classification.
template <int MODE>
__device__ __forceinline__ void frame_quad(const FrameK &k, uint32_t xq, uint32_t y, uint4 *__restrict__ rgba8,
                                           float4 *__restrict__ f32, int4 *__restrict__ hit, const uint2 *__restrict__ row_span)
{
    const uint32_t wq = k.width >> 2;
    const float fy = __uint2float_rn(y);
    const float uvy = screen_uv<MODE>(fy, k.div_y, k.inv_div_y);
    uint32_t px8[4];
    int32_t hid[4];
    const size_t q = (size_t)y * wq + xq;
    if (!ModeTraits<MODE>::fast) {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const float uvx = screen_uv<MODE>(__uint2float_rn(xq * 4u + i), k.div_x, k.inv_div_x);
            float r, g, b;
            const bool h = csmain_shade<MODE>(k, uvx, uvy, r, g, b);
            px8[i] = pack_rgba8(r, g, b, 1.0f);
            hid[i] = h ? 0 : -1;
            if (f32) f32[q * 4 + i] = make_float4(r, g, b, 1.0f);
        }
    } else {
        bool lean = false;                                             // Test.hlsl:36 for all four pixels
        if (row_span) {
            const uint2 sp = __ldg(row_span + y);
            lean = xq < sp.x || xq >= sp.y;
        }
        // a miss stores (uv.x, uv.y, 0, 1): G and alpha bytes are the same for the whole quad
        const uint32_t ga_miss = (unorm8(uvy) << 8) | 0xFF000000u;
#pragma unroll
        for (int i = 0; i < 4; i += 2) {   // two pixels per FP32x2 instruction
            const F2 px = make_float2(__uint2float_rn(xq * 4u + i), __uint2float_rn(xq * 4u + i + 1));
            F2 r, g, b;
            bool h0 = false, h1 = false;
            const F2 uvx2 = screen_uv2(px, k.div_x, k.inv_div_x);
            if (lean) {
                r = uvx2;
                g = splat(uvy);
                b = splat(0.0f);
                px8[i] = unorm8(r.x) | ga_miss;
                px8[i + 1] = unorm8(r.y) | ga_miss;
            } else {
                csmain_shade2<MODE>(k, uvx2, splat(uvy), r, g, b, h0, h1);
                px8[i] = pack_rgba8(r.x, g.x, b.x, 1.0f);
                px8[i + 1] = pack_rgba8(r.y, g.y, b.y, 1.0f);
            }
            hid[i] = h0 ? 0 : -1;
            hid[i + 1] = h1 ? 0 : -1;
            if (f32) {
                f32[q * 4 + i] = make_float4(r.x, g.x, b.x, 1.0f);
                f32[q * 4 + i + 1] = make_float4(r.y, g.y, b.y, 1.0f);
            }
        }
    }
    rgba8[q] = make_uint4(px8[0], px8[1], px8[2], px8[3]);
    if (hit) hit[q] = make_int4(hid[0], hid[1], hid[2], hid[3]);
}

// Test.hlsl:7 for two lanes with per-lane divisor constants dn = -D, dh = RN(1/D), dl = RN(1/D - dh).
// DIV2: RN(p/D) = fma(p, dh, RN(p*dl)) — division by a constant through its reciprocal split in two floats (Brisebarre,
// Muller & Raina 2004), one instruction less than the Markstein form; correct for every binary32 p only for divisors that
// pass the exhaustive check made when the context is created (selftest.cu which = 5: every float in [2^-24, 2^17)
// against __fdiv_rn; 1080, 2160 and every divisor tried pass), otherwise FrameK::div2 = 0 selects the 3-instruction form.
// Pins four integers in registers as they are: without it the compiler re-derives the forward differences of
// csmain_accumulate's RNG state as a dozen separate induction variables (8 adds per sample instead of 3.5).
__device__ __forceinline__ void keep_form(uint32_t &a, uint32_t &b, uint32_t &c, uint32_t &d)
{
    asm("" : "+r"(a), "+r"(b), "+r"(c), "+r"(d));
}

__device__ __forceinline__ uint32_t add_on_fma(uint32_t a, uint32_t b, uint32_t one)
{
    uint32_t d;
    asm("mad.lo.u32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(one), "r"(b));
    return d;
}

template <bool DIV2>
__device__ __forceinline__ F2 screen_uv_axes(F2 p, F2 dn, F2 dh, F2 dl)
{
    if (DIV2) return ffma2(ffma2(p, dh, fmul2(p, dl)), splat(2.0f), splat(-1.0f));
    return ffma2(div_rn_uniform2(p, dn, dh), splat(2.0f), splat(-1.0f));
}

// ---- 1 spp frame, 4 pixels per thread, one STG.128 of packed RGBA8 per thread -------------
// Requires width % 4 == 0 (16 B aligned rows).  Block (32,8): a warp covers 128 consecutive
// pixels of one row = 512 contiguous bytes of RGBA8.
template <int MODE>
__global__ void __launch_bounds__(256) csmain_frame_vec4(const FrameK k, uint4 *__restrict__ rgba8,
                                                         float4 *__restrict__ f32, int4 *__restrict__ hit,
                                                         const uint2 *__restrict__ row_span)
{
    const uint32_t xq = blockIdx.x * blockDim.x + threadIdx.x;   // quad index along x
    const uint32_t y = shard_row(k.row_offset, k.row_stride, k.row_block, blockIdx.y * blockDim.y + threadIdx.y);   // row shard (multi-GPU tiles)
    if (xq >= (k.width >> 2) || y >= k.height) return;           // the guard Test.hlsl:39 lacks
    frame_quad<MODE>(k, xq, y, rgba8, f32, hit, row_span);
}

// One block per image row: the span of quad columns that are NOT certain misses (empty span = (wq, 0) -> all lean).
__global__ void __launch_bounds__(256) row_span_kernel(const FrameK k, uint2 *__restrict__ row_span)
{
    __shared__ uint32_t lo, hi;
    const uint32_t y = blockIdx.x, wq = k.width >> 2;
    if (threadIdx.x == 0) {
        lo = wq;
        hi = 0;
    }
    __syncthreads();
    const float fy = __uint2float_rn(y);
    for (uint32_t xq = threadIdx.x; xq < wq; xq += blockDim.x) {
        const float fx0 = __uint2float_rn(xq * 4u);
        if (!certain_miss_span(k, fx0, fx0 + 3.0f, fy)) {
            atomicMin(&lo, xq);
            atomicMax(&hi, xq + 1u);
        }
    }
    __syncthreads();
    if (threadIdx.x == 0) row_span[y] = make_uint2(lo, hi);
}

cudaError_t launch_row_span(const FrameK &k, uint2 *row_span, cudaStream_t s)
{
    row_span_kernel<<<k.height, 256, 0, s>>>(k, row_span);
    return cudaGetLastError();
}

// ---- n_frames literal frames in ONE persistent launch ---------------------------------------------------------
// Renderer.ComputeFrame is one ~2 us kernel per frame (Renderer.cs:1006-1040), so per-frame launches are bound by launch
// latency, not by the machine.  Here one resident wave of blocks loops over 1024-pixel row chunks (256 threads x one quad)
// of every frame: frame f+1 starts on an SM the moment that SM runs out of chunks of frame f, and the ramp-up / drain of
// a launch is paid once per call instead of once per frame.  Every frame is computed and stored in full; all frames write
// the same image (as the reference's frames do), so the final contents are those of a single frame.
template <int MODE>
__global__ void __launch_bounds__(256) csmain_frames_persistent(const FrameK k, uint4 *__restrict__ rgba8, uint32_t chunks_per_row,
                                                                uint32_t total_chunks, uint32_t step_rows, uint32_t step_cols,
                                                                const uint2 *__restrict__ row_span)
{
    // chunk c = blockIdx.x + i * gridDim.x of the (frame, row, chunk-of-row) sequence; (y, cx) advance by the grid size
    // decomposed on the host into whole rows + columns, so the loop carries no division (all block-uniform)
    const uint32_t wq = k.width >> 2;
    uint32_t y = blockIdx.x / chunks_per_row, cx = blockIdx.x - y * chunks_per_row;
    y %= k.height;
    for (uint32_t c = blockIdx.x; c < total_chunks; c += gridDim.x) {
        const uint32_t xq = cx * 256u + threadIdx.x;
        if (xq < wq) frame_quad<MODE>(k, xq, y, rgba8, nullptr, nullptr, row_span);
        y += step_rows;
        cx += step_cols;
        if (cx >= chunks_per_row) {
            cx -= chunks_per_row;
            ++y;
        }
        if (y >= k.height) y -= k.height;      // the next frame: same image
    }
}

// ---- 1 spp frame, 1 pixel per thread (any width) ------------------------------------------
template <int MODE>
__global__ void __launch_bounds__(256) csmain_frame_scalar(const FrameK k, uint32_t *__restrict__ rgba8,
                                                           float4 *__restrict__ f32, int32_t *__restrict__ hit)
{
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = shard_row(k.row_offset, k.row_stride, k.row_block, blockIdx.y * blockDim.y + threadIdx.y);   // row shard (multi-GPU tiles)
    if (x >= k.width || y >= k.height) return;
    float r, g, b;
    const bool h = csmain_shade<MODE>(k, screen_uv<MODE>(__uint2float_rn(x), k.div_x, k.inv_div_x),
                                      screen_uv<MODE>(__uint2float_rn(y), k.div_y, k.inv_div_y), r, g, b);
    const size_t i = (size_t)y * k.width + x;
    rgba8[i] = pack_rgba8(r, g, b, 1.0f);
    if (f32) f32[i] = make_float4(r, g, b, 1.0f);
    if (hit) hit[i] = h ? 0 : -1;
}

// rows this context renders (ptb_internal.h shard_row)
static uint32_t shard_rows(const FrameK &k)
{
    return shard_row_count(k.height, k.row_offset, k.row_stride, k.row_block);
}

static int pick_mode(const FrameK &k)
{
    if (k.spirv) return MODE_SPIRV;
    return k.fast ? (k.on_axis ? MODE_FAST_ON_AXIS : MODE_FAST) : MODE_GENERIC;
}

template <int MODE>
static cudaError_t launch_frame_t(const FrameK &k, const Buffers &b, cudaStream_t s)
{
    if (shard_rows(k) == 0) return cudaSuccess;
    const dim3 block(32, 8);
    if ((k.width & 3u) == 0) {
        const dim3 grid(((k.width >> 2) + 31) / 32, (shard_rows(k) + 7) / 8);
        csmain_frame_vec4<MODE><<<grid, block, 0, s>>>(k, reinterpret_cast<uint4 *>(b.rgba8), b.f32,
                                                       reinterpret_cast<int4 *>(b.hit), ModeTraits<MODE>::fast ? b.row_span : nullptr);
    } else {
        const dim3 grid((k.width + 31) / 32, (shard_rows(k) + 7) / 8);
        csmain_frame_scalar<MODE><<<grid, block, 0, s>>>(k, b.rgba8, b.f32, b.hit);
    }
    return cudaGetLastError();
}

template <int MODE>
static cudaError_t launch_frames_t(const FrameK &k, const Buffers &b, uint32_t n_frames, cudaStream_t s)
{
    static int grid_cap = 0;     // one resident wave: SMs x resident blocks of this kernel
    if (grid_cap == 0) {
        int per_sm = 0, dev = 0, sms = 148;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, csmain_frames_persistent<MODE>, 256, 0) != cudaSuccess || per_sm < 1)
            per_sm = 1;
        grid_cap = per_sm * sms;
    }
    const uint32_t chunks_per_row = ((k.width >> 2) + 255u) / 256u;
    const uint64_t per_frame = (uint64_t)chunks_per_row * k.height, total = per_frame * n_frames;
    if (total > 0xFFFFFFFFull) return cudaErrorNotSupported;
    // one resident wave; the grid size is made coprime with the chunks of a frame so that, frame after frame, every block
    // walks through ALL chunk positions (a common factor pins a block to a residue class — e.g. only the left, sphere-
    // covered halves of the rows — and leaves half of the SMs idle: measured 27 % idle cycles with 888 blocks on 2160 chunks)
    uint64_t grid = std::min<uint64_t>(total, (uint64_t)grid_cap);
    if (total > grid)
        while (grid > 1 && std::gcd(grid, per_frame) != 1) --grid;
    const uint32_t step_rows = (uint32_t)((grid / chunks_per_row) % k.height), step_cols = (uint32_t)(grid % chunks_per_row);
    csmain_frames_persistent<MODE><<<(unsigned)grid, 256, 0, s>>>(k, reinterpret_cast<uint4 *>(b.rgba8), chunks_per_row, (uint32_t)total,
                                                                 step_rows, step_cols, ModeTraits<MODE>::fast ? b.row_span : nullptr);
    return cudaGetLastError();
}

cudaError_t launch_csmain_frames(const FrameK &k, const Buffers &b, uint32_t n_frames, cudaStream_t s)
{
    if (n_frames == 0) return cudaSuccess;
    if ((k.width & 3u) != 0 || k.row_stride != 1 || k.row_offset != 0 || b.f32 || b.hit) return cudaErrorNotSupported;
    switch (pick_mode(k)) {
    case MODE_FAST_ON_AXIS: return launch_frames_t<MODE_FAST_ON_AXIS>(k, b, n_frames, s);
    case MODE_FAST: return launch_frames_t<MODE_FAST>(k, b, n_frames, s);
    case MODE_SPIRV: return launch_frames_t<MODE_SPIRV>(k, b, n_frames, s);
    default: return launch_frames_t<MODE_GENERIC>(k, b, n_frames, s);
    }
}

cudaError_t launch_csmain_frame(const FrameK &k, const Buffers &b, cudaStream_t s)
{
    switch (pick_mode(k)) {
    case MODE_FAST_ON_AXIS: return launch_frame_t<MODE_FAST_ON_AXIS>(k, b, s);
    case MODE_FAST: return launch_frame_t<MODE_FAST>(k, b, s);
    case MODE_SPIRV: return launch_frame_t<MODE_SPIRV>(k, b, s);
    default: return launch_frame_t<MODE_GENERIC>(k, b, s);
    }
}

// ---- multi-sample accumulation -------------------------------------------------------------
// One thread per pixel, samples looped in-kernel: the float4 read-modify-write of the
// accumulation buffer (32 B/pixel/launch) is amortised over n_samples, leaving the kernel
// FP32-issue bound.  Block (32,4) keeps a warp on one row (coalesced 512 B accum access) and a
// block on a 32x4 tile (hit/miss is spatially coherent, so few warps diverge).
// RNG: pcg3d keyed (pixel, sample, seed); the pixel and seed lanes of its first LCG step are
// loop-invariant and hoisted.
template <int MODE, bool JITTER, bool DIV2>
__global__ void __launch_bounds__(256) csmain_accumulate(const FrameK k, float4 *__restrict__ accum,
                                                         uint32_t first_sample, uint32_t n_samples, uint32_t seed,
                                                         uint32_t *__restrict__ out8, float spp_total, int bgra)
{
    const uint32_t x = blockIdx.x * blockDim.x + threadIdx.x;
    const uint32_t y = shard_row(k.row_offset, k.row_stride, k.row_block, blockIdx.y * blockDim.y + threadIdx.y);   // row shard (multi-GPU tiles)
    if (x >= k.width || y >= k.height) return;
    const uint32_t pix = y * k.width + x;
    const float fx = __uint2float_rn(x), fy = __uint2float_rn(y);
    // (x - 1) + m with m in [1,2) is the same real sum as x + (m - 1): one exact-rounded add.
    const float fxm1 = fsub(fx, 1.0f), fym1 = fsub(fy, 1.0f);
    const uint32_t vx0 = pix * 1664525u + 1013904223u;
    const uint32_t vz0 = seed * 1664525u + 1013904223u;
    const float uvx0 = screen_uv<MODE>(fx, k.div_x, k.inv_div_x), uvy0 = screen_uv<MODE>(fy, k.div_y, k.inv_div_y);

    float sr = 0.0f, sg = 0.0f, sb = 0.0f, sa = 0.0f;
    uint32_t i = 0;
    constexpr bool FAST = ModeTraits<MODE>::fast;
    const bool lean = FAST && certain_miss(k, fx, fy);   // every sample of this pixel misses
    if (FAST && JITTER) {
        // ---- pcg3d's first round by forward differences -----------------------------------------------------------
        // The RNG key is (pixel, sample, seed): across the samples of ONE pixel only the middle word moves, linearly
        // (y0(s) = s*A + C), and the z word starts from the launch constant z0.  Hence, all mod 2^32,
        //   x1(s) = x0 + y0(s)*z0          is linear in s,      slope alpha = A*z0
        //   y1(s) = y0(s) + z0*x1(s)       is linear in s,      slope beta  = A + z0*alpha
        //   z1(s) = z0 + x1(s)*y1(s)       is quadratic in s:   z1(s+1) - z1(s) = D(s),  D(s+1) = D(s) + 2*alpha*beta
        // so the three multiply-adds of the round (plus the one of y0) become four integer ADDs per sample.  The integer
        // multiply-adds share the FMA-heavy pipe with the FP32x2 arithmetic — the pipe ncu shows saturated
        // (math_pipe_throttle) — while the ALU pipe the adds go to had room.  Same 32-bit values, bit for bit.
        const uint32_t A = 1664525u, C = 1013904223u;
        const uint32_t z0 = vz0, alpha = A * z0, beta = A + z0 * alpha, ab2 = 2u * alpha * beta;      // launch-uniform
        const uint32_t y0 = (first_sample + i) * A + C;
        uint32_t ax = vx0 + y0 * z0;
        uint32_t ay = y0 + z0 * ax;
        uint32_t az = z0 + ax * ay;
        uint32_t dz = ax * beta + ay * alpha + alpha * beta;                                          // D(s)
        uint32_t alpha2 = 2u * alpha, beta2 = 2u * beta, ab4 = 2u * ab2;
        keep_form(alpha2, beta2, ab4, dz);
        if (lean) {
            // certain-miss pixels (most of the frame): Test.hlsl:36 for every sample, nothing else to evaluate.  The two
            // FP32x2 lanes carry the x and the y AXIS of one sample, so the sums (sr, sg) are one packed add per sample —
            // still sequential in sample order, the same additions as the scalar form.
            const F2 base = make_float2(fxm1, fym1), dn = make_float2(-k.div_x, -k.div_y);
            const F2 dh = make_float2(k.inv_div_x, k.inv_div_y), dl = make_float2(k.dl_x, k.dl_y);
            F2 sum = make_float2(0.0f, 0.0f);
            const uint32_t one = k.one;   // 1, from the parameter block so that ptxas keeps the multiply-add form
            const uint32_t pairs = (n_samples - i) >> 1;
            i += 2u * pairs;
#pragma unroll 4
            for (uint32_t pr = pairs; pr != 0; --pr) {
                // The x/y adds go out as multiply-adds by a register holding 1: the ALU pipe (shifts, xors, the float
                // assembly, 16 lanes per scheduler) is the busier one in this loop, the FMA pipes have room.
                uint32_t bx = add_on_fma(ax, alpha, one), by = add_on_fma(ay, beta, one), bz = az + dz;   // sample s + 1
                uint32_t px = ax, py = ay, pz = az, qx = bx, qy = by, qz = bz;
                ax = add_on_fma(ax, alpha2, one);                             // sample s + 2
                ay = add_on_fma(ay, beta2, one);
                az = bz + dz + ab2;
                dz += ab4;
                keep_form(ax, ay, az, dz);
                px ^= px >> 16; py ^= py >> 16; pz ^= pz >> 16;
                qx ^= qx >> 16; qy ^= qy >> 16; qz ^= qz >> 16;
                px += py * pz; qx += qy * qz;
                py += pz * px; qy += qz * qx;
                const F2 pa = fadd2(base, make_float2(bits_to_unit_plus_one(px), bits_to_unit_plus_one(py)));
                const F2 pb = fadd2(base, make_float2(bits_to_unit_plus_one(qx), bits_to_unit_plus_one(qy)));
                sum = fadd2(sum, screen_uv_axes<DIV2>(pa, dn, dh, dl));
                sum = fadd2(sum, screen_uv_axes<DIV2>(pb, dn, dh, dl));
            }
            sr = sum.x;
            sg = sum.y;
        } else {
            // two samples per iteration in FP32x2 lanes; the sums stay sequential in sample order
            const uint32_t pairs = (n_samples - i) >> 1;
            i += 2u * pairs;
            for (uint32_t pr = pairs; pr != 0; --pr) {
                uint32_t bx = ax + alpha, by = ay + beta, bz = az + dz;
                uint32_t px = ax, py = ay, pz = az, qx = bx, qy = by, qz = bz;
                ax += alpha2;
                ay += beta2;
                az = bz + dz + ab2;
                dz += ab4;
                keep_form(ax, ay, az, dz);
                px ^= px >> 16; py ^= py >> 16; pz ^= pz >> 16;
                qx ^= qx >> 16; qy ^= qy >> 16; qz ^= qz >> 16;
                px += py * pz; qx += qy * qz;
                py += pz * px; qy += qz * qx;
                const F2 sx = fadd2(splat(fxm1), make_float2(bits_to_unit_plus_one(px), bits_to_unit_plus_one(qx)));
                const F2 sy = fadd2(splat(fym1), make_float2(bits_to_unit_plus_one(py), bits_to_unit_plus_one(qy)));
                const F2 uvx2 = screen_uv_axes<DIV2>(sx, splat(-k.div_x), splat(k.inv_div_x), splat(k.dl_x));
                const F2 uvy2 = screen_uv_axes<DIV2>(sy, splat(-k.div_y), splat(k.inv_div_y), splat(k.dl_y));
                F2 r, g, b;
                bool h0, h1;
                csmain_shade2<MODE>(k, uvx2, uvy2, r, g, b, h0, h1);
                sr = fadd(fadd(sr, r.x), r.y);
                sg = fadd(fadd(sg, g.x), g.y);
                sb = fadd(fadd(sb, b.x), b.y);   // a miss adds +0.0 to a sum that is never -0.0: identity
            }
        }
    }
    for (; i < n_samples; ++i) {
        float uvx = uvx0, uvy = uvy0;
        if (JITTER) {
            uint32_t vx = vx0, vy = (first_sample + i) * 1664525u + 1013904223u, vz = vz0;
            vx += vy * vz;
            vy += vz * vx;
            vz += vx * vy;
            vx ^= vx >> 16;
            vy ^= vy >> 16;
            vz ^= vz >> 16;
            vx += vy * vz;
            vy += vz * vx;
            uvx = screen_uv<MODE>(fadd(fxm1, bits_to_unit_plus_one(vx)), k.div_x, k.inv_div_x);
            uvy = screen_uv<MODE>(fadd(fym1, bits_to_unit_plus_one(vy)), k.div_y, k.inv_div_y);
        }
        if (lean) {
            sr = fadd(sr, uvx);
            sg = fadd(sg, uvy);
            continue;
        }
        float r, g, b;
        const bool h = csmain_shade<MODE>(k, uvx, uvy, r, g, b);
        sr = fadd(sr, r);
        sg = fadd(sg, g);
        if (!FAST) {
            sb = fadd(sb, b);
            sa = fadd(sa, 1.0f);
        } else if (h) {
            sb = fadd(sb, b);
        }
    }
    // alpha: n_samples sequential additions of 1.0 are exact up to 2^24 (checked by the launcher)
    if (FAST) sa = __uint2float_rn(n_samples);
    float4 acc = accum[pix];
    acc.x = fadd(acc.x, sr);
    acc.y = fadd(acc.y, sg);
    acc.z = fadd(acc.z, sb);
    acc.w = fadd(acc.w, sa);
    accum[pix] = acc;
    if (out8) {   // launch-uniform: fused resolve (ptb_render_resolve) — the operations of resolve_kernel<false, BGRA>
        const uint32_t r8 = unorm8(fdiv(acc.x, spp_total)), g8 = unorm8(fdiv(acc.y, spp_total));
        const uint32_t b8 = unorm8(fdiv(acc.z, spp_total)), a8 = unorm8(fdiv(acc.w, spp_total));
        out8[pix] = bgra ? (b8 | (g8 << 8) | (r8 << 16) | (a8 << 24)) : (r8 | (g8 << 8) | (b8 << 16) | (a8 << 24));
    }
}

// ptb_selftest_exact_math which = 3: brute-force check of certain_miss for this context's parameters.
template <int MODE>
__global__ void __launch_bounds__(256) certain_miss_check_kernel(const FrameK k, unsigned long long *out)
{
    unsigned long long bad = 0, n = 0;
    const uint64_t total = (uint64_t)k.width * k.height;
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t x = (uint32_t)(p % k.width), y = (uint32_t)(p / k.width);
        const float fx = __uint2float_rn(x), fy = __uint2float_rn(y);
        if (!certain_miss(k, fx, fy)) continue;
        for (int jy = 0; jy < 8; ++jy)
            for (int jx = 0; jx < 8; ++jx) {
                // jitter values the RNG can produce: multiples of 2^-23 in [0, 1 - 2^-23]
                const float ox = jx == 7 ? 0.99999988079071044921875f : (float)jx * 0.140625f;
                const float oy = jy == 7 ? 0.99999988079071044921875f : (float)jy * 0.140625f;
                float r, g, b;
                bad += csmain_shade<MODE>(k, screen_uv<MODE>(fadd(fx, ox), k.div_x, k.inv_div_x),
                                          screen_uv<MODE>(fadd(fy, oy), k.div_y, k.inv_div_y), r, g, b) ? 1u : 0u;
                ++n;
            }
    }
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_down_sync(0xFFFFFFFFu, bad, o);
        n += __shfl_down_sync(0xFFFFFFFFu, n, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&out[0], bad);
        atomicAdd(&out[1], n);
    }
}

// ptb_selftest_exact_math which = 4: every pixel of every quad the row-span table marks "certain miss" really misses
// in the exactly rounded evaluation (the frame kernels evaluate integer pixel positions only).
template <int MODE>
__global__ void __launch_bounds__(256) row_span_check_kernel(const FrameK k, const uint2 *__restrict__ row_span, unsigned long long *out)
{
    unsigned long long bad = 0, n = 0;
    const uint32_t wq = k.width >> 2;
    const uint64_t total = (uint64_t)wq * k.height;
    for (uint64_t p = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; p < total; p += (uint64_t)gridDim.x * blockDim.x) {
        const uint32_t xq = (uint32_t)(p % wq), y = (uint32_t)(p / wq);
        const uint2 sp = row_span[y];
        if (!(xq < sp.x || xq >= sp.y)) continue;
        for (uint32_t i = 0; i < 4; ++i) {
            float r, g, b;
            bad += csmain_shade<MODE>(k, screen_uv<MODE>(__uint2float_rn(xq * 4u + i), k.div_x, k.inv_div_x),
                                      screen_uv<MODE>(__uint2float_rn(y), k.div_y, k.inv_div_y), r, g, b) ? 1u : 0u;
            ++n;
        }
    }
    for (int o = 16; o > 0; o >>= 1) {
        bad += __shfl_down_sync(0xFFFFFFFFu, bad, o);
        n += __shfl_down_sync(0xFFFFFFFFu, n, o);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&out[0], bad);
        atomicAdd(&out[1], n);
    }
}

cudaError_t selftest_row_span(const FrameK &k, const uint2 *row_span, unsigned long long *mismatches, unsigned long long *tested,
                              cudaStream_t s)
{
    unsigned long long *d = nullptr, h[2] = {0, 0};
    *mismatches = *tested = 0;
    if (!row_span) return cudaSuccess;             // nothing classified
    cudaError_t e = cudaMalloc(&d, sizeof h);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(d, 0, sizeof h, s);
    if (e == cudaSuccess) {
        if (pick_mode(k) == MODE_FAST_ON_AXIS) row_span_check_kernel<MODE_FAST_ON_AXIS><<<148 * 8, 256, 0, s>>>(k, row_span, d);
        else row_span_check_kernel<MODE_FAST><<<148 * 8, 256, 0, s>>>(k, row_span, d);
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d);
    *mismatches = h[0];
    *tested = h[1];
    return e;
}

cudaError_t selftest_certain_miss(const FrameK &k, unsigned long long *mismatches, unsigned long long *tested, cudaStream_t s)
{
    unsigned long long *d = nullptr, h[2] = {0, 0};
    cudaError_t e = cudaMalloc(&d, sizeof h);
    if (e != cudaSuccess) return e;
    e = cudaMemsetAsync(d, 0, sizeof h, s);
    if (e == cudaSuccess) {
        switch (pick_mode(k)) {
        case MODE_FAST_ON_AXIS: certain_miss_check_kernel<MODE_FAST_ON_AXIS><<<148 * 8, 256, 0, s>>>(k, d); break;
        case MODE_FAST: certain_miss_check_kernel<MODE_FAST><<<148 * 8, 256, 0, s>>>(k, d); break;
        default: certain_miss_check_kernel<MODE_GENERIC><<<148 * 8, 256, 0, s>>>(k, d); break;
        }
        e = cudaGetLastError();
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h, d, sizeof h, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d);
    *mismatches = h[0];
    *tested = h[1];
    return e;
}

template <int MODE>
static cudaError_t launch_acc_t(const FrameK &k, float4 *accum, uint32_t first, uint32_t n, uint32_t seed,
                                bool jitter, cudaStream_t s, uint32_t *out8, float spp_total, bool bgra)
{
    if (shard_rows(k) == 0) return cudaSuccess;
    // block height: 32x4 pixels measured best at 1080p (rows 8 / 4 / 2 / 1 -> 0.1711 / 0.1693 / 0.1729 / 0.1796 ms per
    // 64-spp launch: smaller blocks even out the expensive silhouette tiles); PTB200_ACC_BLOCK_Y overrides for A/B runs
    static const unsigned by = [] {
        const char *e = std::getenv("PTB200_ACC_BLOCK_Y");
        const int v = e ? std::atoi(e) : 4;
        return (unsigned)(v >= 1 && v <= 8 ? v : 4);
    }();
    const dim3 block(32, by);
    const dim3 grid((k.width + 31) / 32, (shard_rows(k) + by - 1) / by);
    if (jitter && k.div2 && ModeTraits<MODE>::fast)
        csmain_accumulate<MODE, true, true><<<grid, block, 0, s>>>(k, accum, first, n, seed, out8, spp_total, bgra ? 1 : 0);
    else if (jitter) csmain_accumulate<MODE, true, false><<<grid, block, 0, s>>>(k, accum, first, n, seed, out8, spp_total, bgra ? 1 : 0);
    else csmain_accumulate<MODE, false, false><<<grid, block, 0, s>>>(k, accum, first, n, seed, out8, spp_total, bgra ? 1 : 0);
    return cudaGetLastError();
}

cudaError_t launch_csmain_accumulate(const FrameK &k, float4 *accum, uint32_t first_sample,
                                     uint32_t n_samples, uint32_t seed, bool jitter, cudaStream_t s,
                                     uint32_t *out8, float spp_total, bool bgra)
{
    int mode = pick_mode(k);
    if (n_samples > (1u << 24) && mode != MODE_SPIRV) mode = MODE_GENERIC;
    switch (mode) {
    case MODE_FAST_ON_AXIS: return launch_acc_t<MODE_FAST_ON_AXIS>(k, accum, first_sample, n_samples, seed, jitter, s, out8, spp_total, bgra);
    case MODE_FAST: return launch_acc_t<MODE_FAST>(k, accum, first_sample, n_samples, seed, jitter, s, out8, spp_total, bgra);
    case MODE_SPIRV: return launch_acc_t<MODE_SPIRV>(k, accum, first_sample, n_samples, seed, jitter, s, out8, spp_total, bgra);
    default: return launch_acc_t<MODE_GENERIC>(k, accum, first_sample, n_samples, seed, jitter, s, out8, spp_total, bgra);
    }
}

}  // namespace ptb
