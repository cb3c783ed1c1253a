/*
 * pt_oracle.c — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).  See pt_oracle.h
 * for the scope, the frozen rounding spec and the parity-pin status
 * ("parity unpinned" w.r.t. reference fixtures — it has none; pinned to SURVEY.md
 * Appendix B known answers).
 *
 * Every function cites the reference file:line it restates
 * (paths relative to /root/reference/RayTracing/).
 *
 * Build: gcc -O2 -ffp-contract=off -fopenmp -shared -fPIC (oracle/Makefile).
 */
#include "pt_oracle.h"

#include <math.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* Threads of the following parallel regions (torchrun exports OMP_NUM_THREADS=1 to its workers; the timed CPU
 * baseline asks for the host's cores explicitly instead of inheriting that). */
void pto_set_threads(int n)
{
#ifdef _OPENMP
    if (n > 0) omp_set_num_threads(n);
#else
    (void)n;
#endif
}

int pto_max_threads(void)
{
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

/* Literals of Assets/Shaders/Source/Ray/Test.hlsl:6,8,12,13; the divisor generalises
 * to (H,H) so that 1080p reproduces Test.hlsl:6 exactly. */
void pto_default_frame(pto_frame_params *p, uint32_t width, uint32_t height)
{
    p->width = width;
    p->height = height;
    p->uv_div_x = (float)height;
    p->uv_div_y = (float)height;
    p->cam[0] = 0.0f; p->cam[1] = 0.0f; p->cam[2] = 1.0f;
    p->centre[0] = 0.0f; p->centre[1] = 0.0f; p->centre[2] = 0.0f;
    p->radius = 0.5f;
}

/* dot(x,y) = ((x0*y0 + x1*y1) + x2*y2), each op rounded (frozen spec). */
static inline float dot3(const float a[3], const float b[3])
{
    float s = a[0] * b[0];
    s = s + a[1] * b[1];
    s = s + a[2] * b[2];
    return s;
}

/* normalize(v) = v / sqrtf(dot(v,v)): three correctly rounded divisions. */
static inline void normalize3(const float v[3], float out[3])
{
    float len = sqrtf(dot3(v, v));
    out[0] = v[0] / len;
    out[1] = v[1] / len;
    out[2] = v[2] / len;
}

/* Test.hlsl:6-37.  px,py stand for (float)id.x,(float)id.y (plus jitter for n1). */
int pto_csmain_sample(const pto_frame_params *p, float px, float py,
                      float colour[4], float *disc_out, float *t_out)
{
    /* :7  uv = ((float2)id.xy / resolution) * 2.0 - 1.0 */
    float uvx = (px / p->uv_div_x) * 2.0f - 1.0f;
    float uvy = (py / p->uv_div_y) * 2.0f - 1.0f;
    /* :10 rayDir = normalize(float3(uv, -1.0)) */
    float raw[3] = { uvx, uvy, -1.0f };
    float d[3];
    normalize3(raw, d);
    /* :16 oc = cameraPos - sphereCenter */
    float oc[3] = { p->cam[0] - p->centre[0], p->cam[1] - p->centre[1], p->cam[2] - p->centre[2] };
    /* :17-19 */
    float a = dot3(d, d);
    float b = 2.0f * dot3(oc, d);
    float c = dot3(oc, oc) - p->radius * p->radius;
    /* :21 discriminant = b*b - 4*a*c */
    float bb = b * b;
    float ac4 = (4.0f * a) * c;
    float disc = bb - ac4;
    if (disc_out) *disc_out = disc;
    if (disc > 0.0f) {                                   /* :24 strict */
        /* :27 t = (-b - sqrt(disc)) / (2*a) */
        float t = ((-b) - sqrtf(disc)) / (2.0f * a);
        if (t_out) *t_out = t;
        /* :28 P = cameraPos + t * rayDir ; :29 normal = normalize(P - centre) */
        float P[3] = { p->cam[0] + t * d[0], p->cam[1] + t * d[1], p->cam[2] + t * d[2] };
        float pc[3] = { P[0] - p->centre[0], P[1] - p->centre[1], P[2] - p->centre[2] };
        float n[3];
        normalize3(pc, n);
        /* :31 color = float4(normal * 0.5 + 0.5, 1.0) */
        colour[0] = n[0] * 0.5f + 0.5f;
        colour[1] = n[1] * 0.5f + 0.5f;
        colour[2] = n[2] * 0.5f + 0.5f;
        colour[3] = 1.0f;
        return 1;
    }
    if (t_out) *t_out = 0.0f;
    /* :36 color = float4(uv.xy, 0.0, 1.0) */
    colour[0] = uvx;
    colour[1] = uvy;
    colour[2] = 0.0f;
    colour[3] = 1.0f;
    return 0;
}

/* ---- the "as-compiled" variant: Assets/Shaders/Compiled/Ray/Test.spirv, instruction by instruction ----
 * The reference checks in the DXC output of Test.hlsl (the module Renderer.cs:356-357 actually loads).  Its
 * operation list differs from the source in four places (asserted word by word against the
 * file in tests/test_spirv_pin.py):
 *   [34]      id * fl(1/1080)          (constant 0x3A72B9D6) instead of id / 1080
 *   [43]      b = 2.0 * d.z            (oc . d folded for oc = (0,0,1); exact, same value as the source form)
 *   [44],[45] disc = Fma(b, b, a * -3) instead of b*b - (4a)*0.75   (ONE rounding less)
 *   [52]      -b - s = Fma(d.z, -2, -s)                             (2*d.z is exact: same value as the source form)
 *   [56]      P = (0,0,1) + t*d, centre subtraction folded away     (centre = 0: same value)
 * Implementation-defined precisions (Normalize, Sqrt, FDiv, Dot order) are fixed as in the frozen spec above.
 * Only defined for the literal scene (cam (0,0,1), centre 0, r 0.5); the divisor generalises to fl(1/div). */
int pto_csmain_sample_spirv(const pto_frame_params *p, float px, float py,
                            float colour[4], float *disc_out, float *t_out)
{
    const float kx = 1.0f / p->uv_div_x, ky = 1.0f / p->uv_div_y;    /* [26] fl(1/1080) = 0x3A72B9D6 */
    float uvx = (px * kx) * 2.0f - 1.0f;                             /* [34] [35] [36] */
    float uvy = (py * ky) * 2.0f - 1.0f;
    float raw[3] = { uvx, uvy, -1.0f };
    float d[3];
    normalize3(raw, d);                                              /* [40] */
    float a = dot3(d, d);                                            /* [41] */
    float b = 2.0f * d[2];                                           /* [43] */
    float m3a = a * -3.0f;                                           /* [44] */
    float disc = fmaf(b, b, m3a);                                    /* [45] */
    if (disc_out) *disc_out = disc;
    if (disc > 0.0f) {                                               /* [46] */
        float s = sqrtf(disc);                                       /* [50] */
        float num = fmaf(d[2], -2.0f, -s);                           /* [51] [52] */
        float t = num / (2.0f * a);                                  /* [53] [54] */
        if (t_out) *t_out = t;
        float P[3] = { 0.0f + d[0] * t, 0.0f + d[1] * t, 1.0f + d[2] * t };   /* [55] [56] */
        float n[3];
        normalize3(P, n);                                            /* [57] */
        colour[0] = n[0] * 0.5f + 0.5f;                              /* [58] [59] */
        colour[1] = n[1] * 0.5f + 0.5f;
        colour[2] = n[2] * 0.5f + 0.5f;
        colour[3] = 1.0f;                                            /* [63] */
        return 1;
    }
    if (t_out) *t_out = 0.0f;
    colour[0] = uvx; colour[1] = uvy; colour[2] = 0.0f; colour[3] = 1.0f;   /* [64] */
    return 0;
}

/* Whole frame of the as-compiled variant (same outputs as pto_csmain_frame). */
int64_t pto_csmain_frame_spirv(const pto_frame_params *p, float *out_f32, uint8_t *out_rgba8,
                               int32_t *out_hit, int nthreads)
{
    const int64_t W = p->width, H = p->height;
    int64_t hits = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads) reduction(+ : hits)
#endif
    for (int64_t y = 0; y < H; ++y) {
        for (int64_t x = 0; x < W; ++x) {
            float col[4];
            int hit = pto_csmain_sample_spirv(p, (float)x, (float)y, col, NULL, NULL);
            int64_t i = y * W + x;
            hits += hit;
            if (out_f32) memcpy(out_f32 + 4 * i, col, sizeof col);
            if (out_rgba8) {
                out_rgba8[4 * i + 0] = pto_unorm8(col[0]);
                out_rgba8[4 * i + 1] = pto_unorm8(col[1]);
                out_rgba8[4 * i + 2] = pto_unorm8(col[2]);
                out_rgba8[4 * i + 3] = pto_unorm8(col[3]);
            }
            if (out_hit) out_hit[i] = hit ? 0 : -1;
        }
    }
    return hits;
}

/* Effective image store of Test.hlsl:39 into an R8G8B8A8Unorm image
 * (Graphics/Renderer.cs:124,167): NaN->0, clamp [0,1], scale 255, round-nearest-even. */
uint8_t pto_unorm8(float v)
{
    if (!(v > 0.0f)) return 0;          /* also catches NaN */
    if (v >= 1.0f) return 255;
    return (uint8_t)rintf(v * 255.0f);
}

/* sRGB OETF evaluated in double, then UNORM8 (Graphics/SwapChain.cs:154-165 selects
 * Format.B8G8R8A8Srgb, so the attachment write applies the OETF). */
uint8_t pto_srgb8(float v)
{
    if (!(v > 0.0f)) return 0;
    if (v >= 1.0f) return 255;
    double l = (double)v;
    double e = (l <= 0.0031308) ? 12.92 * l : 1.055 * pow(l, 1.0 / 2.4) - 0.055;
    return (uint8_t)rint(e * 255.0);
}

int64_t pto_csmain_frame(const pto_frame_params *p, float *out_f32, uint8_t *out_rgba8,
                         int32_t *out_hit, int nthreads)
{
    const int64_t W = p->width, H = p->height;
    int64_t hits = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads) reduction(+ : hits)
#endif
    for (int64_t y = 0; y < H; ++y) {          /* the Parallel.For analogue: rows */
        for (int64_t x = 0; x < W; ++x) {
            float col[4];
            int hit = pto_csmain_sample(p, (float)x, (float)y, col, NULL, NULL);
            int64_t i = y * W + x;
            hits += hit;
            if (out_f32) memcpy(out_f32 + 4 * i, col, sizeof col);
            if (out_rgba8) {
                out_rgba8[4 * i + 0] = pto_unorm8(col[0]);
                out_rgba8[4 * i + 1] = pto_unorm8(col[1]);
                out_rgba8[4 * i + 2] = pto_unorm8(col[2]);
                out_rgba8[4 * i + 3] = pto_unorm8(col[3]);
            }
            if (out_hit) out_hit[i] = hit ? 0 : -1;
        }
    }
    return hits;
}

/* ---- RNG (no reference counterpart: Test.hlsl is deterministic; spec frozen here) ---- */
void pto_pcg3d(uint32_t v[3])
{
    uint32_t x = v[0] * 1664525u + 1013904223u;
    uint32_t y = v[1] * 1664525u + 1013904223u;
    uint32_t z = v[2] * 1664525u + 1013904223u;
    x += y * z; y += z * x; z += x * y;
    x ^= x >> 16; y ^= y >> 16; z ^= z >> 16;
    x += y * z; y += z * x; z += x * y;
    v[0] = x; v[1] = y; v[2] = z;
}

float pto_bits_to_unit(uint32_t w)
{
    return (float)(w >> 9) * 1.1920928955078125e-7f;   /* 2^-23, exact */
}

void pto_accumulate(const pto_frame_params *p, uint32_t first_sample, uint32_t n_samples,
                    uint32_t seed, int jitter, float *accum, int nthreads)
{
    pto_accumulate_v(p, 0, first_sample, n_samples, seed, jitter, accum, nthreads);
}

/* variant 0 = the frozen source-order spec (pto_csmain_sample), 1 = the as-compiled one (pto_csmain_sample_spirv). */
void pto_accumulate_v(const pto_frame_params *p, int variant, uint32_t first_sample, uint32_t n_samples,
                      uint32_t seed, int jitter, float *accum, int nthreads)
{
    const int64_t W = p->width, H = p->height;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 4) num_threads(nthreads)
#endif
    for (int64_t y = 0; y < H; ++y) {
        for (int64_t x = 0; x < W; ++x) {
            const uint32_t pix = (uint32_t)(y * W + x);
            float sum[4] = { 0.0f, 0.0f, 0.0f, 0.0f };
            for (uint32_t k = 0; k < n_samples; ++k) {
                float jx = 0.0f, jy = 0.0f;
                if (jitter) {
                    uint32_t v[3] = { pix, first_sample + k, seed };
                    pto_pcg3d(v);
                    jx = pto_bits_to_unit(v[0]);
                    jy = pto_bits_to_unit(v[1]);
                }
                float col[4];
                if (variant) pto_csmain_sample_spirv(p, (float)x + jx, (float)y + jy, col, NULL, NULL);
                else pto_csmain_sample(p, (float)x + jx, (float)y + jy, col, NULL, NULL);
                sum[0] += col[0]; sum[1] += col[1]; sum[2] += col[2]; sum[3] += col[3];
            }
            float *a = accum + 4 * (int64_t)pix;
            a[0] += sum[0]; a[1] += sum[1]; a[2] += sum[2]; a[3] += sum[3];
        }
    }
}

void pto_resolve(const float *accum, uint32_t width, uint32_t height, uint32_t spp_total,
                 uint32_t flags, uint8_t *out8)
{
    const int64_t n = (int64_t)width * height;
    const float spp = (float)spp_total;
    const int bgra = (flags & PTO_RESOLVE_BGRA) != 0;
    const int srgb = (flags & PTO_RESOLVE_SRGB) != 0;
    /* the same Parallel.For analogue as the tracing loops: every pixel is independent */
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (int64_t i = 0; i < n; ++i) {
        float r = accum[4 * i + 0] / spp, g = accum[4 * i + 1] / spp;
        float b = accum[4 * i + 2] / spp, a = accum[4 * i + 3] / spp;
        uint8_t r8 = srgb ? pto_srgb8(r) : pto_unorm8(r);
        uint8_t g8 = srgb ? pto_srgb8(g) : pto_unorm8(g);
        uint8_t b8 = srgb ? pto_srgb8(b) : pto_unorm8(b);
        uint8_t a8 = pto_unorm8(a);                       /* alpha is always linear */
        out8[4 * i + 0] = bgra ? b8 : r8;
        out8[4 * i + 1] = g8;
        out8[4 * i + 2] = bgra ? r8 : b8;
        out8[4 * i + 3] = a8;
    }
}

/* Assets/Shaders/Source/Basic/PixelShader.hlsl:12 (Texture.Sample), sampler nearest +
 * repeat (Graphics/Renderer.cs:181-191), full-screen quad uv (0,0)..(1,1) with no flip
 * (Renderer.cs:45-56), target B8G8R8A8Srgb (SwapChain.cs:154-165). */
void pto_present(const uint8_t *src_rgba8, uint32_t sw, uint32_t sh,
                 uint8_t *dst_bgra8, uint32_t dw, uint32_t dh, uint32_t flags)
{
#ifdef _OPENMP
#pragma omp parallel for schedule(static)
#endif
    for (int64_t jj = 0; jj < (int64_t)dh; ++jj) {
        const uint32_t j = (uint32_t)jj;
        for (uint32_t i = 0; i < dw; ++i) {
            /* nearest texel of the pixel-centre uv; exact integer form of
             * floor(((i+0.5)/dw) * sw) = floor((2i+1)*sw / (2dw)) */
            uint32_t sx = (uint32_t)((((uint64_t)2 * i + 1) * sw) / ((uint64_t)2 * dw)) % sw;
            uint32_t sy = (uint32_t)((((uint64_t)2 * j + 1) * sh) / ((uint64_t)2 * dh)) % sh;
            const uint8_t *s = src_rgba8 + 4 * ((size_t)sy * sw + sx);
            uint8_t *d = dst_bgra8 + 4 * ((size_t)j * dw + i);
            uint8_t r = s[0], g = s[1], b = s[2];
            if (flags & PTO_RESOLVE_SRGB) {
                r = pto_srgb8((float)r / 255.0f);
                g = pto_srgb8((float)g / 255.0f);
                b = pto_srgb8((float)b / 255.0f);
            }
            d[0] = b; d[1] = g; d[2] = r; d[3] = s[3];
        }
    }
}
