/*
 * pt_oracle.h — CPU ORACLE (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * Plain-C restatement of the reference's tracing path, used only as the checker
 * by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs.  Nothing under pathtracing_b200/ may include, link or load this.
 *
 * Reference path restated: RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:1-40
 * (CSMain), store format per RayTracing/Graphics/Renderer.cs:124 (R8G8B8A8Unorm),
 * present path per Assets/Shaders/Source/Basic/PixelShader.hlsl:8-19 and
 * Graphics/SwapChain.cs:154-165 (B8G8R8A8Srgb).
 *
 * PARITY PIN STATUS: the reference ships no tests, golden vectors or CPU implementation
 *  and its HLSL/Vulkan path cannot run on this box.  What it DOES hold is
 * the compiled module it loads, Assets/Shaders/Compiled/Ray/Test.spirv (sha256 051e3943...c9f0f).
 * The oracle is pinned to that file (tests/test_spirv_pin.py, oracle/spirv_eval.py): the
 * module's words are decoded, its constants (fl(1/1080) = 0x3A72B9D6, -3.0, -2.0, 0.5) and
 * operation order asserted, and its operation list evaluated instruction by instruction in
 * numpy float32.  Result: pto_csmain_frame_spirv (the "as-compiled" variant below) equals that
 * evaluation bit for bit; pto_csmain_frame (the frozen source-order spec the CUDA kernels
 * reproduce) has the IDENTICAL hit mask at 320x180, 1920x1080, 1920x1088 and 3840x2160 and
 * radiance within 3.44e-5 relative at 1080p (north_star bar: 1e-4).  The precision of
 * Normalize / Sqrt / FDiv / Dot is implementation-defined in SPIR-V and fixed here by the
 * rounding spec, so the pin is "same operation list, stated roundings", not a run of a Vulkan
 * driver.  Additionally pinned to the survey-derived known answers of SURVEY.md Appendix B
 * (hit counts, hit bbox, per-pixel float bit patterns).  Stages n1-n8 (jitter, accumulation,
 * scenes) have no reference counterpart at all and stay "parity unpinned".
 *
 * Rounding spec (frozen): IEEE-754 binary32, round-to-nearest-even on every single
 * operation, NO fused multiply-add unless fmaf() is written explicitly, left-to-right
 * evaluation, dot(x,y) = ((x0*y0 + x1*y1) + x2*y2), normalize(v) = v / sqrtf(dot(v,v))
 * (three correctly-rounded divisions).  Build with -O2 -ffp-contract=off, never
 * -ffast-math.
 */
#ifndef PT_ORACLE_H
#define PT_ORACLE_H

#include <stdint.h>
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Parameters of the literal CSMain path.  Defaults (pto_default_frame) reproduce the
 * literals of Test.hlsl:6,8,12,13 for a height-H image with divisor (H,H). */
typedef struct pto_frame_params {
    uint32_t width, height;
    float    uv_div_x, uv_div_y;   /* Test.hlsl:6  resolution = (1080,1080) */
    float    cam[3];               /* Test.hlsl:8  (0,0,1) */
    float    centre[3];            /* Test.hlsl:12 (0,0,0) */
    float    radius;               /* Test.hlsl:13 0.5 */
} pto_frame_params;

void pto_default_frame(pto_frame_params *p, uint32_t width, uint32_t height);

/* One pixel of CSMain (Test.hlsl:6-37) at continuous pixel position (px,py)
 * (px = (float)id.x for the literal path).  Writes colour[4]; optional outs. */
int pto_csmain_sample(const pto_frame_params *p, float px, float py,
                      float colour[4], float *disc_out, float *t_out);

/* Whole frame, 1 spp, no jitter (a1-a5).  Any output pointer may be NULL.
 *   out_f32   : W*H*4 floats (the shader-side float4, Test.hlsl:39)
 *   out_rgba8 : W*H*4 bytes, UNORM8 store conversion (Renderer.cs:124)
 *   out_hit   : W*H int32, 0 = sphere, -1 = miss
 * nthreads <= 0 -> all OpenMP threads.  Returns number of hit pixels. */
int64_t pto_csmain_frame(const pto_frame_params *p, float *out_f32, uint8_t *out_rgba8,
                         int32_t *out_hit, int nthreads);

/* The "as-compiled" variant: the operation list of the reference's checked-in
 * Assets/Shaders/Compiled/Ray/Test.spirv (reciprocal multiply for the uv division, the two Fma's, folded
 * constants), literal scene only.  Same hit mask as the frozen source-order spec at every size tested,
 * radiance within 3.4e-5 relative at 1080p (tests/test_spirv_pin.py). */
int pto_csmain_sample_spirv(const pto_frame_params *p, float px, float py,
                            float colour[4], float *disc_out, float *t_out);
int64_t pto_csmain_frame_spirv(const pto_frame_params *p, float *out_f32, uint8_t *out_rgba8,
                               int32_t *out_hit, int nthreads);

/* f32 -> UNORM8 store conversion: NaN -> 0, clamp to [0,1], *255, RNE. */
uint8_t pto_unorm8(float v);
/* sRGB OETF then UNORM8 (what a B8G8R8A8Srgb attachment does on write). */
uint8_t pto_srgb8(float v);

/* ---- n1: counter-based RNG (frozen spec, restated independently by the CUDA path) ----
 * pcg3d (Jarzynski & Olano, JCGT 2020) keyed (pixel index, sample index, seed); a random
 * 32-bit word w maps to the unit interval as (w >> 9) * 2^-23 (exact, 23 bits). */
void  pto_pcg3d(uint32_t v[3]);
float pto_bits_to_unit(uint32_t w);

/* ---- n1/n2: jittered multi-sample accumulation of the CSMain path ----
 * For s in [first_sample, first_sample+n_samples): v = pcg3d(y*W+x, s, seed);
 * jx = unit(v.x), jy = unit(v.y); colour = csmain((float)x + jx, (float)y + jy).
 * Samples are summed sequentially in s into a float4 partial starting from 0, and
 * the partial is then added to accum[pixel] (one add per launch).
 * jitter == 0 -> jx = jy = 0 (reduces to a1-a5 exactly). */
void pto_accumulate(const pto_frame_params *p, uint32_t first_sample, uint32_t n_samples,
                    uint32_t seed, int jitter, float *accum, int nthreads);
/* variant 0 = pto_accumulate; 1 = the same loop over the as-compiled sample (pto_csmain_sample_spirv). */
void pto_accumulate_v(const pto_frame_params *p, int variant, uint32_t first_sample, uint32_t n_samples,
                      uint32_t seed, int jitter, float *accum, int nthreads);

/* resolve: out = accum * (1/spp) via division accum/ (float)spp, then UNORM8 (flags&1: sRGB
 * OETF first), channel order RGBA (flags&2: BGRA). */
#define PTO_RESOLVE_SRGB 1u
#define PTO_RESOLVE_BGRA 2u
void pto_resolve(const float *accum, uint32_t width, uint32_t height, uint32_t spp_total,
                 uint32_t flags, uint8_t *out8);

/* f1: present-equivalent.  Nearest sample of an RGBA8 UNORM source (sw x sh) onto a
 * dw x dh BGRA8 target; texel centre uv = ((i+0.5)/dw, (j+0.5)/dh), nearest = floor(uv*sw),
 * repeat addressing (Renderer.cs:181-191); flags&1 -> apply sRGB OETF (SwapChain.cs:158). */
void pto_present(const uint8_t *src_rgba8, uint32_t sw, uint32_t sh,
                 uint8_t *dst_bgra8, uint32_t dw, uint32_t dh, uint32_t flags);

int pto_max_threads(void);
void pto_set_threads(int n);

#ifdef __cplusplus
}
#endif
#endif
