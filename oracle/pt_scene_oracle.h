/*
 * pt_scene_oracle.h — CPU ORACLE for the scene path (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 *
 * Covers the north_star stages that have NO reference counterpart (SURVEY.md §8 a', rows
 * n3-n7): scene upload, acceleration structure, closest-hit / any-hit queries, ray-sphere and
 * ray-triangle intersection, Lambert / metal / dielectric scattering, multi-bounce paths.
 * The reference's only tracing arithmetic is RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:6-39
 * (one analytic sphere, normal visualisation, depth 0).  This file generalises exactly those
 * lines (camera ray :6-10, sphere quadratic :16-27, normal colour :28-31, miss colour :36) and
 * adds the rest by a spec FROZEN HERE.  PARITY PIN STATUS: "parity unpinned" with respect to
 * the reference (nothing to pin to), except that the default scene in PTO_SHADE_NORMALS mode
 * must reproduce pto_csmain_frame (pt_oracle.c, pinned to SURVEY App. B) bit for bit — which
 * tests/test_scene_oracle.py checks.
 *
 * Frozen spec (every float op is IEEE binary32 RNE, no contraction, left to right):
 *   camera     uv = (p/div)*2-1 (Test.hlsl:7, p = pixel index + jitter);
 *              raw_i = (uv.x*U_i + uv.y*V_i) + W_i ; d = normalize(raw) (Test.hlsl:10 has
 *              U=(1,0,0) V=(0,1,0) W=(0,0,-1)); origin = cam_o (Test.hlsl:8).
 *   sphere     oc=o-c; a=d.d; b=2*(oc.d); cc=oc.oc-r*r; disc=b*b-(4*a)*cc; hit iff disc>0
 *              (Test.hlsl:16-24); t=(-b-sqrt(disc))/(2*a) (Test.hlsl:27); if t is not inside
 *              (t_min,t_max) the far root (-b+sqrt(disc))/(2*a) is tried (no counterpart).
 *   triangle   Moeller-Trumbore, two-sided: e1=v1-v0, e2=v2-v0, p=d x e2, det=e1.p, det==0 ->
 *              miss, inv=1/det, tv=o-v0, u=(tv.p)*inv in [0,1], q=tv x e1, v=(d.q)*inv >= 0,
 *              u+v <= 1, t=(e2.q)*inv in (t_min,t_max).  NaN fails every test.
 *   closest    smallest t; equal t -> smallest primitive id (spheres 0..ns-1, triangles
 *              ns..ns+nt-1).  The result therefore does not depend on the acceleration
 *              structure or the visiting order, as long as node culling is conservative.
 *   normals    sphere: normalize(P-c) (Test.hlsl:28-29), P=o+t*d; triangle: normalize(e1 x e2).
 *   RNG        pcg3d(pixel, sample, seed + k*0x9E3779B9): k=0 -> jitter (x,y words, identical
 *              to pto_accumulate), k=depth+1 -> (u1,u2,u3) of the scatter at that depth.
 *   scatter    see pt_scene_oracle.c (pto_scatter); sin/cos come from a fixed polynomial so
 *              that CPU and GPU evaluate the same operations.
 *   radiance   sky-lit: a path that escapes adds throughput*sky(d) to the pixel; a path that
 *              is absorbed or still alive after max_depth scatters adds nothing.  Each sample
 *              adds (L.r, L.g, L.b, 1) to the pixel's float4 accumulator, in sample order.
 *   sun        (optional, sun_on != 0; no reference counterpart) next-event estimation towards one
 *              directional light: at every hit that scatters on a LAMBERT material, with n the
 *              normal turned against the incoming ray, c = n.sun_dir (sun_dir used as given);
 *              if c > 0 and pto_any_hit(P, sun_dir, t_min, +inf) finds nothing, the path's radiance
 *              gains L_k = L_k + (thr_k*albedo_k)*(sun_rad_k*c), thr = the throughput BEFORE this
 *              scatter.  With the sun on, L starts at +0 and every contribution (also the sky term
 *              of the escaping segment) is L = L + x, in path order; the sky does not contain the
 *              sun, so nothing is counted twice.  Metal and dielectric hits get no sun term.
 */
#ifndef PT_SCENE_ORACLE_H
#define PT_SCENE_ORACLE_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PTO_MAT_LAMBERT    0u
#define PTO_MAT_METAL      1u
#define PTO_MAT_DIELECTRIC 2u

#define PTO_SHADE_PATH    0u   /* multi-bounce path tracing, sky light */
#define PTO_SHADE_NORMALS 1u   /* Test.hlsl:28-36: hit -> n*0.5+0.5, miss -> (uv,0) */

typedef struct pto_material {   /* 32 bytes, same layout as ptb_material */
    uint32_t type;
    float    albedo[3];
    float    fuzz;
    float    ior;
    float    pad[2];
} pto_material;

typedef struct pto_scene {
    uint32_t width, height;
    float    uv_div_x, uv_div_y;
    float    cam_o[3], cam_u[3], cam_v[3], cam_w[3];
    float    sky_horizon[3], sky_zenith[3];
    float    t_min;
    uint32_t shade_mode;
    uint32_t max_depth;
    uint32_t n_spheres;
    const float    *spheres;       /* n_spheres x (cx,cy,cz,r) */
    const uint32_t *sphere_mat;    /* n_spheres */
    uint32_t n_tris;
    const float    *tris;          /* n_tris x (v0,v1,v2) = 9 floats */
    const uint32_t *tri_mat;       /* n_tris */
    uint32_t n_materials;
    const pto_material *materials;
    uint32_t sun_on;               /* next-event estimation towards a directional light */
    float    sun_dir[3], sun_rad[3];
} pto_scene;

typedef struct pto_accel pto_accel;   /* CPU BVH (median split); NULL everywhere = brute force */

pto_accel *pto_accel_build(const pto_scene *s, uint32_t max_leaf);
void       pto_accel_free(pto_accel *a);

/* Single ray queries.  Return 1 on hit.  accel may be NULL (brute force over all primitives). */
int pto_closest_hit(const pto_scene *s, const pto_accel *a, const float o[3], const float d[3],
                    float t_min, float t_max, int32_t *prim, float *t);
int pto_any_hit(const pto_scene *s, const pto_accel *a, const float o[3], const float d[3],
                float t_min, float t_max);

/* Batch: n rays, origins/dirs n x 3.  mode 0 = closest (out_prim = id or -1, out_t = t or 0),
 * mode 1 = any (out_prim = 1/0, out_t untouched). */
void pto_trace_rays(const pto_scene *s, const pto_accel *a, uint64_t n, const float *origins,
                    const float *dirs, float t_min, float t_max, int mode, int32_t *out_prim,
                    float *out_t, int nthreads);

/* One sample of one pixel.  Writes rgb[3]; returns the primary-ray primitive id (-1 = miss). */
int32_t pto_scene_sample(const pto_scene *s, const pto_accel *a, uint32_t px, uint32_t py,
                         uint32_t sample, uint32_t seed, int jitter, float rgb[3]);

/* Whole frame: sample 0, no jitter.  out_f32 = (rgb,1), out_rgba8 = UNORM8 store conversion,
 * out_hit = primary-ray primitive ids.  Any pointer may be NULL.  Returns #hit pixels. */
int64_t pto_scene_frame(const pto_scene *s, const pto_accel *a, uint32_t seed, float *out_f32,
                        uint8_t *out_rgba8, int32_t *out_hit, int nthreads);

/* accum[pixel] += (rgb,1) for samples first..first+n-1, in sample order. */
/* Debugging aid: one path with every segment recorded as 8 floats (origin, direction, t or +inf, primitive id or -1). */
int pto_scene_sample_path(const pto_scene *s, const pto_accel *a, uint32_t px, uint32_t py, uint32_t sample, uint32_t seed,
                          int jitter, float rgb[3], float *segs, int cap);
void pto_scene_accumulate(const pto_scene *s, const pto_accel *a, uint32_t first_sample,
                          uint32_t n_samples, uint32_t seed, int jitter, float *accum, int nthreads);
/* One interleaved row shard only (ptb_set_row_shard_bands semantics); other rows of accum are untouched. */
void pto_scene_accumulate_rows(const pto_scene *s, const pto_accel *a, uint32_t first_sample,
                               uint32_t n_samples, uint32_t seed, int jitter, float *accum,
                               uint32_t row_offset, uint32_t row_stride, uint32_t row_block, int nthreads);

/* Exposed pieces of the spec (unit-tested on their own, restated by the CUDA path). */
void pto_sincos2pi(float u, float *sn, float *cs);
void pto_unit_vector(float u1, float u2, float v[3]);
/* returns 1 and the new direction / attenuation, or 0 when the path is absorbed */
int  pto_scatter(const pto_material *m, const float d[3], const float n_out[3],
                 float u1, float u2, float u3, float d_new[3], float att[3]);

#ifdef __cplusplus
}
#endif
#endif
