/*
 * pt_scene_oracle.c — CPU ORACLE for the scene path (TEST INFRASTRUCTURE, NOT PRODUCT CODE).
 * Spec, scope and parity-pin status: pt_scene_oracle.h.  The camera ray, the sphere quadratic
 * and the normal/miss colours generalise RayTracing/Assets/Shaders/Source/Ray/Test.hlsl:6-36
 * (paths relative to /root/reference/); everything else has no reference counterpart.
 *
 * Build: gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp (oracle/Makefile).
 */
#include "pt_scene_oracle.h"
#include "pt_oracle.h"

#include <math.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ---------------------------------------------------------------- small vector helpers */
static inline float dot3(const float a[3], const float b[3])
{
    float s = a[0] * b[0];
    s = s + a[1] * b[1];
    s = s + a[2] * b[2];
    return s;
}
static inline void cross3(const float a[3], const float b[3], float o[3])
{
    float x = a[1] * b[2] - a[2] * b[1];
    float y = a[2] * b[0] - a[0] * b[2];
    float z = a[0] * b[1] - a[1] * b[0];
    o[0] = x; o[1] = y; o[2] = z;
}
static inline void normalize3(const float v[3], float o[3])
{
    float len = sqrtf(dot3(v, v));
    o[0] = v[0] / len; o[1] = v[1] / len; o[2] = v[2] / len;
}

/* ---------------------------------------------------------------- primitive tests */
/* Test.hlsl:16-27 generalised (far root when the near one is outside the interval). */
static int hit_sphere(const float *cr, const float o[3], const float d[3], float t_min, float t_max, float *t_out)
{
    float oc[3] = { o[0] - cr[0], o[1] - cr[1], o[2] - cr[2] };   /* :16 */
    float a = dot3(d, d);                                        /* :17 */
    float b = 2.0f * dot3(oc, d);                                /* :18 */
    float cc = dot3(oc, oc) - cr[3] * cr[3];                     /* :19 */
    float disc = b * b - (4.0f * a) * cc;                        /* :21 */
    if (!(disc > 0.0f)) return 0;                                /* :24 strict */
    float s = sqrtf(disc);
    float a2 = 2.0f * a;
    float t = ((-b) - s) / a2;                                   /* :27 */
    if (!(t > t_min && t < t_max)) {
        t = ((-b) + s) / a2;
        if (!(t > t_min && t < t_max)) return 0;
    }
    *t_out = t;
    return 1;
}

static int hit_triangle(const float *v, const float o[3], const float d[3], float t_min, float t_max, float *t_out)
{
    float e1[3] = { v[3] - v[0], v[4] - v[1], v[5] - v[2] };
    float e2[3] = { v[6] - v[0], v[7] - v[1], v[8] - v[2] };
    float p[3], q[3];
    cross3(d, e2, p);
    float det = dot3(e1, p);
    if (!(det != 0.0f)) return 0;          /* zero or NaN */
    float inv = 1.0f / det;
    float tv[3] = { o[0] - v[0], o[1] - v[1], o[2] - v[2] };
    float u = dot3(tv, p) * inv;
    if (!(u >= 0.0f && u <= 1.0f)) return 0;
    cross3(tv, e1, q);
    float w = dot3(d, q) * inv;
    if (!(w >= 0.0f && u + w <= 1.0f)) return 0;
    float t = dot3(e2, q) * inv;
    if (!(t > t_min && t < t_max)) return 0;
    *t_out = t;
    return 1;
}

static inline int hit_prim(const pto_scene *s, uint32_t id, const float o[3], const float d[3], float t_min,
                           float t_max, float *t)
{
    if (id < s->n_spheres) return hit_sphere(s->spheres + 4 * (size_t)id, o, d, t_min, t_max, t);
    return hit_triangle(s->tris + 9 * (size_t)(id - s->n_spheres), o, d, t_min, t_max, t);
}

/* ---------------------------------------------------------------- CPU BVH (median split) */
typedef struct bnode {
    float lo[3], hi[3];
    int32_t left, right;      /* children (internal) */
    uint32_t first, count;    /* leaf range in order[] when count > 0 */
} bnode;

struct pto_accel {
    bnode *nodes;
    uint32_t n_nodes, cap;
    uint32_t *order;
    float *blo, *bhi, *cen;   /* per primitive */
    uint32_t max_leaf;
};

static void prim_bounds(const pto_scene *s, uint32_t id, float lo[3], float hi[3])
{
    if (id < s->n_spheres) {
        const float *c = s->spheres + 4 * (size_t)id;
        float r = fabsf(c[3]);
        for (int k = 0; k < 3; ++k) { lo[k] = c[k] - r; hi[k] = c[k] + r; }
    } else {
        const float *v = s->tris + 9 * (size_t)(id - s->n_spheres);
        for (int k = 0; k < 3; ++k) {
            lo[k] = fminf(v[k], fminf(v[3 + k], v[6 + k]));
            hi[k] = fmaxf(v[k], fmaxf(v[3 + k], v[6 + k]));
        }
    }
}

static int g_axis;
static const float *g_cen;
static int cmp_axis(const void *a, const void *b)
{
    float x = g_cen[3 * (size_t)(*(const uint32_t *)a) + g_axis];
    float y = g_cen[3 * (size_t)(*(const uint32_t *)b) + g_axis];
    if (x < y) return -1;
    if (x > y) return 1;
    uint32_t ia = *(const uint32_t *)a, ib = *(const uint32_t *)b;
    return ia < ib ? -1 : (ia > ib);
}

/* nth_element-free median split: sorts the range on the widest centroid axis (build time is
 * irrelevant for a checker; 1M primitives take a few seconds). */
static int32_t build_rec(pto_accel *a, uint32_t first, uint32_t count, float pad)
{
    if (a->n_nodes == a->cap) return -1;
    int32_t me = (int32_t)a->n_nodes++;
    bnode *n = &a->nodes[me];
    float clo[3] = { INFINITY, INFINITY, INFINITY }, chi[3] = { -INFINITY, -INFINITY, -INFINITY };
    for (int k = 0; k < 3; ++k) { n->lo[k] = INFINITY; n->hi[k] = -INFINITY; }
    for (uint32_t i = first; i < first + count; ++i) {
        uint32_t id = a->order[i];
        for (int k = 0; k < 3; ++k) {
            n->lo[k] = fminf(n->lo[k], a->blo[3 * (size_t)id + k]);
            n->hi[k] = fmaxf(n->hi[k], a->bhi[3 * (size_t)id + k]);
            clo[k] = fminf(clo[k], a->cen[3 * (size_t)id + k]);
            chi[k] = fmaxf(chi[k], a->cen[3 * (size_t)id + k]);
        }
    }
    for (int k = 0; k < 3; ++k) { n->lo[k] -= pad; n->hi[k] += pad; }
    n->left = n->right = -1;
    n->first = first;
    n->count = count;
    if (count <= a->max_leaf) return me;
    int axis = 0;
    if (chi[1] - clo[1] > chi[axis] - clo[axis]) axis = 1;
    if (chi[2] - clo[2] > chi[axis] - clo[axis]) axis = 2;
    g_axis = axis;
    g_cen = a->cen;
    qsort(a->order + first, count, sizeof(uint32_t), cmp_axis);
    uint32_t half = count / 2;
    int32_t l = build_rec(a, first, half, pad);
    int32_t r = build_rec(a, first + half, count - half, pad);
    n = &a->nodes[me];      /* nodes[] is preallocated, but keep the habit */
    n->left = l; n->right = r; n->count = 0;
    return me;
}

pto_accel *pto_accel_build(const pto_scene *s, uint32_t max_leaf)
{
    uint32_t np = s->n_spheres + s->n_tris;
    pto_accel *a = (pto_accel *)calloc(1, sizeof *a);
    if (!a) return NULL;
    a->max_leaf = max_leaf ? max_leaf : 4;
    a->cap = np ? 2 * np : 1;
    a->nodes = (bnode *)malloc(sizeof(bnode) * a->cap);
    a->order = (uint32_t *)malloc(sizeof(uint32_t) * (np ? np : 1));
    a->blo = (float *)malloc(sizeof(float) * 3 * (np ? np : 1));
    a->bhi = (float *)malloc(sizeof(float) * 3 * (np ? np : 1));
    a->cen = (float *)malloc(sizeof(float) * 3 * (np ? np : 1));
    float ext = 0.0f;
    for (uint32_t i = 0; i < np; ++i) {
        a->order[i] = i;
        prim_bounds(s, i, a->blo + 3 * (size_t)i, a->bhi + 3 * (size_t)i);
        for (int k = 0; k < 3; ++k) {
            a->cen[3 * (size_t)i + k] = 0.5f * (a->blo[3 * (size_t)i + k] + a->bhi[3 * (size_t)i + k]);
            ext = fmaxf(ext, fmaxf(fabsf(a->blo[3 * (size_t)i + k]), fabsf(a->bhi[3 * (size_t)i + k])));
        }
    }
    /* conservative culling: boxes grow by 2^-18 of the scene's largest coordinate */
    float pad = ext * 3.814697265625e-6f;
    if (np) build_rec(a, 0, np, pad);
    return a;
}

void pto_accel_free(pto_accel *a)
{
    if (!a) return;
    free(a->nodes); free(a->order); free(a->blo); free(a->bhi); free(a->cen); free(a);
}

static inline int slab(const bnode *n, const float o[3], const float inv[3], float t_min, float t_far)
{
    float t0 = t_min, t1 = t_far;
    for (int k = 0; k < 3; ++k) {
        float a = (n->lo[k] - o[k]) * inv[k];
        float b = (n->hi[k] - o[k]) * inv[k];
        float lo = fminf(a, b), hi = fmaxf(a, b);   /* fminf/fmaxf drop NaN (0*inf) */
        t0 = fmaxf(t0, lo * 0.9999990f);
        t1 = fminf(t1, hi * 1.0000010f);
    }
    return t0 <= t1;
}

int pto_closest_hit(const pto_scene *s, const pto_accel *a, const float o[3], const float d[3],
                    float t_min, float t_max, int32_t *prim, float *t_out)
{
    int32_t best = -1;
    float best_t = t_max;
    uint32_t np = s->n_spheres + s->n_tris;
    if (!a) {
        for (uint32_t id = 0; id < np; ++id) {
            float t;
            if (hit_prim(s, id, o, d, t_min, t_max, &t))
                if (best < 0 || t < best_t) { best = (int32_t)id; best_t = t; }   /* ascending ids: ties keep the smaller */
        }
    } else if (np) {
        float inv[3] = { 1.0f / d[0], 1.0f / d[1], 1.0f / d[2] };
        int32_t stack[128];
        int sp = 0;
        stack[sp++] = 0;
        while (sp) {
            const bnode *n = &a->nodes[stack[--sp]];
            if (!slab(n, o, inv, t_min, best < 0 ? t_max : best_t)) continue;
            if (n->count) {
                for (uint32_t i = n->first; i < n->first + n->count; ++i) {
                    uint32_t id = a->order[i];
                    float t;
                    if (hit_prim(s, id, o, d, t_min, t_max, &t))
                        if (best < 0 || t < best_t || (t == best_t && (int32_t)id < best)) { best = (int32_t)id; best_t = t; }
                }
            } else {
                if (sp + 2 > 128) return -1;
                stack[sp++] = n->left;
                stack[sp++] = n->right;
            }
        }
    }
    if (prim) *prim = best;
    if (t_out) *t_out = best >= 0 ? best_t : 0.0f;
    return best >= 0;
}

int pto_any_hit(const pto_scene *s, const pto_accel *a, const float o[3], const float d[3],
                float t_min, float t_max)
{
    uint32_t np = s->n_spheres + s->n_tris;
    float t;
    if (!a) {
        for (uint32_t id = 0; id < np; ++id)
            if (hit_prim(s, id, o, d, t_min, t_max, &t)) return 1;
        return 0;
    }
    if (!np) return 0;
    float inv[3] = { 1.0f / d[0], 1.0f / d[1], 1.0f / d[2] };
    int32_t stack[128];
    int sp = 0;
    stack[sp++] = 0;
    while (sp) {
        const bnode *n = &a->nodes[stack[--sp]];
        if (!slab(n, o, inv, t_min, t_max)) continue;
        if (n->count) {
            for (uint32_t i = n->first; i < n->first + n->count; ++i)
                if (hit_prim(s, a->order[i], o, d, t_min, t_max, &t)) return 1;
        } else {
            stack[sp++] = n->left;
            stack[sp++] = n->right;
        }
    }
    return 0;
}

void pto_trace_rays(const pto_scene *s, const pto_accel *a, uint64_t n, const float *origins,
                    const float *dirs, float t_min, float t_max, int mode, int32_t *out_prim,
                    float *out_t, int nthreads)
{
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 256) num_threads(nthreads)
#endif
    for (int64_t i = 0; i < (int64_t)n; ++i) {
        if (mode == 0) {
            int32_t p; float t;
            pto_closest_hit(s, a, origins + 3 * i, dirs + 3 * i, t_min, t_max, &p, &t);
            out_prim[i] = p;
            if (out_t) out_t[i] = t;
        } else {
            out_prim[i] = pto_any_hit(s, a, origins + 3 * i, dirs + 3 * i, t_min, t_max);
        }
    }
}

/* ---------------------------------------------------------------- sampling spec */
/* sin/cos of 2*pi*u, u in [0,1): quadrant q = floor(4u), a = (frac(4u)-0.5)*pi/2 in
 * [-pi/4,pi/4), Taylor polynomials in Horner form (each step one multiply then one add),
 * rotated by pi/4 and then by the quadrant.  |error| < 1e-7. */
void pto_sincos2pi(float u, float *sn, float *cs)
{
    float x = u * 4.0f;
    int q = (int)x;
    float f = x - (float)q;
    float a = (f - 0.5f) * 1.57079637f;
    float a2 = a * a;
    float ps = 2.75573192e-6f;                 /* 1/9! */
    ps = ps * a2 + -1.98412698e-4f;            /* -1/7! */
    ps = ps * a2 + 8.33333333e-3f;             /* 1/5! */
    ps = ps * a2 + -1.66666667e-1f;            /* -1/3! */
    ps = ps * a2 + 1.0f;
    float s = a * ps;
    float pc = 2.48015873e-5f;                 /* 1/8! */
    pc = pc * a2 + -1.38888889e-3f;            /* -1/6! */
    pc = pc * a2 + 4.16666667e-2f;             /* 1/4! */
    pc = pc * a2 + -0.5f;
    pc = pc * a2 + 1.0f;
    float c = pc;
    float sb = (s + c) * 0.707106769f;
    float cb = (c - s) * 0.707106769f;
    switch (q & 3) {
    case 0: *sn = sb; *cs = cb; break;
    case 1: *sn = cb; *cs = -sb; break;
    case 2: *sn = -sb; *cs = -cb; break;
    default: *sn = -cb; *cs = sb; break;
    }
}

/* uniform direction on the unit sphere from two unit-interval numbers */
void pto_unit_vector(float u1, float u2, float v[3])
{
    float z = 1.0f - 2.0f * u1;
    float r2 = 1.0f - z * z;
    float r = sqrtf(fmaxf(r2, 0.0f));
    float sn, cs;
    pto_sincos2pi(u2, &sn, &cs);
    v[0] = r * cs;
    v[1] = r * sn;
    v[2] = z;
}

int pto_scatter(const pto_material *m, const float d[3], const float n_out[3],
                float u1, float u2, float u3, float d_new[3], float att[3])
{
    float dn_out = dot3(d, n_out);
    int front = dn_out < 0.0f;
    float n[3] = { front ? n_out[0] : -n_out[0], front ? n_out[1] : -n_out[1], front ? n_out[2] : -n_out[2] };
    float dir[3];
    att[0] = m->albedo[0]; att[1] = m->albedo[1]; att[2] = m->albedo[2];
    if (m->type == PTO_MAT_LAMBERT) {
        float v[3];
        pto_unit_vector(u1, u2, v);
        dir[0] = n[0] + v[0]; dir[1] = n[1] + v[1]; dir[2] = n[2] + v[2];
        if (fabsf(dir[0]) < 1e-8f && fabsf(dir[1]) < 1e-8f && fabsf(dir[2]) < 1e-8f) {
            dir[0] = n[0]; dir[1] = n[1]; dir[2] = n[2];
        }
    } else if (m->type == PTO_MAT_METAL) {
        float v[3], refl[3], r[3];
        pto_unit_vector(u1, u2, v);
        float dn = dot3(d, n);
        float k = 2.0f * dn;
        refl[0] = d[0] - k * n[0]; refl[1] = d[1] - k * n[1]; refl[2] = d[2] - k * n[2];
        normalize3(refl, r);
        dir[0] = r[0] + m->fuzz * v[0]; dir[1] = r[1] + m->fuzz * v[1]; dir[2] = r[2] + m->fuzz * v[2];
        if (!(dot3(dir, n) > 0.0f)) return 0;           /* scattered below the surface: absorbed */
    } else {
        float ri = front ? 1.0f / m->ior : m->ior;
        float dn = dot3(d, n);
        float cs = fminf(-dn, 1.0f);
        float sn = sqrtf(fmaxf(1.0f - cs * cs, 0.0f));
        int cannot = ri * sn > 1.0f;
        float r0 = (1.0f - ri) / (1.0f + ri);
        r0 = r0 * r0;
        float mm = 1.0f - cs;
        float m2 = mm * mm;
        float m5 = (m2 * m2) * mm;
        float refl = r0 + (1.0f - r0) * m5;
        if (cannot || refl > u3) {
            float k = 2.0f * dn;
            dir[0] = d[0] - k * n[0]; dir[1] = d[1] - k * n[1]; dir[2] = d[2] - k * n[2];
        } else {
            float perp[3] = { ri * (d[0] + cs * n[0]), ri * (d[1] + cs * n[1]), ri * (d[2] + cs * n[2]) };
            float k = -sqrtf(fabsf(1.0f - dot3(perp, perp)));
            dir[0] = perp[0] + k * n[0]; dir[1] = perp[1] + k * n[1]; dir[2] = perp[2] + k * n[2];
        }
    }
    normalize3(dir, d_new);
    return 1;
}

/* ---------------------------------------------------------------- paths */
static void prim_normal(const pto_scene *s, int32_t id, const float P[3], float n[3])
{
    if ((uint32_t)id < s->n_spheres) {
        const float *c = s->spheres + 4 * (size_t)id;
        float pc[3] = { P[0] - c[0], P[1] - c[1], P[2] - c[2] };   /* Test.hlsl:29 */
        normalize3(pc, n);
    } else {
        const float *v = s->tris + 9 * (size_t)((uint32_t)id - s->n_spheres);
        float e1[3] = { v[3] - v[0], v[4] - v[1], v[5] - v[2] };
        float e2[3] = { v[6] - v[0], v[7] - v[1], v[8] - v[2] };
        float g[3];
        cross3(e1, e2, g);
        normalize3(g, n);
    }
}

static int32_t scene_sample_core(const pto_scene *s, const pto_accel *a, uint32_t px, uint32_t py,
                                 uint32_t sample, uint32_t seed, int jitter, float rgb[3], float *segs, int cap, int *nseg);

int32_t pto_scene_sample(const pto_scene *s, const pto_accel *a, uint32_t px, uint32_t py,
                         uint32_t sample, uint32_t seed, int jitter, float rgb[3])
{
    return scene_sample_core(s, a, px, py, sample, seed, jitter, rgb, NULL, 0, NULL);
}

/* Debugging aid: the same path, recording every segment as 8 floats (origin, direction, t or +inf on a miss, primitive id
 * or -1); returns the number of segments written (<= cap). */
int pto_scene_sample_path(const pto_scene *s, const pto_accel *a, uint32_t px, uint32_t py, uint32_t sample, uint32_t seed,
                          int jitter, float rgb[3], float *segs, int cap)
{
    int n = 0;
    scene_sample_core(s, a, px, py, sample, seed, jitter, rgb, segs, cap, &n);
    return n;
}

static int32_t scene_sample_core(const pto_scene *s, const pto_accel *a, uint32_t px, uint32_t py,
                                 uint32_t sample, uint32_t seed, int jitter, float rgb[3], float *segs, int cap, int *nseg)
{
    const uint32_t pix = py * s->width + px;
    float jx = 0.0f, jy = 0.0f;
    if (jitter) {
        uint32_t v[3] = { pix, sample, seed };
        pto_pcg3d(v);
        jx = pto_bits_to_unit(v[0]);
        jy = pto_bits_to_unit(v[1]);
    }
    float fx = (float)px + jx, fy = (float)py + jy;
    float uvx = (fx / s->uv_div_x) * 2.0f - 1.0f;               /* Test.hlsl:7 */
    float uvy = (fy / s->uv_div_y) * 2.0f - 1.0f;
    float raw[3], d[3], o[3];
    for (int k = 0; k < 3; ++k) {
        raw[k] = (uvx * s->cam_u[k] + uvy * s->cam_v[k]) + s->cam_w[k];
        o[k] = s->cam_o[k];                                     /* Test.hlsl:8 */
    }
    normalize3(raw, d);                                         /* Test.hlsl:10 */
    float thr[3] = { 1.0f, 1.0f, 1.0f };
    int32_t primary = -1;
    rgb[0] = rgb[1] = rgb[2] = 0.0f;
    for (uint32_t depth = 0;; ++depth) {
        int32_t prim; float t;
        int hit = pto_closest_hit(s, a, o, d, s->t_min, INFINITY, &prim, &t);
        if (segs && *nseg < cap) {
            float *g = segs + 8 * (*nseg)++;
            g[0] = o[0]; g[1] = o[1]; g[2] = o[2]; g[3] = d[0]; g[4] = d[1]; g[5] = d[2];
            g[6] = hit ? t : INFINITY; g[7] = hit ? (float)prim : -1.0f;
        }
        if (depth == 0) primary = hit ? prim : -1;
        float P[3], n_out[3];
        if (hit) {
            for (int k = 0; k < 3; ++k) P[k] = o[k] + t * d[k];  /* Test.hlsl:28 */
            prim_normal(s, prim, P, n_out);
        }
        if (s->shade_mode == PTO_SHADE_NORMALS) {
            if (hit) {                                          /* Test.hlsl:31 */
                rgb[0] = n_out[0] * 0.5f + 0.5f; rgb[1] = n_out[1] * 0.5f + 0.5f; rgb[2] = n_out[2] * 0.5f + 0.5f;
            } else {                                            /* Test.hlsl:36 */
                rgb[0] = uvx; rgb[1] = uvy; rgb[2] = 0.0f;
            }
            return primary;
        }
        if (!hit) {
            float w = (d[1] + 1.0f) * 0.5f;
            for (int k = 0; k < 3; ++k) {
                float c = (1.0f - w) * s->sky_horizon[k] + w * s->sky_zenith[k];
                if (s->sun_on) rgb[k] = rgb[k] + thr[k] * c;
                else rgb[k] = thr[k] * c;
            }
            return primary;
        }
        if (depth == s->max_depth) return primary;              /* bounce budget spent: black */
        uint32_t mid = (uint32_t)prim < s->n_spheres ? s->sphere_mat[prim] : s->tri_mat[(uint32_t)prim - s->n_spheres];
        const pto_material *m = &s->materials[mid < s->n_materials ? mid : 0];
        uint32_t v[3] = { pix, sample, seed + (depth + 1u) * 0x9E3779B9u };
        pto_pcg3d(v);
        float u1 = pto_bits_to_unit(v[0]), u2 = pto_bits_to_unit(v[1]), u3 = pto_bits_to_unit(v[2]);
        float dn[3], att[3];
        if (s->sun_on && m->type == PTO_MAT_LAMBERT) {          /* next-event estimation (spec: pt_scene_oracle.h "sun") */
            int front = dot3(d, n_out) < 0.0f;
            float n[3] = { front ? n_out[0] : -n_out[0], front ? n_out[1] : -n_out[1], front ? n_out[2] : -n_out[2] };
            float c = dot3(n, s->sun_dir);
            if (c > 0.0f && !pto_any_hit(s, a, P, s->sun_dir, s->t_min, INFINITY))
                for (int k = 0; k < 3; ++k) rgb[k] = rgb[k] + (thr[k] * m->albedo[k]) * (s->sun_rad[k] * c);
        }
        if (!pto_scatter(m, d, n_out, u1, u2, u3, dn, att)) return primary;
        for (int k = 0; k < 3; ++k) { thr[k] = thr[k] * att[k]; o[k] = P[k]; d[k] = dn[k]; }
    }
}

int64_t pto_scene_frame(const pto_scene *s, const pto_accel *a, uint32_t seed, float *out_f32,
                        uint8_t *out_rgba8, int32_t *out_hit, int nthreads)
{
    const int64_t W = s->width, H = s->height;
    int64_t hits = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads) reduction(+ : hits)
#endif
    for (int64_t y = 0; y < H; ++y)
        for (int64_t x = 0; x < W; ++x) {
            float c[4];
            int32_t prim = pto_scene_sample(s, a, (uint32_t)x, (uint32_t)y, 0, seed, 0, c);
            c[3] = 1.0f;
            int64_t i = y * W + x;
            hits += prim >= 0;
            if (out_f32) memcpy(out_f32 + 4 * i, c, sizeof c);
            if (out_rgba8)
                for (int k = 0; k < 4; ++k) out_rgba8[4 * i + k] = pto_unorm8(c[k]);
            if (out_hit) out_hit[i] = prim;
        }
    return hits;
}

void pto_scene_accumulate(const pto_scene *s, const pto_accel *a, uint32_t first_sample,
                          uint32_t n_samples, uint32_t seed, int jitter, float *accum, int nthreads)
{
    pto_scene_accumulate_rows(s, a, first_sample, n_samples, seed, jitter, accum, 0, 1, 1, nthreads);
}

/* The same for one interleaved row shard only (the rows ptb_set_row_shard_bands selects: bands of row_block rows,
 * band g = y / row_block belongs to the shard when g % row_stride == row_offset); other rows of accum are untouched. */
void pto_scene_accumulate_rows(const pto_scene *s, const pto_accel *a, uint32_t first_sample,
                               uint32_t n_samples, uint32_t seed, int jitter, float *accum,
                               uint32_t row_offset, uint32_t row_stride, uint32_t row_block, int nthreads)
{
    const int64_t W = s->width, H = s->height;
    if (row_stride == 0) row_stride = 1;
    if (row_block == 0) row_block = 1;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#pragma omp parallel for schedule(dynamic, 1) num_threads(nthreads)
#endif
    for (int64_t y = 0; y < H; ++y)
        for (int64_t x = 0; x < W; ++x) {
            if (((uint32_t)y / row_block) % row_stride != row_offset) break;
            float *acc = accum + 4 * (y * W + x);
            for (uint32_t k = 0; k < n_samples; ++k) {
                float c[3];
                pto_scene_sample(s, a, (uint32_t)x, (uint32_t)y, first_sample + k, seed, jitter, c);
                acc[0] += c[0]; acc[1] += c[1]; acc[2] += c[2]; acc[3] += 1.0f;
            }
        }
}
