/* oracle/rt_oracle.c — TEST INFRASTRUCTURE ONLY.  See rt_oracle.h.
 *
 * Every routine cites the lines of /root/reference/kernels/raytracer.cl ("cl:") it restates.
 * Floating point: single IEEE binary32 operations in the order written (build: -O2 -ffp-contract=off).
 * The implementation-defined OpenCL builtins are pinned as in oracle/cl_shim.h:
 *   dot = x*x' + y*y' + z*z' left to right; normalize(v) = v * (1/sqrt(dot(v,v))); vector / scalar is a
 *   component-wise IEEE division; pow(x,5) = powf(x, 5.0f).
 */
#include "rt_oracle.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#define T_MIN 0.001f            /* cl:3 */
#define T_MAX 1000.0f           /* cl:4 */
#define TRI_EPS 0.0000001f      /* cl:1 */
#define TABLE_N 100000u         /* cl:7 */

/* cl:23 / include/scene.h:30 */
enum { M_REFRACTIVE = 0, M_REFLECTIVE = 1, M_DIELECTRIC = 2, M_DIFFUSE = 3, M_TEXTURED = 4, M_LIGHT = 5 };

typedef struct { float x, y, z; } v3;

/* Host AoS records (include/scene.h:32-81); float3 members are 16 bytes. */
typedef struct { int32_t type; int32_t _p0[3]; float color[4]; float extra; float _p1[3]; } a_material; /* 48 */
typedef struct { float pos[4]; float r; uint32_t mat; float _p[2]; } a_sphere;                          /* 32 */
typedef struct { float pos[4]; float normal[4]; uint32_t mat; float _p[3]; } a_plane;                   /* 48 */
typedef struct { float pos[4]; float p1[4]; float p2[4]; float r1, r2; uint32_t mat; float _p; } a_lens; /* 64 */
typedef struct { uint32_t vertex_anchor, index_anchor, face_count, texture_id; } a_mesh;                /* 16 */
typedef struct { uint32_t mesh_anchor, mesh_count, mat; } a_model;                                      /* 12 */

typedef char chk_material[sizeof(a_material) == 48 ? 1 : -1];
typedef char chk_sphere[sizeof(a_sphere) == 32 ? 1 : -1];
typedef char chk_plane[sizeof(a_plane) == 48 ? 1 : -1];
typedef char chk_lens[sizeof(a_lens) == 64 ? 1 : -1];

static inline v3 V(float x, float y, float z) { v3 r = {x, y, z}; return r; }
static inline v3 vld(const float* p) { return V(p[0], p[1], p[2]); }
static inline v3 vadd(v3 a, v3 b) { return V(a.x + b.x, a.y + b.y, a.z + b.z); }
static inline v3 vsub(v3 a, v3 b) { return V(a.x - b.x, a.y - b.y, a.z - b.z); }
static inline v3 vscale(v3 a, float s) { return V(a.x * s, a.y * s, a.z * s); }
static inline v3 vneg(v3 a) { return V(-a.x, -a.y, -a.z); }
static inline v3 vdivs(v3 a, float s) { return V(a.x / s, a.y / s, a.z / s); }
static inline float vdot(v3 a, v3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
static inline v3 vcross(v3 a, v3 b) {
    return V(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
static inline v3 vnormalize(v3 a) {
    float inv = 1.0f / sqrtf(vdot(a, a));
    return vscale(a, inv);
}
static inline float fmin_cl(float a, float b) { return (b < a) ? b : a; } /* OpenCL min(x,y) = y < x ? y : x */
static inline v3 vmin(v3 a, v3 b) { return V(fmin_cl(a.x, b.x), fmin_cl(a.y, b.y), fmin_cl(a.z, b.z)); }

/* cl:127 — (x - MAX)(x - MIN) <= 0 */
static inline int in_range(float x) { return (x - T_MAX) * (x - T_MIN) <= 0.0f; }

typedef struct {
    const a_material* mat;
    const a_sphere* sph;
    const a_plane* pla;
    const a_lens* len;
    const float* vtx;      /* 4 floats per vertex */
    const float* uv;       /* 2 floats per vertex */
    const uint32_t* idx;
    const a_mesh* mesh;
    const a_model* model;
    uint32_t n_sph, n_pla, n_len, n_model;
} scene_t;

static scene_t scene_from_desc(const orc_scene_desc* d) {
    scene_t s;
    s.mat = (const a_material*)d->materials;
    s.sph = (const a_sphere*)d->spheres;
    s.pla = (const a_plane*)d->planes;
    s.len = (const a_lens*)d->lenses;
    s.vtx = (const float*)d->vertices;
    s.uv = (const float*)d->uvs;
    s.idx = (const uint32_t*)d->indices;
    s.mesh = (const a_mesh*)d->meshes;
    s.model = (const a_model*)d->models;
    s.n_sph = d->n_spheres; s.n_pla = d->n_planes; s.n_len = d->n_lenses; s.n_model = d->n_models;
    return s;
}

/* A candidate hit before shading data is derived. */
typedef struct {
    float t;
    uint32_t id;
    float bu, bv;        /* barycentrics for triangles */
    uint32_t mesh;       /* mesh index for triangles */
    uint32_t model;      /* model index for triangles */
    uint32_t lens_side;  /* 0 -> sphere p1/r1, 1 -> sphere p2/r2 */
} cand_t;

/* cl:149-174 hitSphere: near root if in range, else far root if in range. */
static int sphere_t(const a_sphere* s, v3 o, v3 d, float* t_out) {
    v3 oc = vsub(vld(s->pos), o);
    float b = vdot(oc, d);
    float c = vdot(oc, oc) - s->r * s->r;
    float dis = b * b - c;
    if (dis > 0) {
        float sq = sqrtf(dis);
        float t = b - sq;
        if (in_range(t)) { *t_out = t; return 1; }
        t = b + sq;
        if (in_range(t)) { *t_out = t; return 1; }
    }
    return 0;
}

/* cl:176-194 hitPlane */
static int plane_t(const a_plane* p, v3 o, v3 d, float* t_out) {
    v3 n = vld(p->normal);
    float a = vdot(d, n);
    float b = vdot(vsub(vld(p->pos), o), n);
    float t = b / a;
    if (in_range(t)) { *t_out = t; return 1; }
    return 0;
}

/* cl:196-255 hitLens: intersection of two balls. */
static int lens_t(const a_lens* l, v3 o, v3 d, float* t_out, uint32_t* side) {
    v3 oc = vsub(vld(l->p1), o);
    float b1 = vdot(oc, d);
    float c = vdot(oc, oc) - l->r1 * l->r1;
    float dis1 = b1 * b1 - c;
    oc = vsub(vld(l->p2), o);
    float b2 = vdot(oc, d);
    c = vdot(oc, oc) - l->r2 * l->r2;
    float dis2 = b2 * b2 - c;
    if (dis1 > 0 && dis2 > 0) {
        float d1 = sqrtf(dis1), d2 = sqrtf(dis2);
        float t1A = b1 - d1, t1B = b1 + d1, t2A = b2 - d2, t2B = b2 + d2;
        float t;
        uint32_t sd;
        if ((t1B < t2A) || (t2B < t1A)) return 0;                       /* cl:220 */
        else if (T_MIN <= t1A || T_MIN <= t2A) {                        /* cl:221-231 outside */
            if (t2A <= t1A) { sd = 0; t = t1A; } else { sd = 1; t = t2A; }
        } else if (T_MIN <= t1B && T_MIN <= t2B) {                      /* cl:232-242 inside */
            if (t1B <= t2B) { sd = 0; t = t1B; } else { sd = 1; t = t2B; }
        } else return 0;                                                /* cl:243 */
        if (t <= T_MAX) { *t_out = t; *side = sd; return 1; }           /* cl:245 */
    }
    return 0;
}

static inline v3 mesh_vertex(const scene_t* s, const a_mesh* m, uint32_t i) { /* cl:93-95 */
    return vld(s->vtx + 4 * (size_t)(m->vertex_anchor + s->idx[m->index_anchor + i]));
}

/* cl:257-289 hitTriangle (Moller-Trumbore) + the front-face test of cl:295.
 * Returns 1 when the triangle is hit in range AND faces the ray. */
static int triangle_t(const scene_t* s, const a_mesh* m, uint32_t f, v3 o, v3 d, float* t_out, float* bu, float* bv) {
    v3 A = mesh_vertex(s, m, 3 * f), B = mesh_vertex(s, m, 3 * f + 1), C = mesh_vertex(s, m, 3 * f + 2);
    v3 e1 = vsub(B, A), e2 = vsub(C, A);
    v3 h = vcross(d, e2);
    float a = vdot(e1, h);
    if (a > -TRI_EPS && a < TRI_EPS) return 0;
    float inv = 1.0f / a;
    v3 sv = vsub(o, A);
    float u = inv * vdot(sv, h);
    if (u < 0.0f || u > 1.0f) return 0;
    v3 q = vcross(sv, e1);
    float v = inv * vdot(d, q);
    if (v < 0.0f || u + v > 1.0f) return 0;
    float t = inv * vdot(e2, q);
    if (!in_range(t)) return 0;
    v3 n = vnormalize(vcross(e1, e2));                                  /* cl:285 */
    if (!(vdot(n, d) < 0.0f)) return 0;                                 /* cl:295 */
    *t_out = t; *bu = u; *bv = v;
    return 1;
}

/* cl:322-360 hitScene with cl:305-320 hitModel and cl:291-303 hitMeshOut folded in.
 * Strict `<` everywhere: on equal t the earlier kind / lower index wins. */
static int nearest(const scene_t* s, v3 o, v3 d, uint32_t mesh_mode, cand_t* best) {
    float t;
    best->t = T_MAX; best->id = ORC_ID_MISS; best->bu = best->bv = 0.0f; best->mesh = 0; best->model = 0; best->lens_side = 0;
    for (uint32_t i = 0; i < s->n_sph; i++)
        if (sphere_t(s->sph + i, o, d, &t) && t < best->t) { best->t = t; best->id = (ORC_KIND_SPHERE << 28) | i; }
    for (uint32_t i = 0; i < s->n_pla; i++)
        if (plane_t(s->pla + i, o, d, &t) && t < best->t) { best->t = t; best->id = (ORC_KIND_PLANE << 28) | i; }
    for (uint32_t i = 0; i < s->n_len; i++) {
        uint32_t side;
        if (lens_t(s->len + i, o, d, &t, &side) && t < best->t) {
            best->t = t; best->id = (ORC_KIND_LENS << 28) | i; best->lens_side = side;
        }
    }
    for (uint32_t mi = 0; mi < s->n_model; mi++) {
        const a_model* mo = s->model + mi;
        cand_t mb;
        mb.t = T_MAX; mb.id = ORC_ID_MISS; mb.bu = mb.bv = 0.0f; mb.mesh = 0;
        for (uint32_t k = 0; k < mo->mesh_count; k++) {
            const a_mesh* me = s->mesh + mo->mesh_anchor + k;
            float bu, bv;
            if (mesh_mode == ORC_MESH_FIRST_FRONT) {
                for (uint32_t f = 0; f < me->face_count; f++) {
                    if (triangle_t(s, me, f, o, d, &t, &bu, &bv)) {         /* first front-facing hit wins */
                        if (t < mb.t) { mb.t = t; mb.id = (ORC_KIND_TRIANGLE << 28) | (me->index_anchor / 3 + f);
                                        mb.bu = bu; mb.bv = bv; mb.mesh = mo->mesh_anchor + k; }
                        break;
                    }
                }
            } else {                                                        /* nearest front-facing, ties -> lowest face */
                for (uint32_t f = 0; f < me->face_count; f++)
                    if (triangle_t(s, me, f, o, d, &t, &bu, &bv) && t < mb.t) {
                        mb.t = t; mb.id = (ORC_KIND_TRIANGLE << 28) | (me->index_anchor / 3 + f);
                        mb.bu = bu; mb.bv = bv; mb.mesh = mo->mesh_anchor + k;
                    }
            }
        }
        if (mb.id != ORC_ID_MISS && mb.t < best->t) { *best = mb; best->model = mi; best->lens_side = 0; }
    }
    return best->id != ORC_ID_MISS;
}

/* Shading data of the winning candidate: p, normal, uv, mat_ID, texture_ID exactly as the primitive
 * routines fill HPI (cl:158-161, 185-188, 246-249, 281-285, 296, 314). */
static void fill_hit(const scene_t* s, v3 o, v3 d, const cand_t* c, orc_hit* h) {
    uint32_t kind = c->id >> 28, i = c->id & 0x0FFFFFFFu;
    v3 p = vadd(o, vscale(d, c->t));                                     /* cl:141-143 */
    v3 n = V(0, 0, 0);
    h->t = c->t; h->id = c->id;
    h->uv[0] = h->uv[1] = 0.0f; h->texture_id = 0; h->mat_id = 0;
    if (kind == ORC_KIND_SPHERE) {
        const a_sphere* sp = s->sph + i;
        n = vdivs(vsub(p, vld(sp->pos)), sp->r);
        h->mat_id = sp->mat;
    } else if (kind == ORC_KIND_PLANE) {
        const a_plane* pl = s->pla + i;
        v3 pn = vld(pl->normal);
        float a = vdot(d, pn);
        float sg = (a != a) ? 0.0f : (a > 0.0f ? 1.0f : (a < 0.0f ? -1.0f : a));
        n = vscale(vneg(pn), sg);                                        /* cl:187 */
        h->mat_id = pl->mat;
    } else if (kind == ORC_KIND_LENS) {
        const a_lens* l = s->len + i;
        if (c->lens_side == 0) n = vdivs(vsub(p, vld(l->p1)), l->r1);
        else n = vdivs(vsub(p, vld(l->p2)), l->r2);
        h->mat_id = l->mat;
    } else if (kind == ORC_KIND_TRIANGLE) {
        const a_mesh* me = s->mesh + c->mesh;
        uint32_t f = i - me->index_anchor / 3;
        v3 A = mesh_vertex(s, me, 3 * f), B = mesh_vertex(s, me, 3 * f + 1), C = mesh_vertex(s, me, 3 * f + 2);
        n = vnormalize(vcross(vsub(B, A), vsub(C, A)));
        if (s->uv) {                                                     /* cl:97-103: same anchor + index as positions */
            const float* ua = s->uv + 2 * (size_t)(me->vertex_anchor + s->idx[me->index_anchor + 3 * f]);
            const float* ub = s->uv + 2 * (size_t)(me->vertex_anchor + s->idx[me->index_anchor + 3 * f + 1]);
            const float* uc = s->uv + 2 * (size_t)(me->vertex_anchor + s->idx[me->index_anchor + 3 * f + 2]);
            float w = 1.0f - c->bu - c->bv;
            h->uv[0] = ua[0] * w + ub[0] * c->bu + uc[0] * c->bv;
            h->uv[1] = ua[1] * w + ub[1] * c->bu + uc[1] * c->bv;
        }
        h->texture_id = me->texture_id;
        h->mat_id = s->model[c->model].mat;
    }
    h->p[0] = p.x; h->p[1] = p.y; h->p[2] = p.z;
    h->n[0] = n.x; h->n[1] = n.y; h->n[2] = n.z;
}

/* ------------------------------------------------------------------------------------------- RNG */
/* Table mode, cl:113-125.  dot in f32, "* 438.5453" in double, fabs, truncation to uint. */
static inline uint32_t dir_hash(v3 d) {
    float f = d.x * 123.9898f + d.y * 348.233f + d.z * 433.3314f;
    return (uint32_t)fabs((double)f * 438.5453);
}
static v3 table_vec(const float* tab, v3 d, uint32_t s_seed, uint32_t gx, uint32_t gy) {     /* cl:113-118 */
    uint64_t k = (uint64_t)(uint32_t)(s_seed * 2683u) + (uint64_t)gx * 3931u + (uint64_t)gy * 2504u;
    uint32_t seed = (uint32_t)(((uint64_t)dir_hash(d) + k * 3u) % TABLE_N);
    return V(tab[seed], tab[seed + 1], tab[seed + 2]);
}
static float table_uniform(const float* tab, v3 d, uint32_t s_seed, uint32_t gx, uint32_t gy) { /* cl:120-125 */
    uint64_t k = (uint64_t)(uint32_t)(s_seed * 2683u) + (uint64_t)gx * 3931u + (uint64_t)gy;
    uint32_t seed = TABLE_N * 3u + (uint32_t)(((uint64_t)dir_hash(d) + k) % TABLE_N);
    return tab[seed];
}

/* Counter mode (the product's generator).  Integer hash "lowbias32". */
uint32_t orc_rng_mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
uint32_t orc_rng_bounce_key(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce) {
    uint32_t hp = orc_rng_mix32(seed ^ orc_rng_mix32(pixel ^ 0xA511E9B3u));
    uint32_t hs = orc_rng_mix32(hp ^ (sample * 0x9E3779B9u));
    return orc_rng_mix32(hs + bounce * 0x85EBCA6Bu + 0x1B873593u);
}
static inline uint32_t rng_draw(uint32_t hb, uint32_t j) { return orc_rng_mix32(hb + j * 0xC2B2AE35u); }
static inline float u01(uint32_t h) { return (float)(h >> 8) * 5.9604644775390625e-8f; }          /* 2^-24 */
static inline float u16f(uint32_t x) { return (float)x * 1.52587890625e-5f; }                    /* 16 bits -> [0,1)  */
static inline float s16f(uint32_t x) { return (float)x * 3.0517578125e-5f - 1.0f; }              /* 16 bits -> [-1,1) */

/* Uniform in the unit ball with exactly-rounded operations only: Marsaglia disc rejection for the
 * direction (one 32-bit draw = two 16-bit coordinates per try), max of three 16-bit uniforms for the
 * radius (CDF r^3, same law as cbrt(U), src/raytracer.cpp:79).  r3 is supplied by the caller. */
static v3 rng_ball(uint32_t hb, uint32_t base, uint32_t r3bits) {
    float u = 0.0f, v = 0.0f, s = 0.0f;
    for (uint32_t k = 0; k < 16; k++) {
        uint32_t h = rng_draw(hb, base + k);
        u = s16f(h >> 16);
        v = s16f(h & 0xFFFFu);
        s = u * u + v * v;
        if (s < 1.0f) break;
        u = v = s = 0.0f;
    }
    float q = sqrtf(1.0f - s);
    float dx = (2.0f * u) * q, dy = (2.0f * v) * q, dz = 1.0f - 2.0f * s;
    uint32_t hr = rng_draw(hb, base + 16);
    float r = fmaxf(fmaxf(u16f(hr >> 16), u16f(hr & 0xFFFFu)), u16f(r3bits));
    return V(dx * r, dy * r, dz * r);
}
/* The reference reads table floats [seed, seed+2] instead of [3 seed, 3 seed+2] (cl:109-117), so its
 * "random vector" is (a.x,a.y,a.z), (a.y,a.z,b.x) or (a.z,b.x,b.y) of two adjacent ball vectors with
 * probability 1/3 each.  The counter generator draws from that same law.
 * Draw 0: high half -> selector, low half -> third radius uniform of ball a. */
void orc_rng_scatter_vec(uint32_t hb, float* out) {
    uint32_t d0 = rng_draw(hb, 0);
    uint32_t sel = ((d0 >> 16) * 3u) >> 16;
    v3 a = rng_ball(hb, 1, d0 & 0xFFFFu);
    v3 b = rng_ball(hb, 33, rng_draw(hb, 50) >> 16);
    v3 r = sel == 0 ? a : (sel == 1 ? V(a.y, a.z, b.x) : V(a.z, b.x, b.y));
    out[0] = r.x; out[1] = r.y; out[2] = r.z;
}
float orc_rng_uniform(uint32_t hb) { return u01(rng_draw(hb, 0)); }

/* ------------------------------------------------------------------------------------------- texture */
typedef struct { const float* data; int w, h, layers; } tex_t;

static inline const float* texel(const tex_t* t, int i, int j, int l) {
    if (i < 0) i = 0; if (i > t->w - 1) i = t->w - 1;
    if (j < 0) j = 0; if (j > t->h - 1) j = t->h - 1;
    return t->data + 4 * (((size_t)l * t->h + j) * t->w + i);
}
/* cl:105-107 with sampler cl:15: normalized coords, linear filter (OpenCL 1.2 §8.2), taps clamped. */
static v3 texture_rgb(const tex_t* t, float s, float tt, uint32_t layer) {
    float u = s * (float)t->w, v = tt * (float)t->h;
    float fu = floorf(u - 0.5f), fv = floorf(v - 0.5f);
    float a = (u - 0.5f) - fu, b = (v - 0.5f) - fv;
    int i0 = (int)fu, j0 = (int)fv;
    int l = (int)rintf((float)layer);
    if (l < 0) l = 0; if (l > t->layers - 1) l = t->layers - 1;
    const float *t00 = texel(t, i0, j0, l), *t10 = texel(t, i0 + 1, j0, l);
    const float *t01 = texel(t, i0, j0 + 1, l), *t11 = texel(t, i0 + 1, j0 + 1, l);
    float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return V(w00 * t00[0] + w10 * t10[0] + w01 * t01[0] + w11 * t11[0],
             w00 * t00[1] + w10 * t10[1] + w01 * t01[1] + w11 * t11[1],
             w00 * t00[2] + w10 * t10[2] + w01 * t01[2] + w11 * t11[2]);
}

/* ------------------------------------------------------------------------------------------- materials */
typedef struct {
    const scene_t* sc;
    const tex_t* tex;
    const float* table;
    const orc_params* prm;
    uint32_t gx, gy, W;
} ctx_t;

/* cl:362-367 rayReflect (the attenuation of cl:366 is applied by the caller, which knows the type) */
static v3 reflect_dir(v3 d, v3 n) {
    float k = 2.0f * vdot(d, n);
    return vnormalize(vsub(d, vscale(n, k)));
}

/* One iteration of the loop body of getCol (cl:447-483) after a hit.  Returns 1 when the path ends. */
static int shade(const ctx_t* c, const orc_hit* h, v3* o, v3* d, v3* out, uint32_t sample, uint32_t bounce) {
    const a_material* m = c->sc->mat + h->mat_id;
    v3 n = vld(h->n), p = vld(h->p);
    v3 color = vld(m->color);
    uint32_t s_seed = bounce + sample;                                   /* cl:456,472,476 */
    uint32_t hb = 0;
    if (c->prm->rng_mode != ORC_RNG_TABLE)
        hb = orc_rng_bounce_key(c->prm->seed, c->gy * c->W + c->gx, sample, bounce);
    /* ORC_RNG_COUNTER_TABLE: the reference's own table (so its sampling law by construction, misaligned read
     * included, cl:109-117), entered at a counter-keyed uniform index instead of the direction hash. */
    uint32_t tix = (uint32_t)(((uint64_t)hb * TABLE_N) >> 32);
    switch (m->type) {
    case M_DIFFUSE:
    case M_TEXTURED: {                                                   /* cl:393-399 rayScatter */
        v3 rv;
        if (c->prm->rng_mode == ORC_RNG_TABLE) rv = table_vec(c->table, *d, s_seed, c->gx, c->gy);
        else if (c->prm->rng_mode == ORC_RNG_COUNTER_TABLE) rv = V(c->table[tix], c->table[tix + 1], c->table[tix + 2]);
        else { float r3[3]; orc_rng_scatter_vec(hb, r3); rv = vld(r3); }
        *d = vnormalize(vadd(n, rv));
        *o = p;
        *out = vscale(*out, m->extra);
        if (m->type == M_TEXTURED) color = texture_rgb(c->tex, h->uv[0], h->uv[1], h->texture_id);  /* cl:477 */
        *out = vmin(*out, color);
        return 0;
    }
    case M_LIGHT:                                                        /* cl:459-462 */
        *out = vmin(*out, color);
        return 1;
    case M_REFLECTIVE:                                                   /* cl:463-466 */
        *o = p;
        *d = reflect_dir(*d, n);
        *out = vscale(*out, m->extra);
        *out = vmin(*out, color);
        return 0;
    case M_REFRACTIVE:
    case M_DIELECTRIC: {                                                 /* cl:369-391, cl:407-435 */
        v3 nn; float ratio;
        float cai = vdot(*d, n);
        if (cai > 0) { nn = vneg(n); ratio = m->extra; cai = -cai; }
        else { nn = n; ratio = 1.0f / m->extra; }
        int try_refract = 1;
        if (m->type == M_DIELECTRIC) {
            float cosi = -cai;                                           /* cl:401-405 schlick(-cai, ratio) */
            float r0 = (1.0f - ratio) / (1.0f + ratio);
            r0 *= r0;
            float prob = r0 + (1.0f - r0) * powf(1.0f - cosi, 5.0f);
            float u = (c->prm->rng_mode == ORC_RNG_TABLE) ? table_uniform(c->table, *d, s_seed, c->gx, c->gy)
                    : (c->prm->rng_mode == ORC_RNG_COUNTER_TABLE) ? c->table[3u * TABLE_N + tix] : orc_rng_uniform(hb);
            try_refract = prob < u;                                      /* cl:423 */
        }
        float disc = 1.0f - ratio * ratio * (1.0f - cai * cai);
        if (try_refract && disc > 0.0f) {
            *o = p;
            float k = ratio * cai + sqrtf(disc);
            *d = vsub(vscale(*d, ratio), vscale(nn, k));                 /* not normalised, cl:386/428 */
        } else {
            *o = p;
            *d = reflect_dir(*d, nn);
        }
        *out = vmin(*out, color);
        return 0;
    }
    default:
        return 0;                                                        /* no case: ray unchanged */
    }
}

/* cl:444-486 getCol */
static v3 path_radiance(const ctx_t* c, v3 o, v3 d, uint32_t sample, uint64_t* rays) {
    v3 out = V(1.0f, 1.0f, 1.0f);
    for (uint32_t i = 0; i < c->prm->max_depth; i++) {
        cand_t cd;
        orc_hit h;
        (*rays)++;
        if (!nearest(c->sc, o, d, c->prm->mesh_mode, &cd)) { out = V(0.0f, 0.0f, 0.0f); break; }
        fill_hit(c->sc, o, d, &cd, &h);
        if (shade(c, &h, &o, &d, &out, sample, i)) break;
    }
    return out;
}

/* cl:129-139 genInitRay with s = x/W, t = y/H (cl:500-501) */
static v3 primary_dir(const float* cam, int x, int y, int W, int H) {
    float s = (float)x / (float)W, t = (float)y / (float)H;
    v3 llc = vld(cam + 3), hor = vld(cam + 6), ver = vld(cam + 9);
    return vnormalize(vadd(vadd(llc, vscale(hor, s)), vscale(ver, t)));
}

typedef struct {
    ctx_t base;
    scene_t sc;
    tex_t tex;
    const float* cam;
    float* image;
    int W, H;
    int x0, y0, w, h;          /* window of the W x H frame that is rendered; image is w x h */
    uint32_t first, n;
    int next_row;
    uint64_t rays;
    pthread_mutex_t mu;
} job_t;

static void* worker(void* arg) {
    job_t* j = (job_t*)arg;
    uint64_t rays = 0;
    for (;;) {
        pthread_mutex_lock(&j->mu);
        int y0 = j->next_row; j->next_row += 2;
        pthread_mutex_unlock(&j->mu);
        if (y0 >= j->y0 + j->h) break;
        int y1 = y0 + 2 < j->y0 + j->h ? y0 + 2 : j->y0 + j->h;
        for (int y = y0; y < y1; y++)
            for (int x = j->x0; x < j->x0 + j->w; x++) {
                ctx_t c = j->base;
                c.sc = &j->sc; c.tex = &j->tex; c.gx = (uint32_t)x; c.gy = (uint32_t)y; c.W = (uint32_t)j->W;
                v3 o = vld(j->cam), d0 = primary_dir(j->cam, x, y, j->W, j->H);
                float* px = j->image + 4 * ((size_t)(y - j->y0) * j->w + (x - j->x0));
                for (uint32_t k = j->first; k < j->first + j->n; k++) {
                    v3 col = path_radiance(&c, o, d0, k, &rays);
                    if (c.prm->accum_mode == ORC_ACCUM_SUM) {
                        px[0] += col.x; px[1] += col.y; px[2] += col.z; px[3] += 1.0f;
                    } else if (k == 0) {                                 /* cl:509 */
                        px[0] = sqrtf(col.x); px[1] = sqrtf(col.y); px[2] = sqrtf(col.z); px[3] = 1.0f;
                    } else {                                             /* cl:524-531 */
                        float a = (float)k / (float)(k + 1);
                        float pr = px[0] * px[0], pg = px[1] * px[1], pb = px[2] * px[2];
                        px[0] = sqrtf(col.x + (pr - col.x) * a);
                        px[1] = sqrtf(col.y + (pg - col.y) * a);
                        px[2] = sqrtf(col.z + (pb - col.z) * a);
                        px[3] = 1.0f;
                    }
                }
            }
    }
    pthread_mutex_lock(&j->mu);
    j->rays += rays;
    pthread_mutex_unlock(&j->mu);
    return 0;
}

double orc_render(const orc_scene_desc* scene, const float* tex, int tw, int th, int tl, const float* cam12,
                  const float* table, float* image, int W, int H, uint32_t first, uint32_t n,
                  const orc_params* prm, uint64_t* rays) {
    return orc_render_window(scene, tex, tw, th, tl, cam12, table, image, W, H, 0, 0, W, H, first, n, prm, rays);
}

/* Work-items (x, y) in [x0, x0+w) x [y0, y0+h) of the W x H NDRange only (same s = x/W, t = y/H and the same
 * per-pixel RNG keys as in the full frame); image is the w x h window, row 0 = y0. */
double orc_render_window(const orc_scene_desc* scene, const float* tex, int tw, int th, int tl, const float* cam12,
                         const float* table, float* image, int W, int H, int x0, int y0, int w, int h,
                         uint32_t first, uint32_t n, const orc_params* prm, uint64_t* rays) {
    job_t j;
    memset(&j, 0, sizeof j);
    j.sc = scene_from_desc(scene);
    j.tex.data = tex; j.tex.w = tw; j.tex.h = th; j.tex.layers = tl;
    j.base.table = table; j.base.prm = prm;
    j.cam = cam12; j.image = image; j.W = W; j.H = H; j.first = first; j.n = n;
    j.x0 = x0; j.y0 = y0; j.w = w; j.h = h; j.next_row = y0;
    pthread_mutex_init(&j.mu, 0);
    int nt = prm->threads < 1 ? 1 : (prm->threads > 256 ? 256 : prm->threads);
    pthread_t th_[256];
    struct timespec t0, t1;
    clock_gettime(CLOCK_MONOTONIC, &t0);
    for (int i = 1; i < nt; i++) pthread_create(&th_[i], 0, worker, &j);
    worker(&j);
    for (int i = 1; i < nt; i++) pthread_join(th_[i], 0);
    clock_gettime(CLOCK_MONOTONIC, &t1);
    pthread_mutex_destroy(&j.mu);
    if (rays) *rays = j.rays;
    return (double)(t1.tv_sec - t0.tv_sec) + 1e-9 * (double)(t1.tv_nsec - t0.tv_nsec);
}

int orc_intersect(const orc_scene_desc* scene, const float* origin, const float* dir, uint32_t mesh_mode, orc_hit* out) {
    scene_t s = scene_from_desc(scene);
    cand_t c;
    v3 o = vld(origin), d = vld(dir);
    memset(out, 0, sizeof *out);
    if (!nearest(&s, o, d, mesh_mode, &c)) { out->t = T_MAX; out->id = ORC_ID_MISS; return 0; }
    fill_hit(&s, o, d, &c, out);
    return 1;
}

void orc_primary_ids(const orc_scene_desc* scene, const float* cam12, int W, int H, uint32_t mesh_mode,
                     uint32_t* ids, float* ts) {
    scene_t s = scene_from_desc(scene);
    v3 o = vld(cam12);
    for (int y = 0; y < H; y++)
        for (int x = 0; x < W; x++) {
            cand_t c;
            nearest(&s, o, primary_dir(cam12, x, y, W, H), mesh_mode, &c);
            ids[(size_t)y * W + x] = c.id;
            ts[(size_t)y * W + x] = c.t;
        }
}

int orc_bounce(const orc_scene_desc* scene, const float* tex, int tw, int th, int tl, const float* table,
               const orc_params* prm, float* ray6, uint32_t sample, uint32_t bounce, uint32_t gx, uint32_t gy,
               uint32_t W, float* col3) {
    scene_t s = scene_from_desc(scene);
    tex_t t = {tex, tw, th, tl};
    ctx_t c = {&s, &t, table, prm, gx, gy, W};
    v3 o = vld(ray6), d = vld(ray6 + 3), out = V(1.0f, 1.0f, 1.0f);
    cand_t cd;
    orc_hit h;
    int ended;
    if (!nearest(&s, o, d, prm->mesh_mode, &cd)) { out = V(0.0f, 0.0f, 0.0f); ended = 1; }
    else { fill_hit(&s, o, d, &cd, &h); ended = shade(&c, &h, &o, &d, &out, sample, bounce); }
    ray6[0] = o.x; ray6[1] = o.y; ray6[2] = o.z; ray6[3] = d.x; ray6[4] = d.y; ray6[5] = d.z;
    col3[0] = out.x; col3[1] = out.y; col3[2] = out.z;
    return ended;
}

/* ------------------------------------------------------------------------------------------- random table
 * src/raytracer.cpp:67-93.  The reference draws from std::mt19937 through std::normal_distribution<float>
 * and std::uniform_real_distribution<float>; libstdc++ 13's algorithms are restated here (Marsaglia polar
 * method with one saved deviate; generate_canonical<float,24> = one 32-bit draw / 2^32, clamped below 1). */
typedef struct { uint32_t mt[624]; int idx; int have_saved; float saved; } mt_t;

static void mt_seed(mt_t* m, uint32_t seed) {
    m->mt[0] = seed;
    for (int i = 1; i < 624; i++) m->mt[i] = 1812433253u * (m->mt[i - 1] ^ (m->mt[i - 1] >> 30)) + (uint32_t)i;
    m->idx = 624; m->have_saved = 0; m->saved = 0.0f;
}
static uint32_t mt_next(mt_t* m) {
    if (m->idx >= 624) {
        for (int i = 0; i < 624; i++) {
            uint32_t y = (m->mt[i] & 0x80000000u) | (m->mt[(i + 1) % 624] & 0x7fffffffu);
            m->mt[i] = m->mt[(i + 397) % 624] ^ (y >> 1) ^ ((y & 1u) ? 0x9908b0dfu : 0u);
        }
        m->idx = 0;
    }
    uint32_t y = m->mt[m->idx++];
    y ^= y >> 11; y ^= (y << 7) & 0x9d2c5680u; y ^= (y << 15) & 0xefc60000u; y ^= y >> 18;
    return y;
}
static float mt_canonical(mt_t* m) {
    float r = (float)mt_next(m) / 4294967296.0f;
    if (r >= 1.0f) r = nextafterf(1.0f, 0.0f);
    return r;
}
static float mt_uniform01(mt_t* m) { return mt_canonical(m) * 1.0f + 0.0f; }
static float mt_normal(mt_t* m) {
    float ret;
    if (m->have_saved) { m->have_saved = 0; ret = m->saved; }
    else {
        float x, y, r2;
        do {
            x = 2.0f * mt_canonical(m) - 1.0f;
            y = 2.0f * mt_canonical(m) - 1.0f;
            r2 = x * x + y * y;
        } while (r2 > 1.0f || r2 == 0.0f);
        float mult = sqrtf(-2 * logf(r2) / r2);
        m->saved = x * mult; m->have_saved = 1;
        ret = y * mult;
    }
    return ret * 1.0f + 0.0f;
}
void orc_make_table(uint32_t seed, float* out) {
    mt_t m;
    mt_seed(&m, seed);
    for (uint32_t i = 0; i < TABLE_N; i++) {
        float x = mt_normal(&m), y = mt_normal(&m), z = mt_normal(&m);
        float r = cbrtf(mt_uniform01(&m)), u = mt_uniform01(&m);
        float k = r / sqrtf(x * x + y * y + z * z);
        out[3 * i] = x * k; out[3 * i + 1] = y * k; out[3 * i + 2] = z * k;
        out[3 * TABLE_N + i] = u;
    }
}
