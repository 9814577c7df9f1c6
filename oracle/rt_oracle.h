/* oracle/rt_oracle.h — TEST INFRASTRUCTURE ONLY (never linked into or imported by the product).
 *
 * Plain-C restatement ("port") of the reference's per-pixel path-tracing path,
 * /root/reference/kernels/raytracer.cl:93-532, written independently of the CUDA product and pinned
 * bit-for-bit against the reference's own kernel text compiled as C++ (oracle/_ref, see
 * tests/test_oracle_pinning.py).  Compiled with -O2 -ffp-contract=off: every float operation is a single
 * IEEE-754 binary32 operation in source order.
 *
 * Extensions the reference does not have (each is OFF in the reference-exact configuration):
 *   - runtime max depth                 (reference: #define DEPTH 30, kernels/raytracer.cl:5)
 *   - counter-based RNG  (ORC_RNG_COUNTER)  — the product's own generator, specified in DESIGN.md §RNG
 *   - linear-sum accumulation (ORC_ACCUM_SUM) — what the multi-GPU sample split reduces
 *   - nearest-front-facing mesh rule (ORC_MESH_NEAREST) — the brute-force definition a BVH must equal
 */
#ifndef RT_ORACLE_H
#define RT_ORACLE_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Same field order as ref_scene_desc (oracle/ref_harness.cpp) and rt_scene_desc (include/rt_capi.h):
 * pointers to the reference's host AoS layouts, include/scene.h:32-81. */
typedef struct orc_scene_desc {
    const void *materials, *spheres, *planes, *lenses, *vertices, *uvs, *indices, *meshes, *models;
    uint32_t n_materials, n_spheres, n_planes, n_lenses, n_vertices, n_uvs, n_indices, n_meshes, n_models;
} orc_scene_desc;

enum { ORC_RNG_TABLE = 0, ORC_RNG_COUNTER = 1, ORC_RNG_COUNTER_TABLE = 2 };
enum { ORC_ACCUM_REFERENCE = 0, ORC_ACCUM_SUM = 1 };
enum { ORC_MESH_FIRST_FRONT = 0, ORC_MESH_NEAREST = 1 };

#define ORC_ID_MISS 0xFFFFFFFFu
#define ORC_KIND_SPHERE 1u
#define ORC_KIND_PLANE 2u
#define ORC_KIND_LENS 3u
#define ORC_KIND_TRIANGLE 4u

typedef struct orc_params {
    uint32_t max_depth;  /* reference: 30 */
    uint32_t rng_mode;   /* ORC_RNG_* */
    uint32_t seed;       /* counter RNG key */
    uint32_t accum_mode; /* ORC_ACCUM_* */
    uint32_t mesh_mode;  /* ORC_MESH_* */
    int32_t threads;     /* host threads */
} orc_params;

typedef struct orc_hit {
    float t;
    uint32_t id;          /* kind << 28 | index ; ORC_ID_MISS */
    float p[3], n[3], uv[2];
    uint32_t mat_id, texture_id;
} orc_hit;

/* Samples [first, first+n): image is W*H RGBA f32, row 0 = t 0.
 * ORC_ACCUM_REFERENCE: image holds (sqrt(running mean), 1) exactly as trace/retrace leave it.
 * ORC_ACCUM_SUM:       image.rgb += linear radiance of each sample, image.a += 1 per sample.
 * Returns wall seconds; *rays = number of nearest-hit queries (path segments). */
double orc_render(const orc_scene_desc* scene, const float* tex, int tw, int th, int tl, const float* cam12,
                  const float* table, float* image, int W, int H, uint32_t first, uint32_t n,
                  const orc_params* prm, uint64_t* rays);

/* The same for the work-items of a window [x0, x0+w) x [y0, y0+h) of the W x H frame; image is w x h. */
double orc_render_window(const orc_scene_desc* scene, const float* tex, int tw, int th, int tl, const float* cam12,
                         const float* table, float* image, int W, int H, int x0, int y0, int w, int h,
                         uint32_t first, uint32_t n, const orc_params* prm, uint64_t* rays);

int orc_intersect(const orc_scene_desc* scene, const float* origin, const float* dir, uint32_t mesh_mode,
                  orc_hit* out);

void orc_primary_ids(const orc_scene_desc* scene, const float* cam12, int W, int H, uint32_t mesh_mode,
                     uint32_t* ids, float* ts);

/* One bounce of the path loop as work-item (gx,gy) of a W-wide image: ray6 (origin, dir) updated in place,
 * col3 = throughput after the bounce starting from (1,1,1); s_seed = bounce index + sample (table mode);
 * counter mode uses (sample, bounce) separately.  Returns 0 = path continues, 1 = terminated. */
int orc_bounce(const orc_scene_desc* scene, const float* tex, int tw, int th, int tl, const float* table,
               const orc_params* prm, float* ray6, uint32_t sample, uint32_t bounce, uint32_t gx, uint32_t gy,
               uint32_t W, float* col3);

/* Counter RNG primitives, exposed for known-answer tests. */
uint32_t orc_rng_mix32(uint32_t x);
uint32_t orc_rng_bounce_key(uint32_t seed, uint32_t pixel, uint32_t sample, uint32_t bounce);
void orc_rng_scatter_vec(uint32_t hb, float* v3);
float orc_rng_uniform(uint32_t hb);

/* src/raytracer.cpp:67-93 with libstdc++'s mt19937 / normal_distribution / uniform_real_distribution restated. */
void orc_make_table(uint32_t seed, float* out400000);

#ifdef __cplusplus
}
#endif
#endif
