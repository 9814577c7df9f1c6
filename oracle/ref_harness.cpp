// oracle/ref_harness.cpp — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// Host harness around the reference's own kernel text, compiled as C++ (see cl_shim.h and
// oracle/Makefile).  `raytracer_cl.inc` is generated at build time from
// /root/reference/kernels/raytracer.cl by a 4-rule sed and lives in a temp directory; only the
// resulting shared objects are written to oracle/_ref/.
//
// The harness restates the *host* steps of the reference that surround the kernel:
//   - random table generation            /root/reference/src/raytracer.cpp:67-93  (fixed seed instead of random_device)
//   - one NDRange launch per sample      /root/reference/src/raytracer.cpp:127-165 (trace for sample 0, retrace for k>=1)
// and adds instrumentation the reference lacks: a ray counter hooked into hitScene, a primary-hit
// ID dump (re-walking hitScene's loops with the reference's own primitive tests), a single-ray probe
// and a single-bounce probe (getCol with DEPTH = 1).
#include "cl_shim.h"

#include <atomic>
#include <chrono>
#include <cstring>
#include <random>
#include <thread>
#include <vector>

thread_local size_t cl_shim_gid[2] = {0, 0};
thread_local uint cl_shim_depth = 30;
thread_local uint64_t ref_ray_counter = 0;
#define REF_HOOK_RAY() (++ref_ray_counter)

#include "raytracer_cl.inc"   // the reference kernel, generated at build time

static_assert(sizeof(Material) == 48 && sizeof(Sphere) == 32 && sizeof(Plane) == 48 && sizeof(Lens) == 64 &&
              sizeof(Mesh) == 16 && sizeof(Model) == 12, "host AoS layouts (include/scene.h:30-81)");

extern "C" {

// Same field order as rt_scene_desc in include/rt_capi.h (declared separately on purpose).
struct ref_scene_desc {
    const void* materials; const void* spheres; const void* planes; const void* lenses;
    const void* vertices; const void* uvs; const void* indices; const void* meshes; const void* models;
    uint32_t n_materials, n_spheres, n_planes, n_lenses, n_vertices, n_uvs, n_indices, n_meshes, n_models;
};

}  // extern "C"

static Scene make_scene(const ref_scene_desc* d) {
    Scene s;
    s.materials = (const Material*)d->materials;
    s.spheres = (const Sphere*)d->spheres;
    s.planes = (const Plane*)d->planes;
    s.lenses = (const Lens*)d->lenses;
    s.vertex_buffer = (const vec3*)d->vertices;
    s.texture_uv_buffer = (const vec2*)d->uvs;
    s.index_buffer = (const uint*)d->indices;
    s.mesh_buffer = (const Mesh*)d->meshes;
    s.models = (const Model*)d->models;
    s.sphere_count = d->n_spheres;
    s.plane_count = d->n_planes;
    s.lens_count = d->n_lenses;
    s.model_count = d->n_models;
    return s;
}

template <class F>
static void parallel_rows(int H, int nthreads, F&& body) {
    if (nthreads < 1) nthreads = 1;
    std::atomic<int> next(0);
    auto worker = [&]() {
        for (;;) {
            int y0 = next.fetch_add(4);
            if (y0 >= H) break;
            int y1 = y0 + 4 < H ? y0 + 4 : H;
            for (int y = y0; y < y1; ++y) body(y);
        }
    };
    std::vector<std::thread> pool;
    for (int t = 1; t < nthreads; ++t) pool.emplace_back(worker);
    worker();
    for (auto& t : pool) t.join();
}

extern "C" {

// src/raytracer.cpp:67-93 with a caller-supplied seed in place of std::random_device.
void ref_make_table(uint32_t seed, float* out /* 400000 floats */) {
    const int N = RANDOM_BUFFER_SIZE;
    std::mt19937 rng(seed);
    std::uniform_real_distribution<float> dis(0.0f, 1.0f);
    std::normal_distribution<float> ndis(0.0f, 1.0f);
    for (int i = 0; i < N; i++) {
        float x = ndis(rng), y = ndis(rng), z = ndis(rng), r = std::cbrt(dis(rng)), u = dis(rng);
        float len_inv_r = r / std::sqrt(x * x + y * y + z * z);
        out[3 * i] = x * len_inv_r;
        out[3 * i + 1] = y * len_inv_r;
        out[3 * i + 2] = z * len_inv_r;
        out[3 * N + i] = u;
    }
}

// Samples [first, first+n) of the progressive loop: sample 0 -> `trace`, k >= 1 -> `retrace`
// (src/raytracer.cpp:127-165).  `image` is the RGBA32F texture, W*H*4 floats, row 0 = t 0.
// Returns wall seconds of the sample loop; *rays receives the number of hitScene calls.
double ref_render(const ref_scene_desc* desc, const float* tex, int tw, int th, int tl,
                  const float* cam12, const float* table, float* image, int W, int H,
                  uint32_t first, uint32_t n, uint32_t depth, int nthreads, uint64_t* rays) {
    Scene scene = make_scene(desc);
    image2d_t img{image, W, H};
    image2d_array_t texture{tex, tw, th, tl};
    std::atomic<uint64_t> total(0);
    auto t0 = std::chrono::steady_clock::now();
    for (uint32_t k = first; k < first + n; ++k) {
        parallel_rows(H, nthreads, [&](int y) {
            cl_shim_depth = depth;
            uint64_t before = ref_ray_counter;
            for (int x = 0; x < W; ++x) {
                cl_shim_gid[0] = (size_t)x;
                cl_shim_gid[1] = (size_t)y;
                if (k == 0) trace(img, cam12, table, &scene, texture);
                else retrace(img, img, cam12, table, &scene, texture, k);
            }
            total.fetch_add(ref_ray_counter - before, std::memory_order_relaxed);
        });
    }
    auto t1 = std::chrono::steady_clock::now();
    if (rays) *rays = total.load();
    return std::chrono::duration<double>(t1 - t0).count();
}

// One hitScene call (kernels/raytracer.cl:322-360) on an arbitrary ray.
// out9 = {t, p.xyz, normal.xyz, uv.xy}; ids = {mat_ID, texture_ID}. Returns 1 on hit.
int ref_hit(const ref_scene_desc* desc, const float* origin, const float* dir, float* out9, uint32_t* ids) {
    Scene scene = make_scene(desc);
    Ray r;
    r.origin = vec3(origin[0], origin[1], origin[2]);
    r.dir = vec3(dir[0], dir[1], dir[2]);
    HPI hpi;
    std::memset(&hpi, 0, sizeof hpi);
    bool hit = hitScene(&r, &scene, &hpi);
    if (!hit) return 0;
    out9[0] = hpi.t;
    out9[1] = hpi.p.x; out9[2] = hpi.p.y; out9[3] = hpi.p.z;
    out9[4] = hpi.normal.x; out9[5] = hpi.normal.y; out9[6] = hpi.normal.z;
    out9[7] = hpi.uv.x; out9[8] = hpi.uv.y;
    ids[0] = hpi.mat_ID; ids[1] = hpi.texture_ID;
    return 1;
}

// Which primitive does hitScene pick?  Re-walks the loops of hitScene/hitModel/hitMeshOut
// (kernels/raytracer.cl:291-360) calling the reference's own primitive tests, and records the winner.
// id = kind<<28 | index, kind 1 sphere, 2 plane, 3 lens, 4 triangle (global face = index_anchor/3 + face);
// miss = 0xFFFFFFFF.
static uint32_t identify(const Ray* r, const Scene* scene, float* t_out) {
    uint32_t best = 0xFFFFFFFFu;
    float hit_min = MAX_DISTANCE;
    HPI h;
    for (uint i = 0; i < scene->sphere_count; i++)
        if (hitSphere(r, scene->spheres + i, &h) && h.t < hit_min) { hit_min = h.t; best = (1u << 28) | i; }
    for (uint i = 0; i < scene->plane_count; i++)
        if (hitPlane(r, scene->planes + i, &h) && h.t < hit_min) { hit_min = h.t; best = (2u << 28) | i; }
    for (uint i = 0; i < scene->lens_count; i++)
        if (hitLens(r, scene->lenses + i, &h) && h.t < hit_min) { hit_min = h.t; best = (3u << 28) | i; }
    for (uint m = 0; m < scene->model_count; m++) {
        const Model* model = scene->models + m;
        float model_min = MAX_DISTANCE;
        uint32_t model_best = 0xFFFFFFFFu;
        for (uint k = 0; k < model->mesh_count; k++) {
            const Mesh* mesh = scene->mesh_buffer + model->mesh_anchor + k;
            for (uint f = 0; f < mesh->face_count; f++) {
                if (hitTriangle(r, scene, mesh, 3 * f, 3 * f + 1, 3 * f + 2, &h) && dot(h.normal, r->dir) < 0.0f) {
                    if (h.t < model_min) { model_min = h.t; model_best = (4u << 28) | (mesh->index_anchor / 3 + f); }
                    break;  // first front-facing hit of this mesh wins (hitMeshOut)
                }
            }
        }
        if (model_best != 0xFFFFFFFFu && model_min < hit_min) { hit_min = model_min; best = model_best; }
    }
    *t_out = hit_min;
    return best;
}

// Primary-hit id and t for every pixel (the primary ray is sample independent: kernels/raytracer.cl:500-505).
// Returns the number of pixels where the re-walk disagrees with hitScene's own t (must be 0).
int ref_primary_ids(const ref_scene_desc* desc, const float* cam12, int W, int H, uint32_t* ids, float* ts) {
    Scene scene = make_scene(desc);
    int bad = 0;
    for (int y = 0; y < H; ++y)
        for (int x = 0; x < W; ++x) {
            float s = (float)x / (float)W, t = (float)y / (float)H;
            vec3 pos = getVec(cam12, 0);
            Ray r = genInitRay(cam12, &pos, s, t);
            float tt;
            uint32_t id = identify(&r, &scene, &tt);
            HPI hpi;
            bool hit = hitScene(&r, &scene, &hpi);
            if (hit != (id != 0xFFFFFFFFu) || (hit && hpi.t != tt)) bad++;
            ids[(size_t)y * W + x] = id;
            ts[(size_t)y * W + x] = hit ? hpi.t : MAX_DISTANCE;
        }
    return bad;
}

// identify() for an arbitrary ray.
uint32_t ref_identify(const ref_scene_desc* desc, const float* origin, const float* dir, float* t_out) {
    Scene scene = make_scene(desc);
    Ray r;
    r.origin = vec3(origin[0], origin[1], origin[2]);
    r.dir = vec3(dir[0], dir[1], dir[2]);
    return identify(&r, &scene, t_out);
}

// One bounce: getCol with DEPTH = 1 (kernels/raytracer.cl:444-486) on an arbitrary ray, as work-item
// (gx, gy) with seed s_seed (= bounce index + sample).  ray6 = origin.xyz, dir.xyz is updated in place;
// col3 receives the throughput after the bounce (starts from (1,1,1)).
void ref_bounce(const ref_scene_desc* desc, const float* tex, int tw, int th, int tl, const float* table,
                float* ray6, uint32_t s_seed, uint32_t gx, uint32_t gy, float* col3) {
    Scene scene = make_scene(desc);
    image2d_array_t texture{tex, tw, th, tl};
    cl_shim_depth = 1;
    cl_shim_gid[0] = gx;
    cl_shim_gid[1] = gy;
    Ray r;
    r.origin = vec3(ray6[0], ray6[1], ray6[2]);
    r.dir = vec3(ray6[3], ray6[4], ray6[5]);
    col c = getCol(&r, table, &scene, texture, s_seed);
    ray6[0] = r.origin.x; ray6[1] = r.origin.y; ray6[2] = r.origin.z;
    ray6[3] = r.dir.x; ray6[4] = r.dir.y; ray6[5] = r.dir.z;
    col3[0] = c.x; col3[1] = c.y; col3[2] = c.z;
}

const char* ref_build_info(void) {
#ifdef REF_BUILD_INFO
    return REF_BUILD_INFO;
#else
    return "unknown";
#endif
}

}  // extern "C"
