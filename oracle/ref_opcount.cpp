// oracle/ref_opcount.cpp — TEST / MEASUREMENT INFRASTRUCTURE ONLY (never linked into the product).
//
// The reference's kernel text compiled as C++ (as in ref_harness.cpp) with `float` replaced by a class that counts
// every floating-point operation it performs.  This is how SURVEY.md §8(d) defines the algorithmic work of a path
// segment ("op-counting float in the oracle probe": FP32 issue slots per ray = add/sub + mul + compare + div + sqrt +
// pow, minus the mul->add pairs an FMA fuses); tools/opcount.py runs it for the BASELINE scenes, bench.py quotes the
// results as the roofline numerator.
#include <cmath>
#include <cstddef>
#include <cstdint>
#include <cstring>
#include <type_traits>

typedef float real_f;   // the machine float, spelled before `float` is redefined below

struct OpCounts { uint64_t add, mul, cmp, div, sqrt, pow, dmul, fusable; };
static thread_local OpCounts g_ops;

// A float that counts.  An add/sub that consumes the unconsumed result of a multiplication is an FMA on the device (one
// issue slot for the pair); "unconsumed product" is carried in the lowest mantissa bit of the value itself, so that
// cfloat stays 4 bytes and every struct and buffer of the kernel keeps its layout.  Values are therefore off by up to
// one ulp per operation — irrelevant for operation COUNTS averaged over a frame (only knife-edge branches can flip) —
// and the harness clears that bit in all inputs.
static inline real_f cf_tag(real_f v, bool fresh) {
    if (!(v == v) || v == 0.0f || std::isinf(v)) return v;
    uint32_t u;
    std::memcpy(&u, &v, 4);
    u = (u & ~1u) | (fresh ? 1u : 0u);
    std::memcpy(&v, &u, 4);
    return v;
}
static inline bool cf_fresh(real_f v) {
    if (!(v == v) || v == 0.0f || std::isinf(v)) return false;
    uint32_t u;
    std::memcpy(&u, &v, 4);
    return (u & 1u) != 0u;
}
struct cfloat {
    real_f v;
    cfloat() : v(0.0f) {}
    template <class T, class = typename std::enable_if<std::is_arithmetic<T>::value>::type>
    cfloat(T x) : v(cf_tag((real_f)x, false)) {}
    explicit operator double() const { return (double)v; }
    explicit operator int() const { return (int)v; }
    explicit operator unsigned() const { return (unsigned)v; }
    cfloat& operator+=(cfloat b);
    cfloat& operator-=(cfloat b);
    cfloat& operator*=(cfloat b);
    cfloat& operator/=(cfloat b);
};
static inline cfloat mk_res(real_f v, bool fresh) { cfloat r; r.v = cf_tag(v, fresh); return r; }
#define CF_FUSE(a, b) do { if (cf_fresh((a).v) || cf_fresh((b).v)) ++g_ops.fusable; } while (0)
static inline cfloat operator+(cfloat a, cfloat b) { ++g_ops.add; CF_FUSE(a, b); return mk_res(a.v + b.v, false); }
static inline cfloat operator-(cfloat a, cfloat b) { ++g_ops.add; CF_FUSE(a, b); return mk_res(a.v - b.v, false); }
static inline cfloat operator*(cfloat a, cfloat b) { ++g_ops.mul; return mk_res(a.v * b.v, true); }
static inline cfloat operator/(cfloat a, cfloat b) { ++g_ops.div; return mk_res(a.v / b.v, false); }
static inline cfloat operator-(cfloat a) { cfloat r; r.v = -a.v; return r; }    // a sign flip is an operand modifier (keeps the tag)
inline cfloat& cfloat::operator+=(cfloat b) { *this = *this + b; return *this; }
inline cfloat& cfloat::operator-=(cfloat b) { *this = *this - b; return *this; }
inline cfloat& cfloat::operator*=(cfloat b) { *this = *this * b; return *this; }
inline cfloat& cfloat::operator/=(cfloat b) { *this = *this / b; return *this; }
// the reference's one double-precision multiply (kernels/raytracer.cl:114,121: float * double literal)
template <class T, class = typename std::enable_if<std::is_same<T, double>::value>::type>
static inline double operator*(cfloat a, T b) { ++g_ops.dmul; return (double)a.v * b; }
#define CF_CMP(op) static inline bool operator op(cfloat a, cfloat b) { ++g_ops.cmp; return a.v op b.v; }
CF_CMP(<) CF_CMP(>) CF_CMP(<=) CF_CMP(>=) CF_CMP(==) CF_CMP(!=)
#undef CF_CMP
static inline cfloat sqrtf(cfloat x) { ++g_ops.sqrt; return mk_res(::sqrtf(x.v), false); }
static inline cfloat powf(cfloat x, cfloat y) { ++g_ops.pow; return mk_res(::powf(x.v, y.v), false); }
static inline cfloat floorf(cfloat x) { return mk_res(::floorf(x.v), false); }   // conversions / rounding: not FP32-pipe arithmetic
static inline cfloat rintf(cfloat x) { return mk_res(::rintf(x.v), false); }
static_assert(sizeof(cfloat) == 4, "cfloat must keep the layout of float");

#define float cfloat
#include "cl_shim.h"
thread_local size_t cl_shim_gid[2] = {0, 0};
thread_local uint cl_shim_depth = 30;
thread_local uint64_t ref_ray_counter = 0;
#define REF_HOOK_RAY() (++ref_ray_counter)
#include "raytracer_cl.inc"   // the reference kernel, generated at build time (oracle/Makefile)
#undef float

static_assert(sizeof(Material) == 48 && sizeof(Sphere) == 32 && sizeof(Plane) == 48 && sizeof(Lens) == 64 &&
              sizeof(Mesh) == 16 && sizeof(Model) == 12, "host AoS layouts (include/scene.h:30-81)");

extern "C" {

struct ref_scene_desc {   // same field order as rt_scene_desc in include/rt_capi.h
    const void* materials; const void* spheres; const void* planes; const void* lenses;
    const void* vertices; const void* uvs; const void* indices; const void* meshes; const void* models;
    uint32_t n_materials, n_spheres, n_planes, n_lenses, n_vertices, n_uvs, n_indices, n_meshes, n_models;
};

// Clears the tag bit of n floats in place (the caller passes private copies of the float parts of its inputs).
void ref_opcount_clean(float* p, size_t n) {
    for (size_t i = 0; i < n; ++i) p[i] = cf_tag(p[i], false);
}

// Samples [first, first+n) of the pixels [x0,x1) x [y0,y1) of a W x H frame, one thread, as render()/renderAgain() would
// run them (sample 0 -> trace, k >= 1 -> retrace).  `image` is a scratch W*H*4 float image.  counts = {add/sub, mul,
// compare, div, sqrt, pow, double mul, FMA-fusable pairs, rays}.
void ref_opcount_render(const ref_scene_desc* d, const float* tex, int tw, int th, int tl, const float* cam12,
                        const float* table, float* image, int W, int H, int x0, int y0, int x1, int y1,
                        uint32_t first, uint32_t n, uint32_t depth, uint64_t counts[9]) {
    Scene s;
    s.materials = (const Material*)d->materials; s.spheres = (const Sphere*)d->spheres;
    s.planes = (const Plane*)d->planes; s.lenses = (const Lens*)d->lenses;
    s.vertex_buffer = (const vec3*)d->vertices; s.texture_uv_buffer = (const vec2*)d->uvs;
    s.index_buffer = (const uint*)d->indices; s.mesh_buffer = (const Mesh*)d->meshes; s.models = (const Model*)d->models;
    s.sphere_count = d->n_spheres; s.plane_count = d->n_planes; s.lens_count = d->n_lenses; s.model_count = d->n_models;
    image2d_t img{(cfloat*)image, W, H};
    image2d_array_t texture{(const cfloat*)tex, tw, th, tl};
    std::memset(&g_ops, 0, sizeof g_ops);
    ref_ray_counter = 0;
    cl_shim_depth = depth;
    for (uint32_t k = first; k < first + n; ++k)
        for (int y = y0; y < y1; ++y)
            for (int x = x0; x < x1; ++x) {
                cl_shim_gid[0] = (size_t)x; cl_shim_gid[1] = (size_t)y;
                if (k == 0) trace(img, (const cfloat*)cam12, (const cfloat*)table, &s, texture);
                else retrace(img, img, (const cfloat*)cam12, (const cfloat*)table, &s, texture, k);
            }
    const uint64_t out[9] = {g_ops.add, g_ops.mul, g_ops.cmp, g_ops.div, g_ops.sqrt, g_ops.pow, g_ops.dmul, g_ops.fusable, ref_ray_counter};
    std::memcpy(counts, out, sizeof out);
}

}  // extern "C"
