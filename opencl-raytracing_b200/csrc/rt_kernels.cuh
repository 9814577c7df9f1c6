// csrc/rt_kernels.cuh — the path-tracing kernels (sm_100a), templated on the arithmetic policy (rt_math.cuh):
//   k_*_strict = <Exact>: bit-exact against the oracle (oracle/rt_oracle.c == the reference kernel as C++)
//   k_*_fast   = <Fused>: FMA contraction + MUFU approximations; primary hits are certified against the
//                Exact result (a lane whose primary id differs re-runs the Exact intersection), so primary-hit
//                ids are identical to the reference's by construction.
//
// Behaviour follows /root/reference/kernels/raytracer.cl ("cl:" below); the structure does not:
//   * one launch renders MANY samples of a pixel; a lane that finishes a path immediately regenerates the
//     next sample's primary ray (persistent lanes), so every trip of the main loop is one nearest-hit query
//     for (almost) all 32 lanes of the warp;
//   * primitives are read from a shared-memory copy of a float4 "test stream" with warp-uniform addresses;
//   * triangle meshes: a conservative bounding-sphere test and the reference's LAST triangle test (front
//     facing, cl:295) run first, per lane; the surviving (ray, triangle) pairs of the whole warp are then
//     compacted with ballot/scan and redistributed over the 32 lanes, so the expensive Moller-Trumbore
//     arithmetic runs at full lane occupancy (warp-level wavefront).  All tests are side-effect free, so the
//     reordering does not change any result;
//   * hit-point data (position, normal, uv, material) is derived once, for the winning primitive only.
#pragma once
#include "rt_bvh.h"
#include "rt_device.h"
#include "rt_math.cuh"
#include <algorithm>
#include <cuda_fp16.h>

namespace rtb {

#define RT_T_MIN 0.001f        // cl:3
#define RT_T_MAX 1000.0f       // cl:4
#define RT_TRI_EPS 0.0000001f  // cl:1
#define RT_TABLE_N 100000u     // cl:7
#define RT_FULL 0xffffffffu

__device__ __forceinline__ bool in_range(float t) { return t >= RT_T_MIN && t <= RT_T_MAX; }  // == cl:127

struct Hit {
    float t;
    uint32_t id;     // kind << 28 | index (triangles: reference-global face number)
    float bu, bv;    // barycentrics (triangles)
    uint32_t aux;    // lens: side; triangle: index into the triangle stream
};

// Shared memory layout:
//   [test stream | mesh table (3 x uint4 each) | triangle plane offsets (n_tri floats, padded to x4) | per-warp lists (1 KB / warp)]
__device__ __forceinline__ uint32_t pd_f4(const DevScene& sc) { return (sc.n_tri + 3u) >> 2; }
__device__ __forceinline__ void stage_scene(const DevScene& sc, float4* smem) {
    // the global buffer has the same layout as the shared-memory copy: [stream | mesh table | plane offsets]
    const uint32_t n = sc.stream_f4 + sc.n_mesh * 3u + pd_f4(sc);
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) smem[i] = __ldg(sc.stream + i);
    __syncthreads();
}
__device__ __forceinline__ const float* smem_tri_pd(const DevScene& sc, const float4* smem) {
    return reinterpret_cast<const float*>(smem + sc.stream_f4 + 3u * sc.n_mesh);
}
__device__ __forceinline__ uint8_t* warp_list(const DevScene& sc, float4* smem) {
    return reinterpret_cast<uint8_t*>(smem + sc.stream_f4 + 3u * sc.n_mesh + pd_f4(sc)) + (threadIdx.x >> 5) * 1024u;
}

// ---------------------------------------------------------------------------------- analytic primitives
// cl:149-255 (spheres, planes, lenses) in hitScene order (cl:327-349); strict `<` keeps the earlier one on ties.
// cl:149-174: near root if it is in range, else the far root; the caller applies t >= T_MIN and the best-so-far test.
template <class A>
__device__ __forceinline__ bool sphere_candidate(const float4 s, V3 o, V3 d, float& t) {
    const V3 oc = mk(A::sub(s.x, o.x), A::sub(s.y, o.y), A::sub(s.z, o.z));
    const float b = dot<A>(oc, d);
    const float c = A::sub(dot<A>(oc, oc), s.w);
    const float dis = A::msub(b, b, c);
    if (!(dis > 0.0f)) return false;
    const float sq = A::sqrt(dis);
    // near root unless it is below T_MIN; a near root beyond T_MAX makes the reference try the (still farther) far root
    // and reject it (cl:163-171) — here the caller's t < best.t <= T_MAX test rejects it, same outcome
    const float t0 = A::sub(b, sq), t1 = A::add(b, sq);
    t = (t0 >= RT_T_MIN) ? t0 : t1;
    return t >= RT_T_MIN;
}
// cl:196-255: intersection of two balls.  Every accepted branch yields t >= T_MIN; t <= T_MAX (cl:245) is implied
// by the caller's best-so-far test.
template <class A>
__device__ __forceinline__ bool lens_candidate(const float4 l1, const float4 l2, V3 o, V3 d, float& t, uint32_t& side) {
    V3 oc = mk(A::sub(l1.x, o.x), A::sub(l1.y, o.y), A::sub(l1.z, o.z));
    const float b1 = dot<A>(oc, d);
    const float dis1 = A::msub(b1, b1, A::sub(dot<A>(oc, oc), l1.w));
    oc = mk(A::sub(l2.x, o.x), A::sub(l2.y, o.y), A::sub(l2.z, o.z));
    const float b2 = dot<A>(oc, d);
    const float dis2 = A::msub(b2, b2, A::sub(dot<A>(oc, oc), l2.w));
    if (!(dis1 > 0.0f && dis2 > 0.0f)) return false;
    const float d1 = A::sqrt(dis1), d2 = A::sqrt(dis2);
    const float t1A = A::sub(b1, d1), t1B = A::add(b1, d1), t2A = A::sub(b2, d2), t2B = A::add(b2, d2);
    t = -1.0f;
    side = 0u;
    if ((t1B < t2A) || (t2B < t1A)) { /* misses the intersection of the two balls, cl:220 */ }
    else if (RT_T_MIN <= t1A || RT_T_MIN <= t2A) {          // outside, cl:221-231
        if (t2A <= t1A) { side = 0u; t = t1A; } else { side = 1u; t = t2A; }
    } else if (RT_T_MIN <= t1B && RT_T_MIN <= t2B) {        // inside, cl:232-242
        if (t1B <= t2B) { side = 0u; t = t1B; } else { side = 1u; t = t2B; }
    }
    return t >= RT_T_MIN;
}
template <class A>
__device__ __forceinline__ void hit_planes(const DevScene& sc, const float4* __restrict__ S, V3 o, V3 d, Hit& best) {
#pragma unroll 1
    for (uint32_t i = 0; i < sc.n_pla; ++i) {                       // cl:176-194
        const float4 p = S[sc.off_pla + 2 * i], n4 = S[sc.off_pla + 2 * i + 1];
        const V3 n = mk(n4.x, n4.y, n4.z);
        const float a = dot<A>(d, n);
        const float b = dot<A>(mk(A::sub(p.x, o.x), A::sub(p.y, o.y), A::sub(p.z, o.z)), n);
        const float t = A::div(b, a);
        if (t >= RT_T_MIN && t < best.t) { best.t = t; best.id = (kKindPlane << 28) | i; }
    }
}
template <class A>
__device__ __forceinline__ void hit_analytic(const DevScene& sc, const float4* __restrict__ S, V3 o, V3 d, Hit& best) {
    for (uint32_t i = 0; i < sc.n_sph; ++i) {
        float t;
        if (sphere_candidate<A>(S[i], o, d, t) && t < best.t) { best.t = t; best.id = (kKindSphere << 28) | i; }
    }
    hit_planes<A>(sc, S, o, d, best);
#pragma unroll 1
    for (uint32_t i = 0; i < sc.n_len; ++i) {
        float t; uint32_t side;
        if (lens_candidate<A>(S[sc.off_len + 2 * i], S[sc.off_len + 2 * i + 1], o, d, t, side) && t < best.t) {
            best.t = t; best.id = (kKindLens << 28) | i; best.aux = side;
        }
    }
}

// cl:257-289 Moller-Trumbore on one triangle of the stream, WITHOUT the front-face test (done by the caller).
template <class A>
__device__ __forceinline__ bool hit_triangle(const float4 q0, const float4 q1, const float4 q2, V3 o, V3 d, float& t, float& u, float& v) {
    const V3 e1 = mk(q0.w, q1.x, q1.y), e2 = mk(q1.z, q1.w, q2.x);
    const V3 h = cross<A>(d, e2);
    const float a = dot<A>(e1, h);
    if (a > -RT_TRI_EPS && a < RT_TRI_EPS) return false;
    const float f = A::rcp(a);
    const V3 sv = mk(A::sub(o.x, q0.x), A::sub(o.y, q0.y), A::sub(o.z, q0.z));
    u = A::mul(f, dot<A>(sv, h));
    if (u < 0.0f || u > 1.0f) return false;
    const V3 qv = cross<A>(sv, e1);
    v = A::mul(f, dot<A>(d, qv));
    if (v < 0.0f || A::add(u, v) > 1.0f) return false;
    t = A::mul(f, dot<A>(e2, qv));
    return in_range(t);
}
template <class A>
__device__ __forceinline__ bool hit_triangle(const float4* __restrict__ T, V3 o, V3 d, float& t, float& u, float& v) {
    return hit_triangle<A>(T[0], T[1], T[2], o, d, t, u, v);
}
template <class A>
__device__ __forceinline__ bool front_facing(const float4* __restrict__ T, V3 d) {   // cl:285,295
    const float4 q2 = T[2];
    return dot<A>(mk(q2.y, q2.z, q2.w), d) < 0.0f;
}

// Per-lane scan of all mesh instances (cl:291-320 folded into hitScene's order).  Used by the probes of
// single rays and by the Exact re-check of a primary hit; the render loop uses nearest_hit_warp below.
template <class A>
__device__ __forceinline__ void hit_meshes_lane(const DevScene& sc, const float4* __restrict__ S, V3 o, V3 d,
                                                uint32_t mesh_mode, Hit& best) {
    const uint4* M = reinterpret_cast<const uint4*>(S + sc.stream_f4);
    for (uint32_t m = 0; m < sc.n_mesh; ++m) {
        const uint4 mi = M[3 * m];   // tri_first, tri_count, face_id_base, mat_id
        bool open = true;
        for (uint32_t k = 0; k < mi.y; ++k) {
            const float4* T = S + sc.off_tri + 3 * (mi.x + k);
            float t, u, v;
            if (open && front_facing<A>(T, d) && hit_triangle<A>(T, o, d, t, u, v)) {
                if (mesh_mode == 0u) open = false;                  // first front-facing hit ends this mesh
                if (t < best.t) { best.t = t; best.id = (kKindTriangle << 28) | (mi.z + k); best.bu = u; best.bv = v; best.aux = mi.x + k; }
            }
        }
    }
}

template <class A>
__device__ __forceinline__ void nearest_hit_lane(const DevScene& sc, const float4* __restrict__ S, V3 o, V3 d,
                                                 uint32_t mesh_mode, Hit& best) {
    best.t = RT_T_MAX; best.id = kIdMiss; best.bu = 0.0f; best.bv = 0.0f; best.aux = 0u;
    hit_analytic<A>(sc, S, o, d, best);
    hit_meshes_lane<A>(sc, S, o, d, mesh_mode, best);
}
// The Exact re-check of a knife-edge primary hit is cold code: keep it out of the hot instruction stream.
static __device__ __noinline__ Hit nearest_hit_exact_cold(const DevScene& sc, const float4* __restrict__ S, float ox, float oy, float oz,
                                                          float dx, float dy, float dz, uint32_t mesh_mode) {
    Hit best;
    nearest_hit_lane<Exact>(sc, S, mk(ox, oy, oz), mk(dx, dy, dz), mesh_mode, best);
    return best;
}

// ---------------------------------------------------------------------------------- LBVH traversal
// Nearest hit over the finite primitives through the GPU-built LBVH (rt_bvh.cu).  Candidates arrive in traversal
// order, not in hitScene order, so acceptance is by (t, id) lexicographic order — id = kind << 28 | index grows in
// hitScene's scan order (spheres, planes, lenses, triangles; index order inside a kind), which reproduces "on equal t
// the earlier one wins" (strict < at cl:328-352).  Meshes use the nearest-front-facing rule (RT_MESH_NEAREST): the
// reference's first-front-facing-in-face-order rule (cl:291-303) depends on the scan order and cannot be answered by
// a spatial hierarchy.  Box tests are conservative (boxes padded at build time, exit distance scaled up), the
// primitive tests are the same policy-A routines as in the brute-force path, so the winner equals brute force.
__device__ __forceinline__ bool closer(float t, uint32_t id, const Hit& best) {
    return t < best.t || (t == best.t && t < RT_T_MAX && id < best.id);
}
// One leaf = one primitive; its 64-byte record (DevBvh::leaves, leaf order = Morton order) holds the test data of the
// primitive itself, so a leaf visit is one fetch of two sectors next to those of the neighbouring leaves, not an
// index followed by a dependent fetch from the scene's stream.
// One 64-byte record (a node or a leaf) in TWO 256-bit loads (LDG.E.256, sm_100) instead of four 128-bit ones: the LBVH render
// kernel keeps the L1 load/store data pipe busier (68 % of its peak, ncu) than the issue slots (65 %), and three quarters of
// its wavefronts are these fetches — one lane, one record, 32 different lines per instruction.
#ifndef RT_LOAD256
#define RT_LOAD256 1      // ptxas 12.9 crashes on ld.global.nc.v8.f32 in these kernels; v4.u64 compiles to the same LDG.E.ENL2.256
#endif
__device__ __forceinline__ void load64(const float4* __restrict__ p, float4& a, float4& b, float4& c, float4& d) {
#if RT_LOAD256
    unsigned long long u0, u1, u2, u3, v0, v1, v2, v3;
    asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(u0), "=l"(u1), "=l"(u2), "=l"(u3) : "l"(p));
    asm volatile("ld.global.nc.v4.u64 {%0,%1,%2,%3}, [%4];" : "=l"(v0), "=l"(v1), "=l"(v2), "=l"(v3) : "l"(p + 2));
    auto lo = [](unsigned long long x) { return __uint_as_float((unsigned)x); };
    auto hi = [](unsigned long long x) { return __uint_as_float((unsigned)(x >> 32)); };
    a = make_float4(lo(u0), hi(u0), lo(u1), hi(u1)); b = make_float4(lo(u2), hi(u2), lo(u3), hi(u3));
    c = make_float4(lo(v0), hi(v0), lo(v1), hi(v1)); d = make_float4(lo(v2), hi(v2), lo(v3), hi(v3));
#else
    a = __ldg(p); b = __ldg(p + 1); c = __ldg(p + 2); d = __ldg(p + 3);
#endif
}
// Work counters of the instrumented (untimed) LBVH render kernel: the numerator of the roofline of a BVH run is COUNTED
// work, not the brute-force operation count.
struct WorkCount { uint32_t nodes, sph, len, tri, flared_nodes, flared_queries; };
template <class A, bool COUNT = false>
__device__ __forceinline__ void bvh_leaf(const float4* __restrict__ L, V3 o, V3 d, Hit& best, WorkCount* wc = nullptr) {
    float4 r0, r1, r2, r3;
    load64(L, r0, r1, r2, r3);
    const uint32_t id = __float_as_uint(r3.x), idx = __float_as_uint(r3.y), kind = id >> 28;
    if (kind == kKindSphere) {
        float t;
        if (COUNT) ++wc->sph;
        if (sphere_candidate<A>(r0, o, d, t) && closer(t, id, best)) { best.t = t; best.id = id; best.aux = 0u; }
    } else if (kind == kKindLens) {
        float t; uint32_t side;
        if (COUNT) ++wc->len;
        if (lens_candidate<A>(r0, r1, o, d, t, side) && closer(t, id, best)) { best.t = t; best.id = id; best.aux = side; }
    } else {
        float t, u, v;
        if (COUNT) ++wc->tri;
        if (dot<A>(mk(r2.y, r2.z, r2.w), d) < 0.0f && hit_triangle<A>(r0, r1, r2, o, d, t, u, v) && closer(t, id, best)) {   // cl:285,295
            best.t = t; best.id = id; best.bu = u; best.bv = v; best.aux = idx;
        }
    }
}
// The reference's sphere test computes dis = b*b - (oc.oc - r*r) in binary32 (cl:152-155) and ASSUMES a unit direction.
// Two consequences a traversal that must return brute force's nearest hit has to live with:
//  (1) rounding: the absolute error of dis grows like ~12 eps |oc|^2, so from a distance D the test reports hits up to
//      sqrt(r^2 + 7e-7 D^2) - r OUTSIDE the sphere (8.5e-4 D for a small sphere);
//  (2) refracted directions are not normalised (cl:386,428) and sphere normals are (p - c) / r of an inexact hit point
//      (cl:160), so |d|^2 = a drifts away from 1 along glass chains (C5: 8 % of the queries have a - 1 > 1e-3, the
//      largest is 6).  With b = oc.d the test accepts dis > 0  <=>  p^2 < r^2 + (a - 1) b_u^2  (p = distance of the centre
//      from the ray, b_u = its distance along the ray): for a > 1 the sphere "flares" with distance and brute force
//      reports hits on spheres the ray passes at up to sqrt(a - 1) b_u (found by tests/test_gpu_crops.py: 6 % of the
//      pixels of the C5 crop differed before this was handled; the default scene's few big spheres never showed it).
// So every box that can hold a sphere or lens is inflated, before the slab test, by slack x (an upper bound of the
// distance from the ray origin to its farthest corner), slack = ray_slack(d).slack >= sqrt(7e-7 max(1, a) + max(a - 1, 0)), 1e-3 for a unit direction.
// Twelve more instructions per box; triangle-only nodes (BvhNode::flags == 0) do not need it in the fast arm — Moller-
// Trumbore is exact in any parametrisation.
#define RT_SLAB_SLACK 1.0e-3f
#ifndef RT_SLAB_L2
#define RT_SLAB_L2 1
#endif
// Per-ray constants of the box tests (a = |d|^2):
//   ordinary ray (a - 1 <= 3e-6)       slack = 1e-3 (the round-1 value), kappa = 1, sqa = 0: inflated slab, pruned by its entry.
//   THIN flare (3e-6 < a - 1 <= RT_BVH_THIN_FLARE, scene without lenses)   stays in the ordinary code path: the box is inflated
//       by the whole flare, slack >= sqrt(7e-7 a + (a - 1)), which for these rays is a few percent of the distance, and the
//       entry parameter is turned into a valid lower bound of the hit parameter: the point of closest approach to an accepted
//       sphere's centre (parameter t_c) lies inside the inflated box, and t0 >= sqrt(a)(sqrt(a) - sqrt(a - 1)) t_c - r, so
//       t0 >= kappa * t_entry - rho with rho = half the smallest extent of the box (>= r of any sphere inside).
//   WIDE flare: sqa = sqrt(a) > 0, boxes tested by slab_flared_core below (cone test + analytic bound), in trips of their own;
//       sqa < 0  a lens may be inside any flagged box (its two spheres' centres are NOT inside its box, so none of the bounds
//       holds): every flagged box is inflated by the full flare, nothing is pruned by distance — slow, but exact; only small
//       test scenes take this path.
#ifndef RT_BVH_THIN_FLARE
#define RT_BVH_THIN_FLARE 1.0e-2f
#endif
struct RaySlack { float slack, sqa, kappa; };
__device__ __forceinline__ RaySlack ray_slack(V3 d, bool lens_free = true) {
    const float a = fmaf(d.x, d.x, fmaf(d.y, d.y, d.z * d.z));
    RaySlack r;
    const bool flared = !(a - 1.0f <= 3e-6f);                                // NaN counts as flared
    const bool thin = flared && lens_free && a - 1.0f <= RT_BVH_THIN_FLARE;
    // ordinary and thin: shell^2 <= (7e-7 max(1, a) + max(a - 1, 0)) D^2, 2 % margin; wide flare: the shell of Moller-Trumbore
    // only (the spheres have their own test, slab_flared_core)
    r.slack = (flared && !thin) ? RT_SLAB_SLACK * fmaxf(1.0f, sqrtf(a))
                                : fmaxf(RT_SLAB_SLACK, 1.02f * sqrtf(fmaf(7e-7f, fmaxf(1.0f, a), fmaxf(a - 1.0f, 0.0f))));
    r.sqa = (flared && !thin) ? ((lens_free && a < 3e5f) ? sqrtf(a) : -sqrtf(fminf(a, 1e30f))) : 0.0f;
    r.kappa = thin ? 0.998f * sqrtf(a) * (sqrtf(a) - sqrtf(a - 1.0f)) : 1.0f;
    if (!(r.sqa == r.sqa)) { r.sqa = -1e15f; r.slack = RT_SLAB_SLACK; r.kappa = 1.0f; }
    return r;
}
// `thin`: some lane of the warp may carry a thin flare (kappa != 1) — a warp-uniform switch where the caller has one, so that
// warps without such a ray do not issue the rho / kappa instructions at all.
__device__ __forceinline__ float2 node_rho(float packed) {                // BvhNode::rho -> (child 0, child 1)
    const uint32_t u = __float_as_uint(packed);
    return make_float2(__half2float(__ushort_as_half((unsigned short)(u & 0xffffu))), __half2float(__ushort_as_half((unsigned short)(u >> 16))));
}
template <bool ROBUST>
__device__ __forceinline__ bool slab(const float lo[3], const float hi[3], V3 o, V3 inv, float t_best, float& t_near, float slack = RT_SLAB_SLACK,
                                     float kappa = 1.0f, bool thin = true, float rho = 0.0f) {
    float l0 = lo[0] - o.x, h0 = hi[0] - o.x, l1 = lo[1] - o.y, h1 = hi[1] - o.y, l2 = lo[2] - o.z, h2 = hi[2] - o.z;
    if (ROBUST) {
        // distance from the origin to the farthest corner of the box (>= |oc| of any centre inside): the Euclidean one — the
        // sum of the three components, up to 1.73 times as much, tested 3 times as many leaf boxes of C5 as needed
        const float m0 = fmaxf(fabsf(l0), fabsf(h0)), m1 = fmaxf(fabsf(l1), fabsf(h1)), m2 = fmaxf(fabsf(l2), fabsf(h2));
#if RT_SLAB_L2
        float far_corner;                                                  // approximate root (2 ulp): covered by the 1.00001
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(far_corner) : "f"(fmaf(m0, m0, fmaf(m1, m1, m2 * m2))));
        const float dl = slack * 1.00001f * far_corner;
#else
        const float dl = slack * (m0 + m1 + m2);
#endif
        l0 -= dl; h0 += dl; l1 -= dl; h1 += dl; l2 -= dl; h2 += dl;
    }
    const float ax = l0 * inv.x, bx = h0 * inv.x;
    const float ay = l1 * inv.y, by = h1 * inv.y;
    const float az = l2 * inv.z, bz = h2 * inv.z;
    t_near = fmaxf(fmaxf(fminf(ax, bx), fminf(ay, by)), fmaxf(fminf(az, bz), 0.0f));
    const float t_far = fminf(fminf(fmaxf(ax, bx), fmaxf(ay, by)), fmaxf(az, bz)) * 1.0000004f;   // Ize 2013: robust exit
    const bool pass = t_near <= t_far;
    // thin flare: bound of t, not the entry; rho = BvhNode::rho of this child (>= the radius of any sphere inside the box)
    if (ROBUST && thin) t_near = kappa != 1.0f ? fmaxf(fmaf(kappa, t_near, -rho), 0.0f) : t_near;
    return pass && t_near <= t_best;
}
// Box test of a FLARED ray (cold path; 8 % of the queries of C5, none of the default scene).  The accepted hits of such a ray
// are not near the spheres they belong to (t0 = b - sqrt(dis) of an aligned sphere at distance L is ~0.6 L / |d| for a = 1.3)
// and lie in a cone around the ray, so the slab test neither bounds their parameter nor decides whether a box matters.
// Both answers follow from the test itself.  For a sphere of radius r <= rho whose centre lies in the box, at distance L
// from the origin (L_min <= L <= L_max), with b = oc.d <= b_max = max over the box:
//   accepted        =>  c = L^2 - r^2 < 0 (origin inside)  or  b > 0 and b^2 > c - (rounding of dis, <= 1.4e-6 a L^2)
//   its parameter   t0 >= sqrt(a) L - sqrt((a - 1) L^2 + r^2)      (minimum over b, at b = sqrt(a) L; grows with L, falls with r)
// rho = half the smallest extent of the box (a sphere inside a box is no larger).  The lower bound orders the children and
// prunes against the best hit, which keeps a flared ray as local as any other; without the two tests every box of the tree
// passed (a step of C5 went from 0.2 s to 44 s).  Triangles in the same subtree are exact in any parametrisation: the
// ordinary inflated slab decides for them, and the box passes if either test passes.
// `tri`: the subtree may hold triangles (scene-wide: the scene has any), `curved`: it may hold spheres.
__device__ __forceinline__ bool slab_flared_core(const float lo[3], const float hi[3], V3 o, V3 d, V3 inv, float t_best, float slack, float sqa,
                                                 bool curved, bool tri, float& t_near_out) {
    bool pass = false;
    float t_near = 3.0e38f;
    if (tri || !curved) {                                                  // triangles: exact in any parametrisation
        float t_tri;
        pass = slab<true>(lo, hi, o, inv, t_best, t_tri, slack);
        if (pass) t_near = t_tri;
    }
    if (curved) {
        const float l0 = lo[0] - o.x, h0 = hi[0] - o.x, l1 = lo[1] - o.y, h1 = hi[1] - o.y, l2 = lo[2] - o.z, h2 = hi[2] - o.z;
        const float ex = fmaxf(fmaxf(l0, -h0), 0.0f), ey = fmaxf(fmaxf(l1, -h1), 0.0f), ez = fmaxf(fmaxf(l2, -h2), 0.0f);
        const float Lmin2 = fmaf(ex, ex, fmaf(ey, ey, ez * ez));
        const float mx = fmaxf(fabsf(l0), fabsf(h0)), my = fmaxf(fabsf(l1), fabsf(h1)), mz = fmaxf(fabsf(l2), fabsf(h2));
        const float Lmax2 = fmaf(mx, mx, fmaf(my, my, mz * mz));
        const float rho = 0.5f * fminf(fminf(h0 - l0, h1 - l1), h2 - l2), rho2 = rho * rho;
        const float a = sqa * sqa;
        const float bmax = fmaxf(l0 * d.x, h0 * d.x) + fmaxf(l1 * d.y, h1 * d.y) + fmaxf(l2 * d.z, h2 * d.z);
        const float c_min = Lmin2 - rho2;
        if (c_min <= 0.0f || (bmax > 0.0f && fmaf(bmax, bmax * 1.0002f, 3e-6f * a * Lmax2) >= c_min * 0.9998f)) {
            // lower bound of t; 0.998 covers the rounding shell of the reference's test (<= 8.5e-4 sqrt(a) L in t) and this
            // arithmetic.  Boxes too large for a bound (c_min <= 0) are ordered by their distance: 1e-6 L <= 0.001 <= any t.
            const float L = sqrtf(Lmin2);
            float lb = 1e-6f * L;
            // sqrt(a) L - sqrt((a - 1) L^2 + rho'^2) without the cancellation; rho'^2 = rho^2 + the rounding shell of dis
            const float rho2s = fmaf(3e-6f * a, Lmin2, rho2);
            if (Lmin2 > rho2s) lb = fmaxf(lb, 0.999f * (Lmin2 - rho2s) / fmaf(sqa, L, sqrtf(fmaf(a - 1.0f, Lmin2, rho2s))));
            if (lb <= t_best) { pass = true; t_near = fminf(t_near, lb); }
        }
    }
    t_near_out = t_near;
    return pass;
}
static __device__ __noinline__ bool slab_flared(float lx, float ly, float lz, float hx, float hy, float hz, float ox, float oy, float oz,
                                                float dx_, float dy_, float dz_, float ix, float iy, float iz, float t_best, float slack, float sqa,
                                                bool curved, float* t_near_out) {
    const float lo[3] = {lx, ly, lz}, hi[3] = {hx, hy, hz};
    float tn;
    const bool pass = slab_flared_core(lo, hi, mk(ox, oy, oz), mk(dx_, dy_, dz_), mk(ix, iy, iz), t_best, slack, sqa, curved, true, tn);
    *t_near_out = tn;
    return pass;
}
// One child box of a node in the robust path: ordinary rays take the inflated slab, flared rays the cold routine above.
__device__ __forceinline__ bool slab_robust(const float lo[3], const float hi[3], V3 o, V3 d, V3 inv, float t_best, float& t_near,
                                            const RaySlack rs, bool curved, float rho) {
    if (rs.sqa == 0.0f) return slab<true>(lo, hi, o, inv, t_best, t_near, rs.slack, rs.kappa, true, rho);
    if (rs.sqa < 0.0f) {                                                   // flared, lenses possible: full flare inflation, no distance pruning
        const float a = rs.sqa * rs.sqa;
        float tn;
        const bool pass = slab<true>(lo, hi, o, inv, 3.0e38f, tn, curved ? fmaxf(rs.slack, 1.02f * sqrtf(fmaf(7e-7f, a, a - 1.0f))) : rs.slack);
        t_near = curved ? 0.0f : tn;
        return pass && t_near <= t_best;
    }
    return slab_flared(lo[0], lo[1], lo[2], hi[0], hi[1], hi[2], o.x, o.y, o.z, d.x, d.y, d.z, inv.x, inv.y, inv.z, t_best, rs.slack, rs.sqa, curved, &t_near);
}
template <class A>
__device__ __forceinline__ void hit_bvh(const DevScene& sc, const float4* __restrict__ S, const DevBvh& bvh, V3 o, V3 d, Hit& best) {
    if (bvh.n_leaves == 0u) return;
    if (bvh.root < 0) { bvh_leaf<A>(bvh.leaves + 4 * (~bvh.root), o, d, best); return; }
    const V3 inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
    const RaySlack rs = ray_slack(d, sc.n_len == 0u);
    int stack[kBvhStack];
    int sp = 0, node = bvh.root;
    for (;;) {
        const float4* N = reinterpret_cast<const float4*>(bvh.nodes + node);
        float4 n0, n1, n2, n3;
        load64(N, n0, n1, n2, n3);
        const float lo0[3] = {n0.x, n0.y, n0.z}, hi0[3] = {n1.x, n1.y, n1.z}, lo1[3] = {n2.x, n2.y, n2.z}, hi1[3] = {n3.x, n3.y, n3.z};
        const int c0 = __float_as_int(n0.w), c1 = __float_as_int(n1.w);
        float t0, t1;
        const uint32_t fl = __float_as_uint(n2.w);
        const float2 rho = node_rho(n3.w);
        bool h0 = slab_robust(lo0, hi0, o, d, inv, best.t, t0, rs, (fl & 1u) != 0u, rho.x), h1 = slab_robust(lo1, hi1, o, d, inv, best.t, t1, rs, (fl & 2u) != 0u, rho.y);
        if (h0 && c0 < 0) { bvh_leaf<A>(bvh.leaves + 4 * (~c0), o, d, best); h0 = false; }
        if (h1 && c1 < 0) { bvh_leaf<A>(bvh.leaves + 4 * (~c1), o, d, best); h1 = false; }
        if (h0 && h1) {                                   // both internal: nearer first, the other on the stack
            const bool first0 = t0 <= t1;
            if (sp < kBvhStack) stack[sp++] = first0 ? c1 : c0;
            node = first0 ? c0 : c1;
        } else if (h0) node = c0;
        else if (h1) node = c1;
        else {
            if (sp == 0) break;
            node = stack[--sp];
        }
    }
}
template <class A>
__device__ __forceinline__ void nearest_hit_bvh(const DevScene& sc, const float4* __restrict__ S, const DevBvh& bvh, V3 o, V3 d, Hit& best) {
    best.t = RT_T_MAX; best.id = kIdMiss; best.bu = 0.0f; best.bv = 0.0f; best.aux = 0u;
    hit_planes<A>(sc, S, o, d, best);
    hit_bvh<A>(sc, S, bvh, o, d, best);
}
static __device__ __noinline__ Hit nearest_hit_exact_cold_bvh(const DevScene& sc, const float4* __restrict__ S, const DevBvh& bvh,
                                                              float ox, float oy, float oz, float dx, float dy, float dz) {
    Hit best;
    nearest_hit_bvh<Exact>(sc, S, bvh, mk(ox, oy, oz), mk(dx, dy, dz), best);
    return best;
}

// ---------------------------------------------------------------------------------- warp-cooperative meshes
// All 32 lanes call this convergently (inactive lanes pass active = false).  For each mesh instance:
//   1. per lane: conservative ray / bounding-sphere test against the lane's current best t (a "maybe" filter;
//      it can only skip triangles that could not be accepted, so it is not parity relevant);
//   2. per lane, per 32-triangle chunk: front-facing test (cl:295) -> list of surviving triangle slots in the
//      warp's shared-memory scratch;
//   3. warp scan of the list lengths; the (lane, triangle) items of the whole warp are dealt out to the 32
//      lanes round by round; each lane fetches its owner's ray with shuffles and runs Moller-Trumbore;
//   4. owners collect their hits in face order with ballot + shuffles, applying the reference's rule
//      (mesh_mode 0: the first front-facing hit of the mesh wins, cl:291-303) or the nearest rule (mode 1).
template <class A>
__device__ __forceinline__ void hit_meshes_warp(const DevScene& sc, const float4* __restrict__ S, uint8_t* __restrict__ list,
                                                V3 o, V3 d, bool active, uint32_t mesh_mode, Hit& best) {
    const uint32_t lane = threadIdx.x & 31u;
    const uint4* M = reinterpret_cast<const uint4*>(S + sc.stream_f4);
    for (uint32_t m = 0; m < sc.n_mesh; ++m) {
        const uint4 mi = M[3 * m];                      // tri_first, tri_count, face_id_base, mat_id
        const float4 bs = S[sc.stream_f4 + 3 * m + 2];  // bounding sphere (c, r), inflated on the host
        bool need = active;
        if (need) {
            const float ocx = bs.x - o.x, ocy = bs.y - o.y, ocz = bs.z - o.z;
            const float a = fmaf(d.x, d.x, fmaf(d.y, d.y, d.z * d.z));
            const float b = fmaf(ocx, d.x, fmaf(ocy, d.y, ocz * d.z));
            const float cc = fmaf(ocx, ocx, fmaf(ocy, ocy, ocz * ocz)) - bs.w * bs.w;
            const float disc = fmaf(b, b, -a * cc);
            // 2 % slack on the entry distance: the filter must never reject a triangle that could be accepted
            float sq;                                                      // approximate root: the 2 % slack covers it
            asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(fmaxf(disc, 0.0f)));
            need = disc >= 0.0f && !(cc > 0.0f && b < 0.0f) && (b - sq) * 0.98f - 1e-3f * a < best.t * a;
        }
        if (!__any_sync(RT_FULL, need)) continue;
        bool open = need;
        for (uint32_t c0 = 0; c0 < mi.y; c0 += 32u) {
            const uint32_t chunk = min(32u, mi.y - c0);
            const float4* T0 = S + sc.off_tri + 3u * (mi.x + c0);
            uint32_t cnt = 0u;
            if (open) {
                // A front-facing triangle can only be hit at t > 0 from the front side of its plane.  pdm is the
                // plane offset n.A lowered by a safety margin on the host, o_margin covers the rounding of n.o, so
                // this conservative filter only drops triangles whose Moller-Trumbore t would be negative.
                const float* PD = smem_tri_pd(sc, S) + mi.x + c0;
                const float o_margin = 4e-6f * (1.0f + fabsf(o.x) + fabsf(o.y) + fabsf(o.z));
                for (uint32_t k = 0; k < chunk; ++k) {
                    const float4 q2 = T0[3u * k + 2u];
                    const bool front = dot<A>(mk(q2.y, q2.z, q2.w), d) < 0.0f;                            // cl:285,295
                    const bool ahead = fmaf(q2.y, o.x, fmaf(q2.z, o.y, fmaf(q2.w, o.z, o_margin))) >= PD[k];
                    if (front && ahead) { list[cnt * 32u + lane] = (uint8_t)k; ++cnt; }
                }
            }
            uint32_t incl = cnt;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                const uint32_t v = __shfl_up_sync(RT_FULL, incl, off);
                if ((int)lane >= off) incl += v;
            }
            const uint32_t total = __shfl_sync(RT_FULL, incl, 31);
            const uint32_t excl = incl - cnt;
            __syncwarp();
            for (uint32_t base = 0; base < total; base += 32u) {
                const uint32_t j = base + lane;
                const bool valid = j < total;
                uint32_t owner = 0u;                    // smallest lane L with incl[L] > j
#pragma unroll
                for (uint32_t step = 16u; step > 0u; step >>= 1) {
                    const uint32_t v = __shfl_sync(RT_FULL, incl, (owner + step - 1u) & 31u);
                    if (v <= j) owner += step;
                }
                owner &= 31u;
                const uint32_t oex = __shfl_sync(RT_FULL, excl, owner);
                const uint32_t k = valid ? (uint32_t)list[(j - oex) * 32u + owner] : 0u;
                const V3 oo = mk(__shfl_sync(RT_FULL, o.x, owner), __shfl_sync(RT_FULL, o.y, owner), __shfl_sync(RT_FULL, o.z, owner));
                const V3 dd = mk(__shfl_sync(RT_FULL, d.x, owner), __shfl_sync(RT_FULL, d.y, owner), __shfl_sync(RT_FULL, d.z, owner));
                float t = 0.0f, u = 0.0f, v = 0.0f;
                const bool hit = valid && hit_triangle<A>(T0 + 3u * k, oo, dd, t, u, v);
                const uint32_t cand = __ballot_sync(RT_FULL, hit);
                // my items occupy lanes [excl - base, incl - base) of this round
                const int lo = max((int)excl - (int)base, 0), hi = min((int)incl - (int)base, 32);
                uint32_t mine = 0u;
                if (hi > lo) mine = cand & ((hi >= 32 ? 0xffffffffu : ((1u << hi) - 1u)) & ~((1u << lo) - 1u));
                if (!open) mine = 0u;
                while (__any_sync(RT_FULL, mine != 0u)) {
                    const uint32_t src = mine ? (uint32_t)(__ffs((int)mine) - 1) : lane;
                    const float tw = __shfl_sync(RT_FULL, t, src), uw = __shfl_sync(RT_FULL, u, src), vw = __shfl_sync(RT_FULL, v, src);
                    const uint32_t kw = __shfl_sync(RT_FULL, k, src);
                    if (mine) {
                        if (tw < best.t) {
                            best.t = tw; best.id = (kKindTriangle << 28) | (mi.z + c0 + kw);
                            best.bu = uw; best.bv = vw; best.aux = mi.x + c0 + kw;
                        }
                        if (mesh_mode == 0u) { open = false; mine = 0u; }   // first front-facing hit ends this mesh
                        else mine &= mine - 1u;
                    }
                }
            }
            __syncwarp();
        }
    }
}

template <class A>
__device__ __forceinline__ void nearest_hit_warp(const DevScene& sc, const float4* __restrict__ S, uint8_t* __restrict__ list,
                                                 V3 o, V3 d, bool active, uint32_t mesh_mode, Hit& best) {
    best.t = RT_T_MAX; best.id = kIdMiss; best.bu = 0.0f; best.bv = 0.0f; best.aux = 0u;
    if (active) hit_analytic<A>(sc, S, o, d, best);
    __syncwarp();
    hit_meshes_warp<A>(sc, S, list, o, d, active, mesh_mode, best);
}

// ------------------------------------------------------------------------------------------------ RNG
// Table mode, cl:113-125: index = (hash(dir) + k) % 100000, hash = (uint)fabs(double(dot(dir, C)) * 438.5453).
// The 64-bit sum of the reference is reduced modulo 100000 term by term (same residue, 32-bit arithmetic).
template <class A>
__device__ __forceinline__ uint32_t dir_hash(V3 d) {
    const float f = dot<A>(d, mk(123.9898f, 348.233f, 433.3314f));
    return (uint32_t)fabs((double)f * 438.5453);
}
struct PixelRng {
    uint32_t vec_base;   // (gx*3931 + gy*2504) % N      cl:115
    uint32_t uni_base;   // (gx*3931 + gy) % N           cl:122
    uint32_t hp;         // counter generator: per-pixel key
};
__device__ __forceinline__ uint32_t mix32(uint32_t x) {
    x ^= x >> 16; x *= 0x7feb352du; x ^= x >> 15; x *= 0x846ca68bu; x ^= x >> 16;
    return x;
}
__device__ __forceinline__ PixelRng make_pixel_rng(uint32_t gx, uint32_t gy, uint32_t W, uint32_t seed) {
    PixelRng r;
    r.vec_base = (gx * 3931u + gy * 2504u) % RT_TABLE_N;
    r.uni_base = (gx * 3931u + gy) % RT_TABLE_N;
    r.hp = mix32(seed ^ mix32((gy * W + gx) ^ 0xA511E9B3u));
    return r;
}
template <class A>
__device__ __forceinline__ V3 table_vec(const float* __restrict__ tab, const PixelRng& pr, V3 d, uint32_t s_seed) {
    const uint32_t k = ((s_seed * 2683u) % RT_TABLE_N + pr.vec_base) % RT_TABLE_N;
    const uint32_t seed = (dir_hash<A>(d) % RT_TABLE_N + (k * 3u) % RT_TABLE_N) % RT_TABLE_N;
    return mk(__ldg(tab + seed), __ldg(tab + seed + 1), __ldg(tab + seed + 2));
}
template <class A>
__device__ __forceinline__ float table_uniform(const float* __restrict__ tab, const PixelRng& pr, V3 d, uint32_t s_seed) {
    const uint32_t k = ((s_seed * 2683u) % RT_TABLE_N + pr.uni_base) % RT_TABLE_N;
    const uint32_t seed = 3u * RT_TABLE_N + (dir_hash<A>(d) % RT_TABLE_N + k) % RT_TABLE_N;
    return __ldg(tab + seed);
}
// Counter mode: key(pixel) -> key(sample) -> key(bounce) -> draw j, all lowbias32 mixes.
__device__ __forceinline__ uint32_t bounce_key(uint32_t hp, uint32_t sample, uint32_t bounce) {
    const uint32_t hs = mix32(hp ^ (sample * 0x9E3779B9u));
    return mix32(hs + bounce * 0x85EBCA6Bu + 0x1B873593u);
}
__device__ __forceinline__ uint32_t draw(uint32_t hb, uint32_t j) { return mix32(hb + j * 0xC2B2AE35u); }
__device__ __forceinline__ float u01(uint32_t h) { return (float)(h >> 8) * 5.9604644775390625e-8f; }
__device__ __forceinline__ float u16f(uint32_t x) { return (float)x * 1.52587890625e-5f; }            // 16 bits -> [0,1)
template <class A>
__device__ __forceinline__ float s16f(uint32_t x) { return A::sub(A::mul((float)x, 3.0517578125e-5f), 1.0f); }   // -> [-1,1)
// Uniform in the unit ball from exactly-rounded operations: Marsaglia disc rejection (one draw = two 16-bit
// coordinates per try) + max of three 16-bit uniforms for the radius (CDF r^3, the law of cbrt(U)).
template <class A>
__device__ __forceinline__ V3 rng_ball(uint32_t hb, uint32_t base, uint32_t r3bits) {
    float u = 0.0f, v = 0.0f, s = 0.0f;
    for (uint32_t k = 0; k < 16u; ++k) {
        const uint32_t h = draw(hb, base + k);
        u = s16f<A>(h >> 16);
        v = s16f<A>(h & 0xFFFFu);
        s = A::mad(v, v, A::mul(u, u));
        if (s < 1.0f) break;
        u = v = s = 0.0f;
    }
    const float q = A::sqrt(A::sub(1.0f, s));
    const float dx = A::mul(A::mul(2.0f, u), q), dy = A::mul(A::mul(2.0f, v), q), dz = A::sub(1.0f, A::mul(2.0f, s));
    const uint32_t hr = draw(hb, base + 16u);
    const float r = fmaxf(fmaxf(u16f(hr >> 16), u16f(hr & 0xFFFFu)), u16f(r3bits));
    return mk(A::mul(dx, r), A::mul(dy, r), A::mul(dz, r));
}
// Same law as the reference's misaligned table read (cl:109-117): (a.x,a.y,a.z) | (a.y,a.z,b.x) | (a.z,b.x,b.y).
// Draw 0: high half -> selector, low half -> third radius uniform of ball a.
template <class A>
__device__ __forceinline__ V3 counter_vec(uint32_t hb) {
    const uint32_t d0 = draw(hb, 0u);
    const uint32_t sel = ((d0 >> 16) * 3u) >> 16;
    const V3 a = rng_ball<A>(hb, 1u, d0 & 0xFFFFu);
    if (sel == 0u) return a;
    const V3 b = rng_ball<A>(hb, 33u, draw(hb, 50u) >> 16);
    return sel == 1u ? mk(a.y, a.z, b.x) : mk(a.z, b.x, b.y);
}

// ------------------------------------------------------------------------------------------------ texture
// cl:105-107 + sampler cl:15: normalized coordinates, linear filter with full-precision FP32 weights
// (OpenCL 1.2 §8.2), taps clamped to the edge.  Done in software: the texture unit's 8-bit weights would
// not reproduce the reference's filter.
__device__ __forceinline__ float4 texel(const DevTexture& t, int i, int j, int l) {
    i = max(0, min(i, t.w - 1));
    j = max(0, min(j, t.h - 1));
    return __ldg(t.texels + ((size_t)l * t.h + j) * t.w + i);
}
template <class A>
__device__ __forceinline__ V3 texture_rgb(const DevTexture& t, float s, float tt, uint32_t layer) {
    const float u = A::mul(s, (float)t.w), v = A::mul(tt, (float)t.h);
    const float um = A::sub(u, 0.5f), vm = A::sub(v, 0.5f);
    const float fu = floorf(um), fv = floorf(vm);
    const float a = A::sub(um, fu), b = A::sub(vm, fv);
    const int i0 = (int)fu, j0 = (int)fv;
    int l = (int)rintf((float)layer);
    l = max(0, min(l, t.layers - 1));
    const float4 t00 = texel(t, i0, j0, l), t10 = texel(t, i0 + 1, j0, l);
    const float4 t01 = texel(t, i0, j0 + 1, l), t11 = texel(t, i0 + 1, j0 + 1, l);
    const float ia = A::sub(1.0f, a), ib = A::sub(1.0f, b);
    const float w00 = A::mul(ia, ib), w10 = A::mul(a, ib), w01 = A::mul(ia, b), w11 = A::mul(a, b);
    return mk(A::mad(w11, t11.x, A::mad(w01, t01.x, A::mad(w10, t10.x, A::mul(w00, t00.x)))),
              A::mad(w11, t11.y, A::mad(w01, t01.y, A::mad(w10, t10.y, A::mul(w00, t00.y)))),
              A::mad(w11, t11.z, A::mad(w01, t01.z, A::mad(w10, t10.z, A::mul(w00, t00.z)))));
}

// ------------------------------------------------------------------------------------------------ shading
struct ShadeEnv {
    const DevScene* sc;
    const DevTexture* tex;
    const float* table;
    uint32_t rng_mode;
};

template <class A>
__device__ __forceinline__ V3 reflect_dir(V3 d, V3 n) {            // cl:362-365
    const float k = A::mul(2.0f, dot<A>(d, n));
    return A::normalize(vsub<A>(d, vscale<A>(n, k)));
}

// Material of the winning primitive (cl:162, 189, 250, 314) as stored per primitive by the host: material index in the
// low 24 bits, its type (cl:23) in the high 8 — the path loop asks "is it a light?" after every query and must not
// pay a second dependent load for that.
__device__ __forceinline__ uint32_t mat_index(uint32_t packed) { return packed & 0x00FFFFFFu; }
__device__ __forceinline__ int mat_type(uint32_t packed) { return (int)(packed >> 24); }
__device__ __forceinline__ uint32_t hit_material(const DevScene& sc, const float4* __restrict__ S, const Hit& h) {
    const uint32_t kind = h.id >> 28, idx = h.id & 0x0FFFFFFFu;
    if (kind == kKindSphere) return __ldg(sc.sph_mat + idx);
    if (kind == kKindPlane) return __ldg(sc.pla_mat + idx);
    if (kind == kKindLens) return __ldg(sc.len_mat + idx);
    const uint4* M = reinterpret_cast<const uint4*>(S + sc.stream_f4);
    return M[3 * __ldg(sc.tri_mesh + h.aux)].w;
}

// Hit-point data for the winning primitive (cl:158-161, 185-188, 246-249, 281-285, 296, 314) followed by one
// iteration of getCol's switch (cl:454-479).  Returns true when the path ends (light).
template <class A>
__device__ __forceinline__ bool shade(const ShadeEnv& env, const float4* __restrict__ S, const Hit& h, uint32_t mat_id, const PixelRng& pr,
                                      V3& o, V3& d, V3& out, uint32_t sample, uint32_t bounce) {
    const DevScene& sc = *env.sc;
    const uint32_t kind = h.id >> 28, idx = h.id & 0x0FFFFFFFu;
    // issue the loads of this bounce first, so their latency overlaps the hit-point arithmetic below: the material
    // record, and (counter-table RNG) the table entry of this (pixel, sample, bounce) — speculative, used by the
    // diffuse / textured / dielectric branches only
    const float4 m0 = __ldg(sc.mat + 2 * mat_index(mat_id)), m1 = __ldg(sc.mat + 2 * mat_index(mat_id) + 1);
    uint32_t tix = 0u;
    V3 rv_table = mk(0.0f, 0.0f, 0.0f);
    if (env.rng_mode == 2u) {
        tix = __umulhi(bounce_key(pr.hp, sample, bounce), RT_TABLE_N);
        rv_table = mk(__ldg(env.table + tix), __ldg(env.table + tix + 1), __ldg(env.table + tix + 2));
    }
    const V3 p = vmad<A>(d, h.t, o);                                // cl:141-143
    V3 n;
    uint32_t tex_id = 0u;
    float uvx = 0.0f, uvy = 0.0f;
    if (kind == kKindSphere) {
        const float4 s = __ldg(sc.sph_shade + idx);
        n = A::vdiv(vsub<A>(p, mk(s.x, s.y, s.z)), s.w);
    } else if (kind == kKindPlane) {
        const float4 pn = __ldg(sc.pla_normal + idx);
        const V3 nn = mk(pn.x, pn.y, pn.z);
        const float a = dot<A>(d, nn);
        const float sg = (a != a) ? 0.0f : (a > 0.0f ? 1.0f : (a < 0.0f ? -1.0f : a));
        n = vscale<A>(neg(nn), sg);                                 // cl:187
    } else if (kind == kKindLens) {
        const float4 l = __ldg(sc.len_shade + 2 * idx + h.aux);
        n = A::vdiv(vsub<A>(p, mk(l.x, l.y, l.z)), l.w);
    } else {
        const float4 q2 = S[sc.off_tri + 3 * h.aux + 2];
        n = mk(q2.y, q2.z, q2.w);
        const uint32_t mesh = __ldg(sc.tri_mesh + h.aux);
        const uint4* M = reinterpret_cast<const uint4*>(S + sc.stream_f4);
        tex_id = M[3 * mesh + 1].x;
        const float2 ua = __ldg(sc.tri_uv + 3 * h.aux), ub = __ldg(sc.tri_uv + 3 * h.aux + 1), uc = __ldg(sc.tri_uv + 3 * h.aux + 2);
        const float w = A::sub(A::sub(1.0f, h.bu), h.bv);           // cl:101-103
        uvx = A::mad(uc.x, h.bv, A::mad(ub.x, h.bu, A::mul(ua.x, w)));
        uvy = A::mad(uc.y, h.bv, A::mad(ub.y, h.bu, A::mul(ua.y, w)));
    }
    const int type = __float_as_int(m0.x);
    const float extra = m0.y;
    V3 color = mk(m1.x, m1.y, m1.z);
    const uint32_t s_seed = bounce + sample;                        // cl:456,472,476
    // counter modes: key of this (pixel, sample, bounce); computed only inside the branches that draw
    auto bounce_hash = [&]() { return bounce_key(pr.hp, sample, bounce); };

    if (type == 3 || type == 4) {                                   // diffuse / textured: rayScatter cl:393-399
        V3 rv;
        if (env.rng_mode == 2u) rv = rv_table;                      // the reference's table at a counter-keyed uniform index
        else if (env.rng_mode == 0u) rv = table_vec<A>(env.table, pr, d, s_seed);
        else rv = counter_vec<A>(bounce_hash());
        d = A::normalize(vadd<A>(n, rv));
        o = p;
        out = vscale<A>(out, extra);
        if (type == 4) color = texture_rgb<A>(*env.tex, uvx, uvy, tex_id);   // cl:477
        out = vmin(out, color);
        return false;
    }
    if (type == 5) { out = vmin(out, color); return true; }         // light cl:459-462
    if (type == 1) {                                                // reflective cl:463-466
        o = p;
        d = reflect_dir<A>(d, n);
        out = vscale<A>(out, extra);
        out = vmin(out, color);
        return false;
    }
    if (type == 0 || type == 2) {                                   // refractive cl:369-391 / dielectric cl:407-435
        V3 nn; float ratio;
        float cai = dot<A>(d, n);
        if (cai > 0.0f) { nn = neg(n); ratio = extra; cai = -cai; }
        else { nn = n; ratio = A::rcp(extra); }
        bool try_refract = true;
        if (type == 2) {
            float r0 = A::div(A::sub(1.0f, ratio), A::add(1.0f, ratio));   // schlick cl:401-405
            r0 = A::mul(r0, r0);
            const float prob = A::mad(A::sub(1.0f, r0), A::pow5(A::sub(1.0f, -cai)), r0);
            const float u = (env.rng_mode == 2u) ? __ldg(env.table + 3u * RT_TABLE_N + tix)
                          : (env.rng_mode == 0u) ? table_uniform<A>(env.table, pr, d, s_seed) : u01(draw(bounce_hash(), 0u));
            try_refract = prob < u;                                 // cl:423
        }
        const float disc = A::sub(1.0f, A::mul(A::mul(ratio, ratio), A::sub(1.0f, A::mul(cai, cai))));
        o = p;
        if (try_refract && disc > 0.0f) {
            const float k = A::mad(ratio, cai, A::sqrt(disc));
            d = vsub<A>(vscale<A>(d, ratio), vscale<A>(nn, k));     // not normalised, cl:386/428
        } else {
            d = reflect_dir<A>(d, nn);
        }
        out = vmin(out, color);
        return false;
    }
    return false;   // unknown type: the reference's switch has no default, the ray is left unchanged
}

// ------------------------------------------------------------------------------------------------ kernels
// Always instantiated with Exact: a primary ray is computed once per pixel, and sharing it bit for bit between the
// arithmetic arms is what makes the certification of primary hits meaningful.
template <class A>
__device__ __forceinline__ V3 primary_dir(const float* cam, int x, int y, int W, int H) {   // cl:129-139, 500-501
    const float s = A::div((float)x, (float)W), t = A::div((float)y, (float)H);
    const V3 llc = mk(cam[3], cam[4], cam[5]), hor = mk(cam[6], cam[7], cam[8]), ver = mk(cam[9], cam[10], cam[11]);
    return A::normalize(vadd<A>(vadd<A>(llc, vscale<A>(hor, s)), vscale<A>(ver, t)));
}

// trace (cl:496-510) + retrace (cl:512-532) + getCol (cl:444-486) for samples [first, first+n) of every pixel.
// One lane owns one pixel; warps cover 8x4 pixel tiles.  Lanes are persistent over samples: the moment a
// path ends the lane accumulates it and restarts from the primary ray, so each trip of the while loop is one
// nearest-hit query executed by every lane that still has samples left.
// BVH = false: brute force over the shared-memory copy of the scene (small scenes);
// BVH = true : planes brute force + LBVH traversal, scene data read from global memory (large scenes).
template <class A, bool BVH>
__device__ __forceinline__ void trace_body(const TraceLaunch& p, float4* smem) {
    const float4* S = p.sc.stream;
    uint8_t* list = nullptr;
    if (!BVH) {
        stage_scene(p.sc, smem);
        S = smem;
        list = warp_list(p.sc, smem);
    }

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tiles_x = (p.W + 7) >> 3;
    const int tile = blockIdx.x * (kTraceThreads / 32) + warp;
    const int x = (tile % tiles_x) * 8 + (lane & 7);
    const int y = (tile / tiles_x) * 4 + (lane >> 3);
    const bool live = x >= p.rx0 && x < p.rx1 && y >= p.ry0 && y < p.ry1 &&   // the region lies inside the frame
                      (p.split_n <= 1u || blockIdx.x % p.split_n == p.split_i);   // block = slot group of the pixel split

    const V3 cam_o = mk(p.cam[0], p.cam[1], p.cam[2]);
    const V3 d0 = primary_dir<Exact>(p.cam, x, y, p.W, p.H);   // same primary ray in both arithmetic arms
    const PixelRng pr = make_pixel_rng((uint32_t)x, (uint32_t)y, (uint32_t)p.W, p.seed);
    ShadeEnv env; env.sc = &p.sc; env.tex = &p.tex; env.table = p.table; env.rng_mode = p.rng_mode;
    float4* px = p.target + (size_t)y * p.W + x;
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    uint32_t certified = kIdMiss;
    if (live) {
        if (p.accum_mode == 1u || p.first > 0u) acc = *px;
        if (!A::kExact && p.certify_id) certified = __ldg(p.certify_id + (size_t)y * p.W + x);
    }
    const uint32_t k_end = p.first + p.n;
    uint32_t k = live ? p.first : k_end, i = 0u;
    V3 o = cam_o, d = d0, out = mk(1.0f, 1.0f, 1.0f);
    unsigned long long rays = 0ull;

    while (__any_sync(RT_FULL, k < k_end)) {
        const bool active = k < k_end;
        Hit h;
        if (BVH) { if (active) nearest_hit_bvh<A>(p.sc, S, p.bvh, o, d, h); }
        else nearest_hit_warp<A>(p.sc, S, list, o, d, active, p.mesh_mode, h);
        if (active) {
            if (!A::kExact && p.certify_id && i == 0u && h.id != certified)   // knife-edge pixel: take the exact primary hit
                h = BVH ? nearest_hit_exact_cold_bvh(p.sc, S, p.bvh, o.x, o.y, o.z, d.x, d.y, d.z)
                        : nearest_hit_exact_cold(p.sc, S, o.x, o.y, o.z, d.x, d.y, d.z, p.mesh_mode);
            ++rays;
            bool ended;
            if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); ended = true; }   // cl:450-452
            else ended = shade<A>(env, S, h, hit_material(p.sc, S, h), pr, o, d, out, k, i);
            ++i;
            if (ended || i >= p.depth) {
                if (p.accum_mode == 1u) {                   // linear sum (what the sample split reduces)
                    acc.x = A::add(acc.x, out.x); acc.y = A::add(acc.y, out.y); acc.z = A::add(acc.z, out.z); acc.w += 1.0f;
                } else if (k == 0u) {                        // cl:509
                    acc = make_float4(A::sqrt(out.x), A::sqrt(out.y), A::sqrt(out.z), 1.0f);
                } else {                                     // cl:524-531: mix(new, prev^2, k/(k+1))
                    const float a = A::div((float)k, (float)(k + 1u));
                    const float pr_ = A::mul(acc.x, acc.x), pg_ = A::mul(acc.y, acc.y), pb_ = A::mul(acc.z, acc.z);
                    acc = make_float4(A::sqrt(A::mad(A::sub(pr_, out.x), a, out.x)), A::sqrt(A::mad(A::sub(pg_, out.y), a, out.y)),
                                      A::sqrt(A::mad(A::sub(pb_, out.z), a, out.z)), 1.0f);
                }
                ++k; i = 0u; o = cam_o; d = d0; out = mk(1.0f, 1.0f, 1.0f);
            }
        }
    }
    if (live) *px = acc;
    for (int off = 16; off > 0; off >>= 1) rays += __shfl_down_sync(RT_FULL, rays, off);   // one atomic per warp
    if (lane == 0 && rays) atomicAdd(p.rays, rays);
}

// LBVH render kernel.  Traversal lengths differ wildly between rays (tens to hundreds of nodes), so running one whole
// query per trip — as the brute-force kernel does — leaves most lanes idle (measured: 4.9 of 32 lanes active).  Here
// the main loop advances every lane by ONE node per trip; a lane whose traversal is finished waits until at least
// `regen` lanes are waiting (or nobody traverses any more), then the waiting lanes shade, bounce or regenerate the
// next sample's primary ray, and re-enter the traversal loop (persistent lanes with in-loop ray fetch).
template <class A>
__device__ __forceinline__ void trace_body_bvh(const TraceLaunch& p) {
    const float4* S = p.sc.stream;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tiles_x = (p.W + 7) >> 3;
    const int tile = blockIdx.x * (kTraceThreads / 32) + warp;
    const int x = (tile % tiles_x) * 8 + (lane & 7);
    const int y = (tile / tiles_x) * 4 + (lane >> 3);
    const bool live = x >= p.rx0 && x < p.rx1 && y >= p.ry0 && y < p.ry1 &&   // the region lies inside the frame
                      (p.split_n <= 1u || blockIdx.x % p.split_n == p.split_i);   // block = slot group of the pixel split

    const V3 cam_o = mk(p.cam[0], p.cam[1], p.cam[2]);
    const V3 d0 = primary_dir<Exact>(p.cam, x, y, p.W, p.H);
    const PixelRng pr = make_pixel_rng((uint32_t)x, (uint32_t)y, (uint32_t)p.W, p.seed);
    ShadeEnv env; env.sc = &p.sc; env.tex = &p.tex; env.table = p.table; env.rng_mode = p.rng_mode;
    float4* px = p.target + (size_t)y * p.W + x;
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    uint32_t certified = kIdMiss;
    if (live) {
        if (p.accum_mode == 1u || p.first > 0u) acc = *px;
        if (!A::kExact && p.certify_id) certified = __ldg(p.certify_id + (size_t)y * p.W + x);
    }
    const uint32_t k_end = p.first + p.n;
    uint32_t k = live ? p.first : k_end, i = 0u;
    V3 o = cam_o, d = d0, out = mk(1.0f, 1.0f, 1.0f);
    unsigned long long rays = 0ull;
    const uint32_t regen = p.bvh_regen ? p.bvh_regen : 8u;

    Hit h;
    V3 inv = mk(0.0f, 0.0f, 0.0f);
    RaySlack rs; rs.slack = RT_SLAB_SLACK; rs.sqa = 0.0f; rs.kappa = 1.0f;
    int stack[kBvhStack];
    int sp = 0, node = 0;
    bool traversing = false;
    auto start_ray = [&]() {
        h.t = RT_T_MAX; h.id = kIdMiss; h.bu = 0.0f; h.bv = 0.0f; h.aux = 0u;
        hit_planes<A>(p.sc, S, o, d, h);
        sp = 0; node = p.bvh.root; traversing = p.bvh.n_leaves > 0u;
        if (traversing && node < 0) { bvh_leaf<A>(p.bvh.leaves + 4 * (~node), o, d, h); traversing = false; }
        inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
        rs = ray_slack(d, p.sc.n_len == 0u);
    };
    bool has_ray = k < k_end;
    if (has_ray) start_ray();

    while (__any_sync(RT_FULL, has_ray)) {
        for (;;) {                                           // ---- traversal: one node per trip
            if (traversing) {
                const float4* N = reinterpret_cast<const float4*>(p.bvh.nodes + node);
                float4 n0, n1, n2, n3;
                load64(N, n0, n1, n2, n3);
                const float lo0[3] = {n0.x, n0.y, n0.z}, hi0[3] = {n1.x, n1.y, n1.z}, lo1[3] = {n2.x, n2.y, n2.z}, hi1[3] = {n3.x, n3.y, n3.z};
                const int c0 = __float_as_int(n0.w), c1 = __float_as_int(n1.w);
                float t0, t1;
                const uint32_t fl = __float_as_uint(n2.w);
                const float2 rho = node_rho(n3.w);
                bool h0 = slab_robust(lo0, hi0, o, d, inv, h.t, t0, rs, (fl & 1u) != 0u, rho.x), h1 = slab_robust(lo1, hi1, o, d, inv, h.t, t1, rs, (fl & 2u) != 0u, rho.y);
                if (h0 && c0 < 0) { bvh_leaf<A>(p.bvh.leaves + 4 * (~c0), o, d, h); h0 = false; }
                if (h1 && c1 < 0) { bvh_leaf<A>(p.bvh.leaves + 4 * (~c1), o, d, h); h1 = false; }
                if (h0 && h1) {
                    const bool first0 = t0 <= t1;
                    if (sp < kBvhStack) stack[sp++] = first0 ? c1 : c0;
                    node = first0 ? c0 : c1;
                } else if (h0) node = c0;
                else if (h1) node = c1;
                else if (sp > 0) node = stack[--sp];
                else traversing = false;
            }
            const unsigned trav = __ballot_sync(RT_FULL, traversing);
            const unsigned waiting = __ballot_sync(RT_FULL, has_ray && !traversing);
            if (trav == 0u || (uint32_t)__popc(waiting) >= regen) break;
        }
        if (has_ray && !traversing) {                        // ---- shading / regeneration for the waiting lanes
            if (!A::kExact && p.certify_id && i == 0u && h.id != certified)
                h = nearest_hit_exact_cold_bvh(p.sc, S, p.bvh, o.x, o.y, o.z, d.x, d.y, d.z);
            ++rays;
            bool ended;
            if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); ended = true; }
            else ended = shade<A>(env, S, h, hit_material(p.sc, S, h), pr, o, d, out, k, i);
            ++i;
            if (ended || i >= p.depth) {
                if (p.accum_mode == 1u) {
                    acc.x = A::add(acc.x, out.x); acc.y = A::add(acc.y, out.y); acc.z = A::add(acc.z, out.z); acc.w += 1.0f;
                } else if (k == 0u) {
                    acc = make_float4(A::sqrt(out.x), A::sqrt(out.y), A::sqrt(out.z), 1.0f);
                } else {
                    const float a = A::div((float)k, (float)(k + 1u));
                    const float pr_ = A::mul(acc.x, acc.x), pg_ = A::mul(acc.y, acc.y), pb_ = A::mul(acc.z, acc.z);
                    acc = make_float4(A::sqrt(A::mad(A::sub(pr_, out.x), a, out.x)), A::sqrt(A::mad(A::sub(pg_, out.y), a, out.y)),
                                      A::sqrt(A::mad(A::sub(pb_, out.z), a, out.z)), 1.0f);
                }
                ++k; i = 0u; o = cam_o; d = d0; out = mk(1.0f, 1.0f, 1.0f);
            }
            has_ray = k < k_end;
            if (has_ray) start_ray();
        }
    }
    if (live) *px = acc;
    for (int off = 16; off > 0; off >>= 1) rays += __shfl_down_sync(RT_FULL, rays, off);
    if (lane == 0 && rays) atomicAdd(p.rays, rays);
}

// ------------------------------------------------------------------------------------------------ path heads
// See DevHead (rt_device.h): the head of a pixel's path — everything up to its first hit on a surface whose
// material draws random numbers — is the same for every sample, because the reference shoots the same primary ray
// for each sample (cl:500-505, 517-522) and mirrors / refractive surfaces / lights / misses are deterministic
// (cl:454-479).  k_head_* walks it once per pixel, always in Exact arithmetic (so the primary hits of the fast arm
// are the reference's by construction); k_trace_tail_* restarts every sample from the stored state.

// One sample's radiance into the accumulator: linear sum, or the reference's sqrt(mean) chain (cl:509, 524-531).
template <class A>
__device__ __forceinline__ void accumulate_sample(float4& acc, V3 out, uint32_t k, uint32_t accum_mode) {
    if (accum_mode == 1u) {
        acc.x = A::add(acc.x, out.x); acc.y = A::add(acc.y, out.y); acc.z = A::add(acc.z, out.z); acc.w += 1.0f;
    } else if (k == 0u) {
        acc = make_float4(A::sqrt(out.x), A::sqrt(out.y), A::sqrt(out.z), 1.0f);
    } else {
        const float a = A::div((float)k, (float)(k + 1u));
        const float pr_ = A::mul(acc.x, acc.x), pg_ = A::mul(acc.y, acc.y), pb_ = A::mul(acc.z, acc.z);
        acc = make_float4(A::sqrt(A::mad(A::sub(pr_, out.x), a, out.x)), A::sqrt(A::mad(A::sub(pg_, out.y), a, out.y)),
                          A::sqrt(A::mad(A::sub(pb_, out.z), a, out.z)), 1.0f);
    }
}
__device__ __forceinline__ void slot_pixel(uint32_t slot, int W, int& x, int& y) {
    const int tiles_x = (W + 7) >> 3;
    const int tile = (int)(slot >> 5), lane = (int)(slot & 31u);
    x = (tile % tiles_x) * 8 + (lane & 7);
    y = (tile / tiles_x) * 4 + (lane >> 3);
}

template <bool BVH>
__device__ __forceinline__ void head_body(const HeadLaunch& p, float4* smem) {
    const float4* S = p.sc.stream;
    uint8_t* list = nullptr;
    if (!BVH) { stage_scene(p.sc, smem); S = smem; list = warp_list(p.sc, smem); }
    const uint32_t slot = blockIdx.x * kTraceThreads + threadIdx.x;
    int x, y;
    slot_pixel(slot, p.W, x, y);
    const bool live = slot < p.n_slots && x >= p.rx0 && x < p.rx1 && y >= p.ry0 && y < p.ry1 &&
                      (p.split_n <= 1u || blockIdx.x % p.split_n == p.split_i);
    uint32_t seg = 0u;
    V3 o = mk(p.cam[0], p.cam[1], p.cam[2]), d = primary_dir<Exact>(p.cam, x, y, p.W, p.H), out = mk(1.0f, 1.0f, 1.0f);
    uint32_t i = 0u, mat_id = 0u;
    uint8_t flag = kHeadTerminal;
    Hit h; h.t = 0.0f; h.id = kIdMiss; h.bu = h.bv = 0.0f; h.aux = 0u;
    ShadeEnv env; env.sc = &p.sc; env.tex = nullptr; env.table = nullptr; env.rng_mode = 0u;
    PixelRng pr; pr.vec_base = pr.uni_base = pr.hp = 0u;
    bool walking = live;
    while (__any_sync(RT_FULL, walking)) {                   // the warp-cooperative triangle stage needs all 32 lanes
        Hit q;                                               // lanes that stopped keep the hit they stopped at
        if (BVH) { if (walking) nearest_hit_bvh<Exact>(p.sc, S, p.bvh, o, d, q); }
        else nearest_hit_warp<Exact>(p.sc, S, list, o, d, walking, p.mesh_mode, q);
        if (walking) {
            h = q;
            ++seg;
            if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); walking = false; }              // cl:450-452
            else {
                mat_id = hit_material(p.sc, S, h);
                const int type = mat_type(mat_id);
                if (type == 5) {                                                                // light, cl:459-462
                    const float4 m1 = __ldg(p.sc.mat + 2 * mat_index(mat_id) + 1);
                    out = vmin(out, mk(m1.x, m1.y, m1.z));
                    walking = false;
                } else if (type == 2 || type == 3 || type == 4) {                               // draws random numbers
                    flag = kHeadStochastic; walking = false;
                } else {
                    shade<Exact>(env, S, h, mat_id, pr, o, d, out, 0u, i);                      // mirror / refractive
                    if (++i >= p.depth) walking = false;                                        // cl:447: loop exhausted
                }
            }
        }
    }
    if (live) {
        if (flag == kHeadStochastic) {
            p.rec[slot] = make_float4(o.x, o.y, o.z, h.t);
            p.rec[p.n_slots + slot] = make_float4(d.x, d.y, d.z, h.bu);
            p.rec[2u * p.n_slots + slot] = make_float4(out.x, out.y, out.z, h.bv);
            p.rec[3u * p.n_slots + slot] = make_float4(__uint_as_float(i), __uint_as_float(h.id), __uint_as_float(h.aux), __uint_as_float(mat_id));
        } else {
            p.rec[2u * p.n_slots + slot] = make_float4(out.x, out.y, out.z, 0.0f);
        }
        p.flag[slot] = flag;
    } else if (slot < p.n_slots) {
        p.flag[slot] = kHeadDead;
    }
    // class of the head for the ordering of the stochastic list: material type, primitive kind and — brute-force scenes
    // only: with a hierarchy the tile order inside a class is worth more than the primitive — the primitive's low bits
    const bool stoch = live && flag == kHeadStochastic;
    if (slot < p.n_slots)
        p.key[slot] = stoch ? ((((uint32_t)mat_type(mat_id) << 11) | ((h.id >> 28) << 8) | (BVH ? 0u : (h.id & 0xFFu))) & p.key_mask) : kHeadKeyOther;
    const uint32_t n_st = __popc(__ballot_sync(RT_FULL, stoch));
    seg = __reduce_add_sync(RT_FULL, seg);
    if ((threadIdx.x & 31) == 0) {
        if (seg) atomicAdd(p.segments, (unsigned long long)seg);
        if (n_st) atomicAdd(p.segments + 1, (unsigned long long)n_st);
    }
}

// Pixels whose whole path is deterministic: every sample has the colour stored by k_head.
template <class A>
__device__ __forceinline__ void terminal_body(const TraceLaunch& p, uint32_t j) {
    if (j >= p.head.n_slots - p.head.n_stoch) return;
    const uint32_t slot = __ldg(p.head.list + p.head.n_stoch + j);
    if (__ldg(p.head.flag + slot) != kHeadTerminal) return;
    int x, y;
    slot_pixel(slot, p.W, x, y);
    const float4 r2 = __ldg(p.head.rec + 2u * p.head.n_slots + slot);
    const V3 out = mk(r2.x, r2.y, r2.z);
    float4* px = p.target + (size_t)y * p.W + x;
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (p.accum_mode == 1u || p.first > 0u) acc = *px;
    const bool black = out.x == 0.0f && out.y == 0.0f && out.z == 0.0f;
    if (black && p.accum_mode == 1u && acc.w == floorf(acc.w) && acc.w >= 0.0f && acc.w + (float)p.n <= 16777216.0f) {
        acc.w += (float)p.n;                                   // n exact additions of 1.0f; x + 0.0f == x
    } else if (black && p.accum_mode == 0u && acc.x == 0.0f && acc.y == 0.0f && acc.z == 0.0f) {
        acc = make_float4(0.0f, 0.0f, 0.0f, 1.0f);             // the sqrt(mean) chain of zeros stays zero
    } else {
        for (uint32_t k = p.first; k < p.first + p.n; ++k) accumulate_sample<A>(acc, out, k, p.accum_mode);
    }
    *px = acc;
}

#ifndef RT_STEAL
#define RT_STEAL 1                  // brute-force render kernel, fast arm: idle lanes take samples from busy ones
#endif
// trace + retrace + getCol for samples [first, first+n) of the stochastic pixels, restarting from the path heads.
// Every lane entering a trip holds a pending hit on a surface that is neither a miss nor a light: the trip shades it,
// traces the bounced ray, and closes the path on a miss (cl:450-452), a light (cl:459-462) or at the depth limit.  So
// every live lane does exactly one nearest-hit query per trip; a finished path restarts from the stored head in place.
//
// Work items.  One lane renders all n samples of one stochastic pixel — or, when the host asks for sample chunks
// (TraceLaunch::chunks > 1, linear sums only), chunk_spp of them.  Grid block -> (chunk, block of kTraceThreads
// consecutive list entries), in list order or with the most expensive pixel blocks first (TraceLaunch::block_order): the
// paths of neighbouring image regions differ in length by multiples (glass, mirrors facing mirrors), a block lives as
// long as its slowest lane, and in list order the late, expensive blocks left the SMs idle for a quarter of the kernel
// (ncu: sm__cycles_active.avg / sm__cycles_elapsed = 0.73).  See order_blocks / plan_chunks in rt_capi.cpp.
template <class A>
__device__ __forceinline__ void trace_tail_body(const TraceLaunch& p, float4* smem) {
    const uint32_t chunks = p.chunks > 1u ? p.chunks : 1u;
    const uint32_t pixel_blocks = (p.head.n_stoch + kTraceThreads - 1u) / kTraceThreads;
    const uint32_t stoch_blocks = pixel_blocks * chunks;
    if (blockIdx.x >= stoch_blocks) { terminal_body<A>(p, (blockIdx.x - stoch_blocks) * kTraceThreads + threadIdx.x); return; }
    stage_scene(p.sc, smem);
    const float4* S = smem;
    uint8_t* list = warp_list(p.sc, smem);
    float4* hs = reinterpret_cast<float4*>(reinterpret_cast<uint8_t*>(smem + p.sc.stream_f4 + 3u * p.sc.n_mesh + pd_f4(p.sc)) +
                                           (kTraceThreads / 32) * 1024u) + threadIdx.x;   // [4][kTraceThreads] restart state
    const uint32_t lane = threadIdx.x & 31u;
    // block -> (chunk c, pixel block pb): list order, or the most expensive pixel blocks first (TraceLaunch::block_order)
    uint32_t c, pb;
    if (p.block_order) { c = blockIdx.x % chunks; pb = __ldg(p.block_order + blockIdx.x / chunks); }
    else { c = blockIdx.x / pixel_blocks; pb = blockIdx.x - c * pixel_blocks; }
    const uint32_t j = pb * kTraceThreads + threadIdx.x;
    const bool live = j < p.head.n_stoch;
    const uint32_t slot = live ? __ldg(p.head.list + j) : 0u;
    int x, y;
    slot_pixel(slot, p.W, x, y);
#pragma unroll
    for (uint32_t r = 0; r < 4u; ++r) hs[r * kTraceThreads] = __ldg(p.head.rec + r * p.head.n_slots + slot);

    PixelRng pr = make_pixel_rng((uint32_t)x, (uint32_t)y, (uint32_t)p.W, p.seed);
    ShadeEnv env; env.sc = &p.sc; env.tex = &p.tex; env.table = p.table; env.rng_mode = p.rng_mode;
    float4* px = chunks > 1u ? p.partial + ((size_t)c * p.head.n_stoch + j) : p.target + (size_t)y * p.W + x;
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    if (live && chunks == 1u && (p.accum_mode == 1u || p.first > 0u)) acc = *px;
    uint32_t k = p.first, k_end = p.first + p.n;
    if (chunks > 1u) { k = min(k + c * p.chunk_spp, k_end); k_end = min(k + p.chunk_spp, k_end); }
    if (!live) k = k_end;
    uint32_t i, mat_id;
    V3 o, d, out;
    Hit h;
    uint32_t root = lane;                                    // the lane whose pixel this lane works on (sample stealing below)
    auto restart = [&]() {
        const float4* hsp = hs - lane + root;                // restart state of that pixel
        const float4 r0 = hsp[0], r1 = hsp[kTraceThreads], r2 = hsp[2 * kTraceThreads], r3 = hsp[3 * kTraceThreads];
        o = mk(r0.x, r0.y, r0.z); d = mk(r1.x, r1.y, r1.z); out = mk(r2.x, r2.y, r2.z);
        h.t = r0.w; h.bu = r1.w; h.bv = r2.w;
        i = __float_as_uint(r3.x); h.id = __float_as_uint(r3.y); h.aux = __float_as_uint(r3.z); mat_id = __float_as_uint(r3.w);
    };
    restart();
    uint32_t queries = 0u;
    // Sample stealing (fast arm, linear sums).  A warp lives as long as its slowest lane, and the lanes' totals differ: the
    // pixels differ, and so do sums of n random path lengths (by ~2 / sqrt(n) of the mean between the extremes of 32 lanes) —
    // 2 of 32 lanes idle over a launch of 1024 samples per lane (ncu), more for shorter items (one rank of an 8-way sample
    // split).  A lane that has run out of samples takes the upper half of the samples not yet started of the lane that has
    // most left, with that pixel's RNG key and restart state, sums them in an accumulator of its own (`sacc`) and hands the sum
    // to the owner of the pixel (`root`, a lane of this warp) when it steals again or the warp ends.  One steal per trip; which
    // lane steals what depends on the paths alone, so the picture stays reproducible bit for bit, but a pixel's samples are
    // added in another order than sample order — the strict arm and the reference's sqrt(mean) chain never steal.
    const bool steal = RT_STEAL && !A::kExact && p.accum_mode == 1u && p.steal != 0u;
    float4 sacc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    auto hand_over = [&](unsigned from) {                    // the warp: every lane in `from` gives its sacc to the owner of the pixel
        for (; from; from &= from - 1u) {
            const int src = __ffs((int)from) - 1;
            const float4 v = make_float4(__shfl_sync(RT_FULL, sacc.x, src), __shfl_sync(RT_FULL, sacc.y, src),
                                         __shfl_sync(RT_FULL, sacc.z, src), __shfl_sync(RT_FULL, sacc.w, src));
            const uint32_t to = __shfl_sync(RT_FULL, root, src);
            if (lane == to) { acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w; }
            if ((int)lane == src) sacc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        }
    };

    for (;;) {
        const unsigned busy = __ballot_sync(RT_FULL, k < k_end);
        if (busy == 0u) break;
        if (steal && busy != RT_FULL) {                      // warp uniform
            // victim: most samples not yet started (the one in work stays where it is)
            const uint32_t spare = k < k_end ? k_end - k - 1u : 0u;
            const uint32_t top = __reduce_max_sync(RT_FULL, (min(spare, 0x03ffffffu) << 5) | lane);
            if ((top >> 5) != 0u) {
                const int v = (int)(top & 31u), t = __ffs((int)~busy) - 1;
                hand_over(__ballot_sync(RT_FULL, (int)lane == t && root != lane));       // the thief's previous loot first
                const uint32_t take = ((top >> 5) + 1u) >> 1;
                const uint32_t v_end = __shfl_sync(RT_FULL, k_end, v), v_root = __shfl_sync(RT_FULL, root, v);
                const uint32_t vb = __shfl_sync(RT_FULL, pr.vec_base, v), ub = __shfl_sync(RT_FULL, pr.uni_base, v), hp = __shfl_sync(RT_FULL, pr.hp, v);
                if ((int)lane == v) k_end -= take;
                if ((int)lane == t) {
                    k = v_end - take; k_end = v_end;
                    pr.vec_base = vb; pr.uni_base = ub; pr.hp = hp;
                    root = v_root;
                    restart();
                }
            }
        }
        const bool active = k < k_end;
        if (active) { shade<A>(env, S, h, mat_id, pr, o, d, out, k, i); ++i; }
        const bool q = active && i < p.depth;
        nearest_hit_warp<A>(p.sc, S, list, o, d, q, p.mesh_mode, h);
        if (active) {
            bool ended = !q;                                                        // cl:447: depth exhausted, keep out
            if (q) {
                ++queries;
                if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); ended = true; }  // cl:450-452
                else {
                    mat_id = hit_material(p.sc, S, h);
                    if (mat_type(mat_id) == 5) {                                    // light, cl:459-462
                        const float4 m1 = __ldg(p.sc.mat + 2 * mat_index(mat_id) + 1);
                        out = vmin(out, mk(m1.x, m1.y, m1.z));
                        ended = true;
                    }
                }
            }
            if (ended) {
                if (root == lane) accumulate_sample<A>(acc, out, k, p.accum_mode);
                else accumulate_sample<A>(sacc, out, k, p.accum_mode);
                ++k; restart();
            }
        }
    }
    if (steal) hand_over(__ballot_sync(RT_FULL, root != lane));
    if (live) *px = acc;
    queries = __reduce_add_sync(RT_FULL, queries);                                   // < 2^32 per warp launch
    if (lane == 0u && queries) {
        atomicAdd(p.queries, (unsigned long long)queries);
        if (p.block_cost) atomicAdd(p.block_cost + pb, queries);
    }
}

// Warp-cooperative brute force over ALL leaves for the ray of lane `src` (called convergently by the 32 lanes).  For rays the
// hierarchy cannot help: a very wide flare (|d|^2 > RT_BVH_BRUTE_A: the acceptance cone of the reference's sphere test opens
// to nearly the forward half space, cos^2 = 1 / |d|^2), or any flared ray in a scene with lenses (no bound holds for them).
// Such a ray walks most of the tree; done by one lane that is 10^5 dependent fetches (0.1 s for C5 — a handful of them, a few in
// one pixel, made every launch last a second), done by the warp it is n_leaves / 32 coalesced rounds (60 us).  These rays are
// rare (C5: |d|^2 > 16 for 2 in 10^4 queries).  Same leaf tests, winner by (t, id): exactly hitScene's result.
#ifndef RT_BVH_BRUTE_A
#define RT_BVH_BRUTE_A 16.0f
#endif
template <class A, bool COUNT>
static __device__ __noinline__ void warp_brute_force(const float4* __restrict__ leaves, uint32_t n_leaves, int src, float ox, float oy, float oz,
                                                     float dx, float dy, float dz, Hit* out, WorkCount* wc) {
    const uint32_t lane = threadIdx.x & 31u;
    const V3 o = mk(__shfl_sync(RT_FULL, ox, src), __shfl_sync(RT_FULL, oy, src), __shfl_sync(RT_FULL, oz, src));
    const V3 d = mk(__shfl_sync(RT_FULL, dx, src), __shfl_sync(RT_FULL, dy, src), __shfl_sync(RT_FULL, dz, src));
    Hit b; b.t = RT_T_MAX; b.id = kIdMiss; b.bu = 0.0f; b.bv = 0.0f; b.aux = 0u;
    WorkCount dummy; dummy.nodes = dummy.sph = dummy.len = dummy.tri = dummy.flared_nodes = dummy.flared_queries = 0u;
    for (uint32_t i = lane; i < n_leaves; i += 32u) bvh_leaf<A, COUNT>(leaves + 4 * (size_t)i, o, d, b, COUNT ? wc : &dummy);
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
        Hit q;
        q.t = __shfl_xor_sync(RT_FULL, b.t, off); q.id = __shfl_xor_sync(RT_FULL, b.id, off);
        q.bu = __shfl_xor_sync(RT_FULL, b.bu, off); q.bv = __shfl_xor_sync(RT_FULL, b.bv, off); q.aux = __shfl_xor_sync(RT_FULL, b.aux, off);
        if (closer(q.t, q.id, b)) b = q;
    }
    if ((int)lane == src && closer(b.t, b.id, *out)) *out = b;      // *out holds the plane hit (planes are not in the tree)
}

// LBVH flavour of trace_tail_body: lanes advance ONE node per trip of the traversal loop (see trace_body_bvh); a lane
// whose traversal finished waits until `regen` lanes wait, then the waiting lanes close their path on a miss / light,
// or shade the hit and start the bounced ray; finished paths restart from the stored head.
//   * Box test per node: the distance-proportional inflation (slab<true>) is needed only where a sphere or a lens can be
//     hit (BvhNode::flags, set by the builder); the fast arm tests triangle-only nodes with the plain slab.  The strict arm
//     keeps the inflated test everywhere (its equality with brute force is what tests/test_gpu_bvh.py pins).
//   * Leaves: a trip tests at most two leaves per lane through ONE copy of the leaf code (lanes whose first or second
//     child is a leaf run it together).
//   * Traversal stack: in local memory (L1).  RT_BVH_SSTACK > 0 keeps the first levels in shared memory ([level][thread],
//     conflict free) — measured 4 % SLOWER on C4 and C5 (r02a probes: 5.25 vs 5.49 / 3.39 vs 3.60 Grays/s), so it is off.
#ifndef RT_BVH_SSTACK
#define RT_BVH_SSTACK 0
#endif
#ifndef RT_BVH_FLAG_SLAB
#define RT_BVH_FLAG_SLAB 1
#endif
#ifndef RT_BVH_BRUTE_MAX_LEAVES
#define RT_BVH_BRUTE_MAX_LEAVES 262144u   // a scan of more leaves than this costs the warp more than the walk costs the lane
#endif
#ifndef RT_BVH_BRUTE_SQA
#define RT_BVH_BRUTE_SQA 4.0f       // sqrt(RT_BVH_BRUTE_A)
#endif
#ifndef RT_BVH_FLARE_COST
#define RT_BVH_FLARE_COST 96       // flared lanes step when (their number) x (trips they waited) reaches this
#endif
template <class A, bool COUNT>
__device__ __forceinline__ void trace_tail_bvh_body(const TraceLaunch& p) {
    if (blockIdx.x >= p.tail_blocks) { terminal_body<A>(p, (blockIdx.x - p.tail_blocks) * kTraceThreads + threadIdx.x); return; }
    __shared__ float4 hs_all[4 * kTraceThreads];
    float4* hs = hs_all + threadIdx.x;
#if RT_BVH_SSTACK > 0
    __shared__ int sstack_all[RT_BVH_SSTACK * kTraceThreads];
    int* sst = sstack_all + threadIdx.x;
    int lst[kBvhStack - RT_BVH_SSTACK];
#else
    int* sst = nullptr;
    int lst[kBvhStack];
#endif
    const float4* S = p.sc.stream;
    const uint32_t lane = threadIdx.x & 31u;
    ShadeEnv env; env.sc = &p.sc; env.tex = &p.tex; env.table = p.table; env.rng_mode = p.rng_mode;
    // Work items.  One lane renders all samples of a pixel in sample order — or, when the host asks for sample chunks
    // (TraceLaunch::chunks > 1, fast arm + linear sums only), chunk_spp of them: a pixel behind a chain of glass spheres costs
    // hundreds of times the average (its rays are flared and walk large parts of the tree), and with whole pixels as items a
    // few such pixels, each serial in one lane, set the duration of the launch (C5, 32 spp: 1.1 s for 0.2 s of work).  Item
    // q -> pixel q / chunks, chunk q % chunks: the lanes of a warp trace different samples of the same pixels.  Chunk sums go
    // to partial[c][j]; k_combine_chunks adds them in chunk order (deterministic).
    const uint32_t chunks = p.chunks > 1u ? p.chunks : 1u;
    const uint32_t n_items = p.head.n_stoch * chunks;
    uint32_t k_end = p.first + p.n;
    // Persistent lanes over a queue of stochastic pixels: path lengths in large scenes are heavy tailed (mirror and
    // glass chains up to the depth limit), so with one pixel per lane most lanes of a warp idled behind its slowest
    // pixel (ncu: 12.5 of 32 lanes traversing).  A lane renders all samples of its pixel in sample order (the
    // accumulation order stays the reference's) and then takes the next list entry from a global counter.
    PixelRng pr; pr.vec_base = pr.uni_base = pr.hp = 0u;
    float4* px = nullptr;
    float4 acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    uint32_t k = k_end, k_first = k_end, i = 0u, mat_id = 0u;
    V3 o = mk(0.0f, 0.0f, 0.0f), d = mk(0.0f, 0.0f, 1.0f), out = mk(0.0f, 0.0f, 0.0f);
    Hit h; h.t = 0.0f; h.id = kIdMiss; h.bu = h.bv = 0.0f; h.aux = 0u;
    auto restart = [&]() {
        const float4 r0 = hs[0], r1 = hs[kTraceThreads], r2 = hs[2 * kTraceThreads], r3 = hs[3 * kTraceThreads];
        o = mk(r0.x, r0.y, r0.z); d = mk(r1.x, r1.y, r1.z); out = mk(r2.x, r2.y, r2.z);
        h.t = r0.w; h.bu = r1.w; h.bv = r2.w;
        i = __float_as_uint(r3.x); h.id = __float_as_uint(r3.y); h.aux = __float_as_uint(r3.z); mat_id = __float_as_uint(r3.w);
    };
    uint32_t queries = 0u;
    WorkCount wc; wc.nodes = wc.sph = wc.len = wc.tri = wc.flared_nodes = wc.flared_queries = 0u;
    const uint32_t regen = p.bvh_regen ? p.bvh_regen : 8u;
    V3 inv = mk(0.0f, 0.0f, 0.0f);
    RaySlack rs; rs.slack = RT_SLAB_SLACK; rs.sqa = 0.0f; rs.kappa = 1.0f;
    int sp = 0, node = 0;
    bool traversing = false, has_ray = false, have_result = false, exhausted = false;
    uint32_t flare_age = 0u;                                 // warp uniform: trips since the flared lanes last stepped
    auto push = [&](int v) {
        if (sp < RT_BVH_SSTACK) sst[sp * kTraceThreads] = v;
        else if (sp < kBvhStack) lst[sp - RT_BVH_SSTACK] = v;
        else return;
        ++sp;
    };
    auto pop = [&]() { --sp; return sp < RT_BVH_SSTACK ? sst[sp * kTraceThreads] : lst[sp - RT_BVH_SSTACK]; };
    auto start_ray = [&]() {
        h.t = RT_T_MAX; h.id = kIdMiss; h.bu = 0.0f; h.bv = 0.0f; h.aux = 0u;
        hit_planes<A>(p.sc, S, o, d, h);
        sp = 0; node = p.bvh.root; traversing = p.bvh.n_leaves > 0u;
        if (traversing && node < 0) { bvh_leaf<A, COUNT>(p.bvh.leaves + 4 * (~node), o, d, h, &wc); traversing = false; }
        inv = mk(1.0f / d.x, 1.0f / d.y, 1.0f / d.z);
        rs = ray_slack(d, p.sc.n_len == 0u);
        // a direction with an infinite or NaN component hits nothing finite: every sphere / lens root and every Moller-Trumbore
        // parameter of such a ray is infinite or NaN and fails the range test (cl:127) — no need to walk the tree for it
        if (!(fabsf(d.x) + fabsf(d.y) + fabsf(d.z) < 3.0e38f)) traversing = false;
        if (COUNT && rs.sqa != 0.0f) ++wc.flared_queries;
        have_result = true;
    };

    for (;;) {
        const bool need = !exhausted && !has_ray;            // ---- lanes without a pixel: flush, take the next one
        const unsigned nm = __ballot_sync(RT_FULL, need);
        if (nm) {
            if (need && px) *px = acc;
            const int leader = __ffs((int)nm) - 1;
            uint32_t base = 0u;
            if ((int)lane == leader) base = atomicAdd(p.queue, (uint32_t)__popc(nm));
            base = __shfl_sync(RT_FULL, base, leader);
            if (need) {
                const uint32_t q = base + (uint32_t)__popc(nm & ((1u << lane) - 1u));
                if (q < n_items) {
                    const uint32_t j = chunks > 1u ? q / chunks : q, c = q - j * chunks;
                    const uint32_t slot = __ldg(p.head.list + j);
                    int x, y;
                    slot_pixel(slot, p.W, x, y);
#pragma unroll
                    for (uint32_t r = 0; r < 4u; ++r) hs[r * kTraceThreads] = __ldg(p.head.rec + r * p.head.n_slots + slot);
                    pr = make_pixel_rng((uint32_t)x, (uint32_t)y, (uint32_t)p.W, p.seed);
                    if (chunks > 1u) {
                        px = p.partial + ((size_t)c * p.head.n_stoch + j);
                        acc = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                        k_first = min(p.first + c * p.chunk_spp, p.first + p.n);
                        k_end = min(k_first + p.chunk_spp, p.first + p.n);
                    } else {
                        px = p.target + (size_t)y * p.W + x;
                        acc = (p.accum_mode == 1u || p.first > 0u) ? *px : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
                        k_first = p.first;
                    }
                    k = k_first; has_ray = k < k_end; have_result = false; traversing = false;
                    restart();
                } else {
                    exhausted = true; px = nullptr;
                }
            }
        }
        if (!__any_sync(RT_FULL, has_ray)) break;
        if (has_ray && !traversing) {                        // ---- waiting lanes: close, shade, bounce, restart
            if (have_result) {
                ++queries;
                bool ended = false;
                if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); ended = true; }      // cl:450-452
                else {
                    mat_id = hit_material(p.sc, S, h);
                    if (mat_type(mat_id) == 5) {                                        // light, cl:459-462
                        const float4 m1 = __ldg(p.sc.mat + 2 * mat_index(mat_id) + 1);
                        out = vmin(out, mk(m1.x, m1.y, m1.z));
                        ended = true;
                    }
                }
                if (ended) { accumulate_sample<A>(acc, out, k, p.accum_mode); ++k; has_ray = k < k_end; restart(); }
            }
            while (has_ray) {
                shade<A>(env, S, h, mat_id, pr, o, d, out, k, i);
                if (++i < p.depth) { start_ray(); break; }
                accumulate_sample<A>(acc, out, k, p.accum_mode);                         // cl:447: depth exhausted, keep out
                ++k; has_ray = k < k_end; restart();
            }
        }
        {                                                    // ---- rays the hierarchy cannot help: the warp scans the leaves
            unsigned bm = __ballot_sync(RT_FULL, traversing && (rs.sqa < 0.0f || (rs.sqa > RT_BVH_BRUTE_SQA && p.bvh.n_leaves <= RT_BVH_BRUTE_MAX_LEAVES)));
            while (bm) {
                const int src = __ffs((int)bm) - 1;
                bm &= bm - 1u;
                warp_brute_force<A, COUNT>(p.bvh.leaves, p.bvh.n_leaves, src, o.x, o.y, o.z, d.x, d.y, d.z, &h, &wc);
                if ((int)lane == src) traversing = false;
            }
        }
        // warp-uniform switches of the traversal loop (rays only change in the block above): does any lane carry a wide flare
        // (step rule, slab_flared_core), a thin one (rho / kappa arithmetic)?  Most warps of most scenes carry neither.
        const bool warp_flared = __any_sync(RT_FULL, has_ray && rs.sqa != 0.0f), warp_thin = __any_sync(RT_FULL, has_ray && rs.kappa != 1.0f);
        for (;;) {                                           // ---- traversal: one node per trip
            // Wide-flared rays (rs.sqa != 0, slab_flared_core) need three to four times the instructions of an ordinary box test,
            // and a warp pays them in full however few of its lanes are flared.  So flared lanes do not step every trip: they keep
            // their place in the tree and wait until their step is worth its price — enough of them (count x trips waited >=
            // RT_BVH_FLARE_COST, or 12 lanes), or nothing else left to advance.  Waiting is cheap (an idle lane of a warp that goes
            // on), a step for one lane is not.
            bool step = traversing;
            if (warp_flared) {
                const uint32_t tf = (uint32_t)__popc(__ballot_sync(RT_FULL, traversing && rs.sqa != 0.0f));
                const bool any_n = __any_sync(RT_FULL, traversing && rs.sqa == 0.0f);
                const bool do_f = tf != 0u && (!any_n || tf >= 12u || tf * flare_age >= (uint32_t)RT_BVH_FLARE_COST);
                flare_age = (do_f || tf == 0u) ? 0u : flare_age + 1u;
                step = traversing && (rs.sqa == 0.0f || do_f);
            }
            if (step) {
                const bool flare_mode = rs.sqa != 0.0f;
                const float4* N = reinterpret_cast<const float4*>(p.bvh.nodes + node);
                float4 n0, n1, n2, n3;
                load64(N, n0, n1, n2, n3);
                const float lo0[3] = {n0.x, n0.y, n0.z}, hi0[3] = {n1.x, n1.y, n1.z}, lo1[3] = {n2.x, n2.y, n2.z}, hi1[3] = {n3.x, n3.y, n3.z};
                const int c0 = __float_as_int(n0.w), c1 = __float_as_int(n1.w);
                float t0, t1;
                bool h0, h1;
                if (COUNT) { ++wc.nodes; if (flare_mode) ++wc.flared_nodes; }
                const uint32_t fl = __float_as_uint(n2.w);
                const float2 rho = node_rho(n3.w);
                if (flare_mode && fl != 0u) {
                    if (rs.sqa > 0.0f) {
                        h0 = slab_flared_core(lo0, hi0, o, d, inv, h.t, rs.slack, rs.sqa, (fl & 1u) != 0u, p.sc.n_tri != 0u, t0);
                        h1 = slab_flared_core(lo1, hi1, o, d, inv, h.t, rs.slack, rs.sqa, (fl & 2u) != 0u, p.sc.n_tri != 0u, t1);
                    } else {
                        h0 = slab_robust(lo0, hi0, o, d, inv, h.t, t0, rs, (fl & 1u) != 0u, rho.x); h1 = slab_robust(lo1, hi1, o, d, inv, h.t, t1, rs, (fl & 2u) != 0u, rho.y);
                    }
                } else {
                    // ONE instruction stream for every lane: a child whose subtree holds triangles only (flag clear; fast arm) gets
                    // slack 0 — the plain slab test, exact for Moller-Trumbore in any parametrisation — instead of a branch of its
                    // own: with two branches the common half of the test was issued twice, at 13.5 lanes each (ncu, C4).
                    const bool use_flags = RT_BVH_FLAG_SLAB && !A::kExact;
                    const bool cv0 = !use_flags || (fl & 1u) != 0u, cv1 = !use_flags || (fl & 2u) != 0u;
                    h0 = slab<true>(lo0, hi0, o, inv, h.t, t0, cv0 ? rs.slack : 0.0f, cv0 ? rs.kappa : 1.0f, warp_thin, rho.x);
                    h1 = slab<true>(lo1, hi1, o, inv, h.t, t1, cv1 ? rs.slack : 0.0f, cv1 ? rs.kappa : 1.0f, warp_thin, rho.y);
                }
                // children that passed, as one integer each: < 0 a leaf (~slot), > 0 an internal node, 0 none (node 0 is the root,
                // never a child) — min / max instead of pairs of predicates
                const int k0 = h0 ? c0 : 0, k1 = h1 ? c1 : 0;
                int l0 = min(k0, 0), l1 = min(k1, 0);
                if (l0 == 0) { l0 = l1; l1 = 0; }
                while (l0 != 0) {
                    bvh_leaf<A, COUNT>(p.bvh.leaves + 4 * (~l0), o, d, h, &wc);
                    l0 = l1; l1 = 0;
                }
                // the sibling of a leaf that was just hit may now lie behind the hit (t_near is a lower bound of its hits)
                const int i0 = t0 <= h.t ? max(k0, 0) : 0, i1 = t1 <= h.t ? max(k1, 0) : 0;
                if (i0 != 0 && i1 != 0) {
                    const bool first0 = t0 <= t1;
                    push(first0 ? i1 : i0);
                    node = first0 ? i0 : i1;
                } else if ((i0 | i1) != 0) node = i0 | i1;
                else if (sp > 0) node = pop();
                else traversing = false;
            }
            const unsigned trav = __ballot_sync(RT_FULL, traversing);
            const unsigned waiting = __ballot_sync(RT_FULL, !traversing && !exhausted);   // finished a ray, or needs a pixel
            if (trav == 0u || (uint32_t)__popc(waiting) >= regen) break;
        }
    }
    queries = __reduce_add_sync(RT_FULL, queries);
    if (lane == 0u && queries) atomicAdd(p.queries, (unsigned long long)queries);
    if (COUNT) {
        const uint32_t a = __reduce_add_sync(RT_FULL, wc.nodes), b = __reduce_add_sync(RT_FULL, wc.sph),
                       c = __reduce_add_sync(RT_FULL, wc.len), e = __reduce_add_sync(RT_FULL, wc.tri),
                       fn = __reduce_add_sync(RT_FULL, wc.flared_nodes), fq = __reduce_add_sync(RT_FULL, wc.flared_queries);
        if (lane == 0u && p.work) {
            atomicAdd(p.work + 0, (unsigned long long)a); atomicAdd(p.work + 1, (unsigned long long)b);
            atomicAdd(p.work + 2, (unsigned long long)c); atomicAdd(p.work + 3, (unsigned long long)e);
            atomicAdd(p.work + 6, (unsigned long long)fn); atomicAdd(p.work + 7, (unsigned long long)fq);
        }
    }
}

// Nearest hit of arbitrary rays (rays6 != nullptr) or of the primary rays of the frame; same code path as the
// render loop (warp-cooperative meshes, optional certification of primary hits).
template <class A, bool BVH>
__device__ __forceinline__ void probe_hits_body(const ProbeLaunch& p, float4* smem_) {
    const float4* smem = p.sc.stream;
    uint8_t* list = nullptr;
    if (!BVH) { stage_scene(p.sc, smem_); smem = smem_; list = warp_list(p.sc, smem_); }
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < p.n;
    V3 o = mk(0.0f, 0.0f, 0.0f), d = mk(0.0f, 0.0f, 1.0f);
    if (active) {
        if (p.rays6) {
            o = mk(p.rays6[6 * i], p.rays6[6 * i + 1], p.rays6[6 * i + 2]);
            d = mk(p.rays6[6 * i + 3], p.rays6[6 * i + 4], p.rays6[6 * i + 5]);
        } else {
            o = mk(p.cam[0], p.cam[1], p.cam[2]);
            d = primary_dir<Exact>(p.cam, (int)(i % (uint32_t)p.W), (int)(i / (uint32_t)p.W), p.W, p.H);
        }
    }
    Hit h;
    if (BVH) { if (active) nearest_hit_bvh<A>(p.sc, smem, p.bvh, o, d, h); }
    else nearest_hit_warp<A>(p.sc, smem, list, o, d, active, p.mesh_mode, h);
    if (!active) return;
    if (!A::kExact && p.certify_id && h.id != p.certify_id[i])
        h = BVH ? nearest_hit_exact_cold_bvh(p.sc, smem, p.bvh, o.x, o.y, o.z, d.x, d.y, d.z)
                : nearest_hit_exact_cold(p.sc, smem, o.x, o.y, o.z, d.x, d.y, d.z, p.mesh_mode);
    if (p.out_id) p.out_id[i] = h.id;
    if (p.out_t) p.out_t[i] = h.t;
}

// One iteration of the path loop per ray (single-bounce differential test).
template <class A, bool BVH>
__device__ __forceinline__ void probe_bounce_body(const ProbeLaunch& p, float4* smem_) {
    const float4* smem = p.sc.stream;
    uint8_t* list = nullptr;
    if (!BVH) { stage_scene(p.sc, smem_); smem = smem_; list = warp_list(p.sc, smem_); }
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool active = i < p.n;
    V3 o = mk(0.0f, 0.0f, 0.0f), d = mk(0.0f, 0.0f, 1.0f);
    if (active) {
        o = mk(p.rays6[6 * i], p.rays6[6 * i + 1], p.rays6[6 * i + 2]);
        d = mk(p.rays6[6 * i + 3], p.rays6[6 * i + 4], p.rays6[6 * i + 5]);
    }
    Hit h;
    if (BVH) { if (active) nearest_hit_bvh<A>(p.sc, smem, p.bvh, o, d, h); }
    else nearest_hit_warp<A>(p.sc, smem, list, o, d, active, p.mesh_mode, h);
    if (!active) return;
    V3 out = mk(1.0f, 1.0f, 1.0f);
    const PixelRng pr = make_pixel_rng(p.gx[i], p.gy[i], (uint32_t)p.W, p.seed);
    ShadeEnv env; env.sc = &p.sc; env.tex = &p.tex; env.table = p.table; env.rng_mode = p.rng_mode;
    bool ended;
    if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); ended = true; }
    else ended = shade<A>(env, smem, h, hit_material(p.sc, smem, h), pr, o, d, out, p.sample[i], p.bounce[i]);
    p.rays6_out[6 * i] = o.x; p.rays6_out[6 * i + 1] = o.y; p.rays6_out[6 * i + 2] = o.z;
    p.rays6_out[6 * i + 3] = d.x; p.rays6_out[6 * i + 4] = d.y; p.rays6_out[6 * i + 5] = d.z;
    p.col3[3 * i] = out.x; p.col3[3 * i + 1] = out.y; p.col3[3 * i + 2] = out.z;
    p.ended[i] = ended ? 1 : 0;
}

inline size_t staged_bytes(const DevScene& sc) {
    return ((size_t)sc.stream_f4 + 3u * sc.n_mesh + ((sc.n_tri + 3u) >> 2)) * sizeof(float4) + (kTraceThreads / 32) * 1024u;
}

// grid of k_trace_tail_*: blocks of stochastic slots, then blocks over the rest of the list (terminal pixels)
inline size_t tail_work_items(const TraceLaunch& p) {
    const size_t T = kTraceThreads, blocks = ((size_t)p.head.n_stoch + T - 1) / T * (p.chunks > 1u ? p.chunks : 1u);
    return (blocks + (p.head.n_slots - p.head.n_stoch + T - 1) / T) * T;
}
// k_trace_tail_*_bvh: tail_blocks persistent blocks over the queue of stochastic pixels + one block per 128 entries of
// the rest of the list (terminal pixels)
template <class K>
inline cudaError_t launch_tail_bvh(K kernel, TraceLaunch p, cudaStream_t s) {
    int dev = 0, sms = 0, per_sm = 0;
    cudaError_t e;
    if ((e = cudaGetDevice(&dev)) != cudaSuccess) return e;
    if ((e = cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev)) != cudaSuccess) return e;
    if ((e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, kTraceThreads, 0)) != cudaSuccess) return e;
    const uint32_t T = kTraceThreads;
    const uint64_t items = (uint64_t)p.head.n_stoch * (p.chunks > 1u ? p.chunks : 1u);
    const uint32_t cap = (uint32_t)(sms * per_sm), want = (uint32_t)std::min<uint64_t>((items + T - 1u) / T, cap);
    p.tail_blocks = want < cap ? want : cap;
    const uint32_t rest_blocks = (p.head.n_slots - p.head.n_stoch + T - 1u) / T;
    if (p.tail_blocks + rest_blocks == 0u) return cudaSuccess;
    kernel<<<p.tail_blocks + rest_blocks, T, 0, s>>>(p);
    return cudaGetLastError();
}
constexpr size_t kHeadStateBytes = 4u * kTraceThreads * sizeof(float4);   // restart state of k_trace_tail_*

template <class K, class P>
inline cudaError_t launch_staged(K kernel, const P& p, size_t work_items, cudaStream_t s, bool bvh = false, size_t extra = 0) {
    const size_t smem = bvh ? 0 : staged_bytes(p.sc) + extra;
    if (smem > kMaxBlockSmemBytes) return cudaErrorInvalidValue;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    const unsigned blocks = (unsigned)((work_items + kTraceThreads - 1) / kTraceThreads);
    if (blocks == 0) return cudaSuccess;
    kernel<<<blocks, kTraceThreads, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace rtb
