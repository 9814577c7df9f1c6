// csrc/rt_kernels.cuh — the path-tracing kernels (sm_100a).  Included by exactly two translation units:
//   rt_kernels_strict.cu  (RT_STRICT 1; nvcc -fmad=false -prec-div=true -prec-sqrt=true)
//       every float operation is one IEEE binary32 operation in the oracle's order -> bit-exact images
//   rt_kernels_fast.cu    (RT_STRICT 0; nvcc -fmad=true -prec-div=false -prec-sqrt=false)
//       FMA contraction + MUFU approximations -> the throughput arm (statistical parity, identical primary ids)
//
// Behaviour follows /root/reference/kernels/raytracer.cl ("cl:" below); the structure does not:
//   * one launch renders MANY samples of a pixel; a lane that finishes a path immediately regenerates the
//     next sample's primary ray (persistent lanes), so the brute-force primitive loop — 95 % of the work —
//     always runs with the warp converged;
//   * primitives are read from a shared-memory copy of a float4 "test stream" with warp-uniform addresses;
//   * triangles carry host-precomputed edges and unit normal, and the reference's LAST test (front facing,
//     cl:295) runs FIRST — a pure reordering of side-effect-free tests, so results are unchanged;
//   * hit-point data (position, normal, uv, material) is derived once, for the winning primitive only.
#pragma once
#include "rt_device.h"

namespace rtb {
namespace RT_FLAVOUR {

#define RT_T_MIN 0.001f        // cl:3
#define RT_T_MAX 1000.0f       // cl:4
#define RT_TRI_EPS 0.0000001f  // cl:1
#define RT_TABLE_N 100000u     // cl:7

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
__device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
__device__ __forceinline__ float min_cl(float a, float b) { return (b < a) ? b : a; }   // OpenCL min(x, y)
__device__ __forceinline__ V3 vmin(V3 a, V3 b) { return mk(min_cl(a.x, b.x), min_cl(a.y, b.y), min_cl(a.z, b.z)); }

#if RT_STRICT
__device__ __forceinline__ float rcp_(float x) { return 1.0f / x; }
__device__ __forceinline__ float div_(float a, float b) { return a / b; }
__device__ __forceinline__ float sqrt_(float x) { return sqrtf(x); }
__device__ __forceinline__ V3 normalize(V3 v) { float inv = 1.0f / sqrtf(dot(v, v)); return v * inv; }
__device__ __forceinline__ V3 vdiv(V3 v, float s) { return mk(v.x / s, v.y / s, v.z / s); }
// glibc powf (the oracle) is correctly rounded in all but ~1e-8 of cases; so is (float)pow(double)
__device__ __forceinline__ float pow5(float x) { return (float)pow((double)x, 5.0); }
#else
__device__ __forceinline__ float rcp_(float x) { return __fdividef(1.0f, x); }
__device__ __forceinline__ float div_(float a, float b) { return __fdividef(a, b); }
__device__ __forceinline__ float sqrt_(float x) { return sqrtf(x); }
__device__ __forceinline__ V3 normalize(V3 v) { return v * rsqrtf(dot(v, v)); }
__device__ __forceinline__ V3 vdiv(V3 v, float s) { return v * __fdividef(1.0f, s); }
__device__ __forceinline__ float pow5(float x) { float x2 = x * x; return x2 * x2 * x; }
#endif

__device__ __forceinline__ bool in_range(float t) { return t >= RT_T_MIN && t <= RT_T_MAX; }  // == cl:127

struct Hit {
    float t;
    uint32_t id;     // kind << 28 | index (triangles: reference-global face number)
    float bu, bv;    // barycentrics (triangles)
    uint32_t aux;    // lens: side; triangle: index into the triangle stream
};

// Shared-memory view of the scene (built once per block by stage_scene).
struct SmemScene {
    const float4* f4;          // test stream
    const uint4* mesh;         // 3 x uint4 per MeshInst
};

__device__ __forceinline__ void stage_scene(const DevScene& sc, float4* smem) {
    const uint32_t n1 = sc.stream_f4, n2 = sc.n_mesh * 3u;
    const float4* m = reinterpret_cast<const float4*>(sc.mesh);
    for (uint32_t i = threadIdx.x; i < n1; i += blockDim.x) smem[i] = __ldg(sc.stream + i);
    for (uint32_t i = threadIdx.x; i < n2; i += blockDim.x) smem[n1 + i] = __ldg(m + i);
    __syncthreads();
}

// cl:322-360 hitScene, with hitModel (cl:305-320) and hitMeshOut (cl:291-303) folded into one ordered scan.
// Strict `<`: on equal t the earlier kind / lower index wins.
__device__ __forceinline__ void nearest_hit(const DevScene& sc, const float4* __restrict__ S, V3 o, V3 d,
                                            uint32_t mesh_mode, Hit& best) {
    best.t = RT_T_MAX; best.id = kIdMiss; best.bu = 0.0f; best.bv = 0.0f; best.aux = 0u;

    // ---- spheres, cl:149-174
    for (uint32_t i = 0; i < sc.n_sph; ++i) {
        const float4 s = S[i];
        const V3 oc = mk(s.x - o.x, s.y - o.y, s.z - o.z);
        const float b = dot(oc, d);
        const float c = dot(oc, oc) - s.w;
        const float dis = b * b - c;
        if (dis > 0.0f) {
            const float sq = sqrt_(dis);
            float t = b - sq;
            if (!in_range(t)) t = b + sq;
            if (t >= RT_T_MIN && t < best.t) { best.t = t; best.id = (kKindSphere << 28) | i; }
        }
    }
    // ---- planes, cl:176-194
    for (uint32_t i = 0; i < sc.n_pla; ++i) {
        const float4 p = S[sc.off_pla + 2 * i], n4 = S[sc.off_pla + 2 * i + 1];
        const V3 n = mk(n4.x, n4.y, n4.z);
        const float a = dot(d, n);
        const float b = dot(mk(p.x - o.x, p.y - o.y, p.z - o.z), n);
        const float t = div_(b, a);
        if (t >= RT_T_MIN && t < best.t) { best.t = t; best.id = (kKindPlane << 28) | i; }
    }
    // ---- lenses, cl:196-255
    for (uint32_t i = 0; i < sc.n_len; ++i) {
        const float4 l1 = S[sc.off_len + 2 * i], l2 = S[sc.off_len + 2 * i + 1];
        V3 oc = mk(l1.x - o.x, l1.y - o.y, l1.z - o.z);
        const float b1 = dot(oc, d);
        const float dis1 = b1 * b1 - (dot(oc, oc) - l1.w);
        oc = mk(l2.x - o.x, l2.y - o.y, l2.z - o.z);
        const float b2 = dot(oc, d);
        const float dis2 = b2 * b2 - (dot(oc, oc) - l2.w);
        if (dis1 > 0.0f && dis2 > 0.0f) {
            const float d1 = sqrt_(dis1), d2 = sqrt_(dis2);
            const float t1A = b1 - d1, t1B = b1 + d1, t2A = b2 - d2, t2B = b2 + d2;
            float t = -1.0f;
            uint32_t side = 0u;
            if ((t1B < t2A) || (t2B < t1A)) { /* misses the intersection of the two balls */ }
            else if (RT_T_MIN <= t1A || RT_T_MIN <= t2A) {
                if (t2A <= t1A) { side = 0u; t = t1A; } else { side = 1u; t = t2A; }
            } else if (RT_T_MIN <= t1B && RT_T_MIN <= t2B) {
                if (t1B <= t2B) { side = 0u; t = t1B; } else { side = 1u; t = t2B; }
            }
            // every accepted branch yields t >= T_MIN; t <= T_MAX (cl:245) is implied by t < best.t
            if (t >= RT_T_MIN && t < best.t) { best.t = t; best.id = (kKindLens << 28) | i; best.aux = side; }
        }
    }
    // ---- mesh instances, cl:257-320
    const uint4* M = reinterpret_cast<const uint4*>(S + sc.stream_f4);
    for (uint32_t m = 0; m < sc.n_mesh; ++m) {
        const uint4 mi = M[3 * m];   // tri_first, tri_count, face_id_base, mat_id
        bool open = true;            // FIRST_FRONT: no front-facing hit found yet in this mesh
        for (uint32_t k = 0; k < mi.y; ++k) {
            const uint32_t tri = mi.x + k;
            const float4 q0 = S[sc.off_tri + 3 * tri], q1 = S[sc.off_tri + 3 * tri + 1], q2 = S[sc.off_tri + 3 * tri + 2];
            const V3 nrm = mk(q2.y, q2.z, q2.w);
            if (open && dot(nrm, d) < 0.0f) {                       // cl:295 evaluated first
                const V3 e1 = mk(q0.w, q1.x, q1.y), e2 = mk(q1.z, q1.w, q2.x);
                const V3 h = cross(d, e2);
                const float a = dot(e1, h);
                if (!(a > -RT_TRI_EPS && a < RT_TRI_EPS)) {
                    const float f = rcp_(a);
                    const V3 sv = mk(o.x - q0.x, o.y - q0.y, o.z - q0.z);
                    const float u = f * dot(sv, h);
                    if (!(u < 0.0f || u > 1.0f)) {
                        const V3 qv = cross(sv, e1);
                        const float v = f * dot(d, qv);
                        if (!(v < 0.0f || u + v > 1.0f)) {
                            const float t = f * dot(e2, qv);
                            if (in_range(t)) {
                                if (mesh_mode == 0u) open = false;   // first front-facing hit ends this mesh
                                if (t < best.t) {
                                    best.t = t; best.id = (kKindTriangle << 28) | (mi.z + k);
                                    best.bu = u; best.bv = v; best.aux = tri;
                                }
                            }
                        }
                    }
                }
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ RNG
// Table mode, cl:113-125: index = (hash(dir) + k) % 100000, hash = (uint)fabs(double(dot(dir, C)) * 438.5453).
// The 64-bit sum of the reference is reduced modulo 100000 term by term (same residue, 32-bit arithmetic).
__device__ __forceinline__ uint32_t dir_hash(V3 d) {
    const float f = d.x * 123.9898f + d.y * 348.233f + d.z * 433.3314f;
    return (uint32_t)fabs((double)f * 438.5453);
}
struct PixelRng {
    uint32_t gx, gy;
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
    r.gx = gx; r.gy = gy;
    r.vec_base = (gx * 3931u + gy * 2504u) % RT_TABLE_N;
    r.uni_base = (gx * 3931u + gy) % RT_TABLE_N;
    r.hp = mix32(seed ^ mix32((gy * W + gx) ^ 0xA511E9B3u));
    return r;
}
__device__ __forceinline__ V3 table_vec(const float* __restrict__ tab, const PixelRng& pr, V3 d, uint32_t s_seed) {
    const uint32_t k = ((s_seed * 2683u) % RT_TABLE_N + pr.vec_base) % RT_TABLE_N;
    const uint32_t seed = (dir_hash(d) % RT_TABLE_N + (k * 3u) % RT_TABLE_N) % RT_TABLE_N;
    return mk(__ldg(tab + seed), __ldg(tab + seed + 1), __ldg(tab + seed + 2));
}
__device__ __forceinline__ float table_uniform(const float* __restrict__ tab, const PixelRng& pr, V3 d, uint32_t s_seed) {
    const uint32_t k = ((s_seed * 2683u) % RT_TABLE_N + pr.uni_base) % RT_TABLE_N;
    const uint32_t seed = 3u * RT_TABLE_N + (dir_hash(d) % RT_TABLE_N + k) % RT_TABLE_N;
    return __ldg(tab + seed);
}
// Counter mode: key(pixel) -> key(sample) -> key(bounce) -> draw j, all lowbias32 mixes.
__device__ __forceinline__ uint32_t bounce_key(uint32_t hp, uint32_t sample, uint32_t bounce) {
    const uint32_t hs = mix32(hp ^ (sample * 0x9E3779B9u));
    return mix32(hs + bounce * 0x85EBCA6Bu + 0x1B873593u);
}
__device__ __forceinline__ uint32_t draw(uint32_t hb, uint32_t j) { return mix32(hb + j * 0xC2B2AE35u); }
__device__ __forceinline__ float u01(uint32_t h) { return (float)(h >> 8) * 5.9604644775390625e-8f; }
__device__ __forceinline__ float s11(uint32_t h) { return (float)(h >> 8) * 1.1920928955078125e-7f - 1.0f; }
// Uniform in the unit ball from exactly-rounded operations: Marsaglia disc rejection + max of 3 uniforms.
__device__ __forceinline__ V3 rng_ball(uint32_t hb, uint32_t base) {
    float u = 0.0f, v = 0.0f, s = 0.0f;
    for (uint32_t k = 0; k < 16u; ++k) {
        u = s11(draw(hb, base + 2u * k));
        v = s11(draw(hb, base + 2u * k + 1u));
        s = u * u + v * v;
        if (s < 1.0f) break;
        u = v = s = 0.0f;
    }
    const float q = sqrt_(1.0f - s);
    const float dx = (2.0f * u) * q, dy = (2.0f * v) * q, dz = 1.0f - 2.0f * s;
    const float r = fmaxf(fmaxf(u01(draw(hb, base + 32u)), u01(draw(hb, base + 33u))), u01(draw(hb, base + 34u)));
    return mk(dx * r, dy * r, dz * r);
}
// Same law as the reference's misaligned table read (cl:109-117): (a.x,a.y,a.z) | (a.y,a.z,b.x) | (a.z,b.x,b.y).
__device__ __forceinline__ V3 counter_vec(uint32_t hb) {
    const uint32_t sel = __umulhi(draw(hb, 0u), 3u);
    const V3 a = rng_ball(hb, 1u), b = rng_ball(hb, 64u);
    return sel == 0u ? a : (sel == 1u ? mk(a.y, a.z, b.x) : mk(a.z, b.x, b.y));
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
__device__ __forceinline__ V3 texture_rgb(const DevTexture& t, float s, float tt, uint32_t layer) {
    const float u = s * (float)t.w, v = tt * (float)t.h;
    const float fu = floorf(u - 0.5f), fv = floorf(v - 0.5f);
    const float a = (u - 0.5f) - fu, b = (v - 0.5f) - fv;
    const int i0 = (int)fu, j0 = (int)fv;
    int l = (int)rintf((float)layer);
    l = max(0, min(l, t.layers - 1));
    const float4 t00 = texel(t, i0, j0, l), t10 = texel(t, i0 + 1, j0, l);
    const float4 t01 = texel(t, i0, j0 + 1, l), t11 = texel(t, i0 + 1, j0 + 1, l);
    const float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return mk(w00 * t00.x + w10 * t10.x + w01 * t01.x + w11 * t11.x,
              w00 * t00.y + w10 * t10.y + w01 * t01.y + w11 * t11.y,
              w00 * t00.z + w10 * t10.z + w01 * t01.z + w11 * t11.z);
}

// ------------------------------------------------------------------------------------------------ shading
struct ShadeEnv {
    const DevScene* sc;
    const DevTexture* tex;
    const float* table;
    uint32_t rng_mode;
};

__device__ __forceinline__ V3 reflect_dir(V3 d, V3 n) {            // cl:362-365
    const float k = 2.0f * dot(d, n);
    return normalize(d - n * k);
}

// Hit-point data for the winning primitive (cl:158-161, 185-188, 246-249, 281-285, 296, 314) followed by one
// iteration of getCol's switch (cl:454-479).  Returns true when the path ends (light).
__device__ __forceinline__ bool shade(const ShadeEnv& env, const float4* __restrict__ S, const Hit& h, const PixelRng& pr,
                                      V3& o, V3& d, V3& out, uint32_t sample, uint32_t bounce) {
    const DevScene& sc = *env.sc;
    const uint32_t kind = h.id >> 28, idx = h.id & 0x0FFFFFFFu;
    const V3 p = o + d * h.t;                                       // cl:141-143
    V3 n;
    uint32_t mat_id, tex_id = 0u;
    float uvx = 0.0f, uvy = 0.0f;
    if (kind == kKindSphere) {
        const float4 s = __ldg(sc.sph_shade + idx);
        n = vdiv(p - mk(s.x, s.y, s.z), s.w);
        mat_id = __ldg(sc.sph_mat + idx);
    } else if (kind == kKindPlane) {
        const float4 pn = __ldg(sc.pla_normal + idx);
        const V3 nn = mk(pn.x, pn.y, pn.z);
        const float a = dot(d, nn);
        const float sg = (a != a) ? 0.0f : (a > 0.0f ? 1.0f : (a < 0.0f ? -1.0f : a));
        n = neg(nn) * sg;                                           // cl:187
        mat_id = __ldg(sc.pla_mat + idx);
    } else if (kind == kKindLens) {
        const float4 l = __ldg(sc.len_shade + 2 * idx + h.aux);
        n = vdiv(p - mk(l.x, l.y, l.z), l.w);
        mat_id = __ldg(sc.len_mat + idx);
    } else {
        const float4 q2 = S[sc.off_tri + 3 * h.aux + 2];
        n = mk(q2.y, q2.z, q2.w);
        const uint4 mi0 = reinterpret_cast<const uint4*>(S + sc.stream_f4)[3 * __ldg(sc.tri_mesh + h.aux)];
        const uint4 mi1 = reinterpret_cast<const uint4*>(S + sc.stream_f4)[3 * __ldg(sc.tri_mesh + h.aux) + 1];
        mat_id = mi0.w;
        tex_id = mi1.x;
        const float2 ua = __ldg(sc.tri_uv + 3 * h.aux), ub = __ldg(sc.tri_uv + 3 * h.aux + 1), uc = __ldg(sc.tri_uv + 3 * h.aux + 2);
        const float w = 1.0f - h.bu - h.bv;                         // cl:101-103
        uvx = ua.x * w + ub.x * h.bu + uc.x * h.bv;
        uvy = ua.y * w + ub.y * h.bu + uc.y * h.bv;
    }
    const float4 m0 = __ldg(sc.mat + 2 * mat_id), m1 = __ldg(sc.mat + 2 * mat_id + 1);
    const int type = __float_as_int(m0.x);
    const float extra = m0.y;
    V3 color = mk(m1.x, m1.y, m1.z);
    const uint32_t s_seed = bounce + sample;                        // cl:456,472,476
    const uint32_t hb = (env.rng_mode == 1u) ? bounce_key(pr.hp, sample, bounce) : 0u;

    if (type == 3 || type == 4) {                                   // diffuse / textured: rayScatter cl:393-399
        const V3 rv = (env.rng_mode == 0u) ? table_vec(env.table, pr, d, s_seed) : counter_vec(hb);
        d = normalize(n + rv);
        o = p;
        out = out * extra;
        if (type == 4) color = texture_rgb(*env.tex, uvx, uvy, tex_id);   // cl:477
        out = vmin(out, color);
        return false;
    }
    if (type == 5) { out = vmin(out, color); return true; }         // light cl:459-462
    if (type == 1) {                                                // reflective cl:463-466
        o = p;
        d = reflect_dir(d, n);
        out = out * extra;
        out = vmin(out, color);
        return false;
    }
    if (type == 0 || type == 2) {                                   // refractive cl:369-391 / dielectric cl:407-435
        V3 nn; float ratio;
        float cai = dot(d, n);
        if (cai > 0.0f) { nn = neg(n); ratio = extra; cai = -cai; }
        else { nn = n; ratio = rcp_(extra); }
        bool try_refract = true;
        if (type == 2) {
            float r0 = div_(1.0f - ratio, 1.0f + ratio);            // schlick cl:401-405
            r0 *= r0;
            const float prob = r0 + (1.0f - r0) * pow5(1.0f - (-cai));
            const float u = (env.rng_mode == 0u) ? table_uniform(env.table, pr, d, s_seed) : u01(draw(hb, 0u));
            try_refract = prob < u;                                 // cl:423
        }
        const float disc = 1.0f - ratio * ratio * (1.0f - cai * cai);
        o = p;
        if (try_refract && disc > 0.0f) {
            const float k = ratio * cai + sqrt_(disc);
            d = d * ratio - nn * k;                                 // not normalised, cl:386/428
        } else {
            d = reflect_dir(d, nn);
        }
        out = vmin(out, color);
        return false;
    }
    return false;   // unknown type: the reference's switch has no default, the ray is left unchanged
}

// ------------------------------------------------------------------------------------------------ kernels
__device__ __forceinline__ V3 primary_dir(const float* cam, int x, int y, int W, int H) {   // cl:129-139, 500-501
    const float s = div_((float)x, (float)W), t = div_((float)y, (float)H);
    const V3 llc = mk(cam[3], cam[4], cam[5]), hor = mk(cam[6], cam[7], cam[8]), ver = mk(cam[9], cam[10], cam[11]);
    return normalize(llc + hor * s + ver * t);
}

// trace (cl:496-510) + retrace (cl:512-532) + getCol (cl:444-486) for samples [first, first+n) of every pixel.
// One lane owns one pixel; warps cover 8x4 pixel tiles.  Lanes are persistent over samples: the moment a
// path ends the lane accumulates it and restarts from the primary ray, so each trip of the while loop is one
// brute-force nearest-hit query executed by every lane that still has samples left.
extern "C" __global__ void __launch_bounds__(kTraceThreads)
RT_KERNEL_NAME(k_trace)(const __grid_constant__ TraceLaunch p) {
    extern __shared__ float4 smem[];
    stage_scene(p.sc, smem);
    const float4* S = smem;

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int tiles_x = (p.W + 7) >> 3;
    const int tile = blockIdx.x * (kTraceThreads / 32) + warp;
    const int x = (tile % tiles_x) * 8 + (lane & 7);
    const int y = (tile / tiles_x) * 4 + (lane >> 3);
    const bool live = (x < p.W) && (y < p.H);

    unsigned long long rays = 0ull;
    if (live) {
        const V3 cam_o = mk(p.cam[0], p.cam[1], p.cam[2]);
        const V3 d0 = primary_dir(p.cam, x, y, p.W, p.H);
        const PixelRng pr = make_pixel_rng((uint32_t)x, (uint32_t)y, (uint32_t)p.W, p.seed);
        ShadeEnv env; env.sc = &p.sc; env.tex = &p.tex; env.table = p.table; env.rng_mode = p.rng_mode;
        float4* px = p.target + (size_t)y * p.W + x;
        float4 acc = (p.accum_mode == 1u || p.first > 0u) ? *px : make_float4(0.0f, 0.0f, 0.0f, 0.0f);

        uint32_t k = p.first, i = 0u;
        const uint32_t k_end = p.first + p.n;
        V3 o = cam_o, d = d0, out = mk(1.0f, 1.0f, 1.0f);
        while (k < k_end) {
            Hit h;
            nearest_hit(p.sc, S, o, d, p.mesh_mode, h);
            ++rays;
            bool ended;
            if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); ended = true; }   // cl:450-452
            else ended = shade(env, S, h, pr, o, d, out, k, i);
            ++i;
            if (ended || i >= p.depth) {
                if (p.accum_mode == 1u) {                   // linear sum (what the sample split reduces)
                    acc.x += out.x; acc.y += out.y; acc.z += out.z; acc.w += 1.0f;
                } else if (k == 0u) {                        // cl:509
                    acc = make_float4(sqrt_(out.x), sqrt_(out.y), sqrt_(out.z), 1.0f);
                } else {                                     // cl:524-531
                    const float a = div_((float)k, (float)(k + 1u));
                    const float pr_ = acc.x * acc.x, pg_ = acc.y * acc.y, pb_ = acc.z * acc.z;
                    acc = make_float4(sqrt_(out.x + (pr_ - out.x) * a), sqrt_(out.y + (pg_ - out.y) * a),
                                      sqrt_(out.z + (pb_ - out.z) * a), 1.0f);
                }
                ++k; i = 0u; o = cam_o; d = d0; out = mk(1.0f, 1.0f, 1.0f);
            }
        }
        *px = acc;
    }
    // one atomic per warp
    for (int off = 16; off > 0; off >>= 1) rays += __shfl_down_sync(0xffffffffu, rays, off);
    if (lane == 0 && rays) atomicAdd(p.rays, rays);
}

// Nearest hit of arbitrary rays (rays6 != nullptr) or of the primary rays of the frame.
extern "C" __global__ void __launch_bounds__(kTraceThreads)
RT_KERNEL_NAME(k_probe_hits)(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    stage_scene(p.sc, smem);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    V3 o, d;
    if (p.rays6) {
        o = mk(p.rays6[6 * i], p.rays6[6 * i + 1], p.rays6[6 * i + 2]);
        d = mk(p.rays6[6 * i + 3], p.rays6[6 * i + 4], p.rays6[6 * i + 5]);
    } else {
        o = mk(p.cam[0], p.cam[1], p.cam[2]);
        d = primary_dir(p.cam, (int)(i % (uint32_t)p.W), (int)(i / (uint32_t)p.W), p.W, p.H);
    }
    Hit h;
    nearest_hit(p.sc, smem, o, d, p.mesh_mode, h);
    if (p.out_id) p.out_id[i] = h.id;
    if (p.out_t) p.out_t[i] = h.t;
}

// One iteration of the path loop per ray (single-bounce differential test).
extern "C" __global__ void __launch_bounds__(kTraceThreads)
RT_KERNEL_NAME(k_probe_bounce)(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    stage_scene(p.sc, smem);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= p.n) return;
    V3 o = mk(p.rays6[6 * i], p.rays6[6 * i + 1], p.rays6[6 * i + 2]);
    V3 d = mk(p.rays6[6 * i + 3], p.rays6[6 * i + 4], p.rays6[6 * i + 5]);
    V3 out = mk(1.0f, 1.0f, 1.0f);
    const PixelRng pr = make_pixel_rng(p.gx[i], p.gy[i], (uint32_t)p.W, p.seed);
    ShadeEnv env; env.sc = &p.sc; env.tex = &p.tex; env.table = p.table; env.rng_mode = p.rng_mode;
    Hit h;
    nearest_hit(p.sc, smem, o, d, p.mesh_mode, h);
    bool ended;
    if (h.id == kIdMiss) { out = mk(0.0f, 0.0f, 0.0f); ended = true; }
    else ended = shade(env, smem, h, pr, o, d, out, p.sample[i], p.bounce[i]);
    p.rays6_out[6 * i] = o.x; p.rays6_out[6 * i + 1] = o.y; p.rays6_out[6 * i + 2] = o.z;
    p.rays6_out[6 * i + 3] = d.x; p.rays6_out[6 * i + 4] = d.y; p.rays6_out[6 * i + 5] = d.z;
    p.col3[3 * i] = out.x; p.col3[3 * i + 1] = out.y; p.col3[3 * i + 2] = out.z;
    p.ended[i] = ended ? 1 : 0;
}

inline size_t staged_bytes(const DevScene& sc) { return ((size_t)sc.stream_f4 + 3u * sc.n_mesh) * sizeof(float4); }

template <class K, class P>
inline cudaError_t launch_staged(K kernel, const P& p, size_t work_items, cudaStream_t s) {
    const size_t smem = staged_bytes(p.sc);
    if (smem > kMaxStagedBytes) return cudaErrorInvalidValue;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return e;
    }
    const unsigned blocks = (unsigned)((work_items + kTraceThreads - 1) / kTraceThreads);
    if (blocks == 0) return cudaSuccess;
    kernel<<<blocks, kTraceThreads, smem, s>>>(p);
    return cudaGetLastError();
}

}  // namespace RT_FLAVOUR
}  // namespace rtb
