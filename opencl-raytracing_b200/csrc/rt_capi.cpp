// csrc/rt_capi.cpp — host side of the C ABI (include/rt_capi.h): context, device buffers, scene conversion
// from the reference's host AoS layouts to the device test stream, launches, read-backs.
//
// This layer takes over the roles of KernelGL (src/kernelgl.cpp: device/context/queue), of
// RayTracer::createCLBuffers/setKernelArgs/render/renderAgain (src/raytracer.cpp:63-165) and of
// SceneCreator::setupBuffers/createScene (src/scene.cpp:46-87) in the reference.
#include "../../include/rt_capi.h"
#include "rt_bvh.h"
#include "rt_device.h"

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <random>
#include <string>
#include <thread>
#include <vector>

using namespace rtb;

namespace {

thread_local std::string g_last_error;

int fail(int code, const char* fmt, ...) {
    char buf[512];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_last_error = buf;
    return code;
}

#define CU(call)                                                                                      \
    do {                                                                                              \
        cudaError_t e_ = (call);                                                                      \
        if (e_ != cudaSuccess) return fail(RT_ERR_CUDA, "%s: %s", #call, cudaGetErrorString(e_));     \
    } while (0)

// Host AoS records, include/scene.h:32-81 (cl_float3 = 16 bytes)
struct HMaterial { int32_t type; int32_t pad0[3]; float color[4]; float extra; float pad1[3]; };
struct HSphere { float pos[4]; float r; uint32_t mat; float pad[2]; };
struct HPlane { float pos[4]; float normal[4]; uint32_t mat; float pad[3]; };
struct HLens { float pos[4]; float p1[4]; float p2[4]; float r1, r2; uint32_t mat; float pad; };
struct HMesh { uint32_t vertex_anchor, index_anchor, face_count, texture_id; };
struct HModel { uint32_t mesh_anchor, mesh_count, mat; };
static_assert(sizeof(HMaterial) == 48 && sizeof(HSphere) == 32 && sizeof(HPlane) == 48 && sizeof(HLens) == 64 &&
              sizeof(HMesh) == 16 && sizeof(HModel) == 12, "reference host layouts");

template <class T>
struct DevBuf {
    T* p = nullptr;
    size_t n = 0, cap = 0;
    ~DevBuf() { release(); }
    void release() { if (p) cudaFree(p); p = nullptr; n = cap = 0; }
    // Enqueue the copy on s (the pageable source is staged before the call returns); the allocation is reused while it
    // is large enough.  The caller synchronises once after all uploads.
    cudaError_t upload(const std::vector<T>& v, cudaStream_t s) { return upload(v.data(), v.size(), s); }
    cudaError_t upload(const T* src, size_t count, cudaStream_t s) {
        const size_t want = count ? count : 1;
        if (want > cap) {
            cudaError_t e = cudaStreamSynchronize(s);      // kernels of an earlier render may still read the old array
            if (e != cudaSuccess) return e;
            if (p) cudaFree(p);
            p = nullptr; cap = 0;
            e = cudaMalloc(&p, want * sizeof(T));
            if (e != cudaSuccess) { n = 0; return e; }
            cap = want;
        }
        n = count;
        return n ? cudaMemcpyAsync(p, src, n * sizeof(T), cudaMemcpyHostToDevice, s) : cudaSuccess;
    }
};

}  // namespace

struct rt_ctx {
    int device = 0, W = 0, H = 0;
    cudaStream_t own_stream = nullptr, stream = nullptr;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev2 = nullptr;
    float4* d_image = nullptr;
    float4* d_accum_own = nullptr;
    float4* d_accum = nullptr;      // own or caller-supplied
    unsigned long long* d_rays = nullptr;
    unsigned long long* d_head_counts = nullptr;   // inside d_rays
    float* d_table = nullptr;
    float4* d_tex = nullptr;
    DevTexture tex{nullptr, 1, 1, 1};
    bool have_scene = false;
    DevScene sc{};
    DevBuf<float4> b_stream, b_sph_shade, b_pla_normal, b_len_shade, b_mat;
    DevBuf<uint32_t> b_sph_mat, b_pla_mat, b_len_mat, b_tri_mesh;
    DevBuf<float2> b_tri_uv;
    DevBuf<uint32_t> b_tri_face;      // reference-global face number per stream triangle (BVH leaf ids)
    BvhBuffers bvh;                   // GPU-built LBVH (empty while the brute-force path is active)
    void* h_stage = nullptr;          // pinned host staging of the per-triangle arrays (grow-only), see convert_scene
    size_t h_stage_cap = 0;
    size_t staged_bytes = 0;          // shared memory the brute-force path needs for this scene
    uint32_t n_finite = 0;            // spheres + lenses + triangles
    rt_params prm{30u, RT_RNG_TABLE, 0u, RT_ARITH_STRICT, RT_MESH_FIRST_FRONT, RT_ACCEL_AUTO, RT_HEAD_ON};
    uint32_t sample_counter = 0;
    int rx0 = 0, ry0 = 0, rx1 = 0, ry1 = 0;   // render region (rt_set_region); the whole frame by default
    int head_region[6] = {0, 0, 0, 0, 0, 0};
    uint32_t split_n = 1, split_i = 0;        // pixel split (rt_set_pixel_split)
    int64_t max_texture_used = -1;   // highest texture layer referenced by a textured mesh
    // fast arithmetic: Exact primary-hit ids of the current (scene, camera, mesh rule), see ensure_certified()
    uint32_t* d_certify = nullptr;
    float certify_cam[12] = {0};
    uint32_t certify_mesh_mode = 0;
    bool certify_valid = false;
    // path heads of the current (scene, camera, depth, mesh rule, accel), see ensure_head()
    float4* d_head_rec = nullptr;
    uint8_t* d_head_flag = nullptr;
    uint32_t* d_head_list = nullptr;
    uint32_t *d_head_key = nullptr, *d_head_key_sorted = nullptr, *d_head_iota = nullptr;
    void* d_head_scratch = nullptr;
    size_t head_scratch_bytes = 0;
    uint32_t head_n_slots = 0, head_n_stoch = 0;
    uint64_t head_segments = 0;       // sum over the pixels of the segments in their path head
    float head_cam[12] = {0};
    uint32_t head_mesh_mode = 0, head_depth = 0;
    bool head_bvh = false, head_valid = false;
    float4* d_partial = nullptr;      // chunk sums of k_trace_tail_* (TraceLaunch::partial)
    size_t partial_cap = 0;
    int sm_count = 0;
    // pixel blocks of the current path heads by decreasing cost (TraceLaunch::block_order), see order_blocks()
    uint32_t *d_block_cost = nullptr, *d_block_sorted = nullptr, *d_block_iota = nullptr, *d_block_order = nullptr;
    void* d_block_scratch = nullptr;
    size_t block_scratch_bytes = 0;
    uint32_t block_cap = 0;
    bool order_valid = false;
    // peers of the sample split (rt_ipc_open_peers / rt_set_peers_direct)
    int peer_n = 0, peer_rank = 0;
    void* peer_accum[kMaxPeers] = {nullptr};
    void* peer_image0 = nullptr;
    bool peer_opened[kMaxPeers] = {false};   // accum pointers obtained from cudaIpcOpenMemHandle
    bool peer_image_opened = false;
    float4* finalize_target = nullptr;        // rt_set_finalize_target: where rt_finalize* store the image instead of the image buffer
    bool peer_flags_ok = false;               // peers were opened through CUDA IPC: every buffer carries its flag words
    uint32_t peer_epoch = 0;
    uint32_t* d_barrier_timeout = nullptr;
    // Deferred render() / renderAgain() calls (rt_render, rt_render_again): consecutive samples with one camera are
    // collected and rendered by ONE launch when something observes the image or the statistics (flush_pending) — the
    // sqrt(mean) chain of those samples inside one launch is bit-identical to one launch per sample.
    uint32_t pend_first = 0, pend_n = 0;
    float pend_cam[12] = {0};
    bool count_work = false;          // rt_set_work_counting: LBVH renders go through the instrumented (untimed) kernel
    unsigned long long* d_work = nullptr;
    uint64_t work_counts[8] = {0};
    rt_stats stats{};
};

namespace {

struct Guard {   // make the context's device current for the duration of a call
    int prev = -1;
    bool ok = true;
    explicit Guard(const rt_ctx* c) {
        if (cudaGetDevice(&prev) != cudaSuccess) { ok = false; return; }
        if (prev != c->device && cudaSetDevice(c->device) != cudaSuccess) ok = false;
    }
    ~Guard() { if (prev >= 0) cudaSetDevice(prev); }
};

inline void sub3(const float* a, const float* b, float* r) { r[0] = a[0] - b[0]; r[1] = a[1] - b[1]; r[2] = a[2] - b[2]; }

// Reference AoS -> device test stream + shade arrays.  All float precomputation is a single correctly
// rounded operation per value (this file is built with -ffp-contract=off), so e1, e2, r*r and the unit
// normal are bit-identical to what kernels/raytracer.cl computes per ray.
int convert_scene(rt_ctx* c, const rt_scene_desc* d) {
    const HMaterial* mats = (const HMaterial*)d->materials;
    const HSphere* sph = (const HSphere*)d->spheres;
    const HPlane* pla = (const HPlane*)d->planes;
    const HLens* len = (const HLens*)d->lenses;
    const float* vtx = (const float*)d->vertices;
    const float* uvs = (const float*)d->uvs;
    const uint32_t* idx = (const uint32_t*)d->indices;
    const HMesh* meshes = (const HMesh*)d->meshes;
    const HModel* models = (const HModel*)d->models;

    if (d->n_materials == 0) return fail(RT_ERR_INVALID, "scene has no materials");
    if ((d->n_materials && !mats) || (d->n_spheres && !sph) || (d->n_planes && !pla) || (d->n_lenses && !len) ||
        (d->n_vertices && !vtx) || (d->n_indices && !idx) || (d->n_meshes && !meshes) || (d->n_models && !models))
        return fail(RT_ERR_INVALID, "null array with non-zero count");
    for (uint32_t i = 0; i < d->n_materials; ++i)
        if (mats[i].type < 0 || mats[i].type > 5) return fail(RT_ERR_INVALID, "material %u has unknown type %d", i, mats[i].type);
    if (d->n_materials > 0x00FFFFFFu) return fail(RT_ERR_UNSUPPORTED, "more than 2^24 materials");
    auto is_textured = [&](uint32_t m) { return mats[m].type == 4; };
    // per-primitive material word of the device arrays: index in the low 24 bits, type in the high 8 (rt_kernels.cuh)
    auto packed = [&](uint32_t m) { return m | ((uint32_t)mats[m].type << 24); };

    std::vector<float4> stream, sph_shade, pla_normal, len_shade, mat;
    std::vector<uint32_t> sph_mat, pla_mat, len_mat;
    std::vector<MeshInst> inst;

    for (uint32_t i = 0; i < d->n_spheres; ++i) {
        if (sph[i].mat >= d->n_materials) return fail(RT_ERR_INVALID, "sphere %u: material %u does not exist", i, sph[i].mat);
        if (is_textured(sph[i].mat)) return fail(RT_ERR_UNSUPPORTED, "sphere %u: textured material on a non-mesh primitive (uninitialised uv in the reference)", i);
        stream.push_back(make_float4(sph[i].pos[0], sph[i].pos[1], sph[i].pos[2], sph[i].r * sph[i].r));
        sph_shade.push_back(make_float4(sph[i].pos[0], sph[i].pos[1], sph[i].pos[2], sph[i].r));
        sph_mat.push_back(packed(sph[i].mat));
    }
    const uint32_t off_pla = (uint32_t)stream.size();
    for (uint32_t i = 0; i < d->n_planes; ++i) {
        if (pla[i].mat >= d->n_materials) return fail(RT_ERR_INVALID, "plane %u: material %u does not exist", i, pla[i].mat);
        if (is_textured(pla[i].mat)) return fail(RT_ERR_UNSUPPORTED, "plane %u: textured material on a non-mesh primitive", i);
        stream.push_back(make_float4(pla[i].pos[0], pla[i].pos[1], pla[i].pos[2], 0.0f));
        stream.push_back(make_float4(pla[i].normal[0], pla[i].normal[1], pla[i].normal[2], 0.0f));
        pla_normal.push_back(make_float4(pla[i].normal[0], pla[i].normal[1], pla[i].normal[2], 0.0f));
        pla_mat.push_back(packed(pla[i].mat));
    }
    const uint32_t off_len = (uint32_t)stream.size();
    for (uint32_t i = 0; i < d->n_lenses; ++i) {
        if (len[i].mat >= d->n_materials) return fail(RT_ERR_INVALID, "lens %u: material %u does not exist", i, len[i].mat);
        if (is_textured(len[i].mat)) return fail(RT_ERR_UNSUPPORTED, "lens %u: textured material on a non-mesh primitive", i);
        stream.push_back(make_float4(len[i].p1[0], len[i].p1[1], len[i].p1[2], len[i].r1 * len[i].r1));
        stream.push_back(make_float4(len[i].p2[0], len[i].p2[1], len[i].p2[2], len[i].r2 * len[i].r2));
        len_shade.push_back(make_float4(len[i].p1[0], len[i].p1[1], len[i].p1[2], len[i].r1));
        len_shade.push_back(make_float4(len[i].p2[0], len[i].p2[1], len[i].p2[2], len[i].r2));
        len_mat.push_back(packed(len[i].mat));
    }
    const uint32_t off_tri = (uint32_t)stream.size();
    uint32_t n_tri = 0;
    int64_t max_tex = -1;
    // pass 1: validate the model / mesh tables, lay the mesh instances out in hitModel order
    struct Src { uint32_t vertex_anchor, index_anchor, mesh; };
    std::vector<Src> src;
    for (uint32_t mo = 0; mo < d->n_models; ++mo) {
        const HModel& M = models[mo];
        if (M.mat >= d->n_materials) return fail(RT_ERR_INVALID, "ERROR: MATERIAL OF ID: %u DOES NOT EXIST", M.mat);
        if ((uint64_t)M.mesh_anchor + M.mesh_count > d->n_meshes) return fail(RT_ERR_INVALID, "model %u: mesh range out of bounds", mo);
        for (uint32_t k = 0; k < M.mesh_count; ++k) {
            const HMesh& me = meshes[M.mesh_anchor + k];
            if ((uint64_t)me.index_anchor + 3ull * me.face_count > d->n_indices) return fail(RT_ERR_INVALID, "mesh %u: index range out of bounds", M.mesh_anchor + k);
            if ((uint64_t)n_tri + me.face_count > 0x0FFFFFFFull) return fail(RT_ERR_UNSUPPORTED, "more than 2^28 triangles");
            if (is_textured(M.mat)) max_tex = std::max<int64_t>(max_tex, me.texture_id);
            MeshInst mi{};
            mi.tri_first = n_tri; mi.tri_count = me.face_count; mi.face_id_base = me.index_anchor / 3u;
            mi.mat_id = packed(M.mat); mi.texture_id = me.texture_id;
            inst.push_back(mi);
            src.push_back(Src{me.vertex_anchor, me.index_anchor, M.mesh_anchor + k});
            n_tri += me.face_count;
        }
    }
    // pass 2: the per-triangle records, written in place; long meshes are split over the host threads (a 1.3 M-triangle
    // mesh took 0.6 s on one core, which is what an upload-per-frame caller of a large scene waits for)
    // They go straight into pinned host memory owned by the context (no page faults or zero fill per upload, and the
    // copies to the device run at PCIe speed), laid out as the device wants it:
    // all = [test stream | mesh-instance table (3 x 16 B each) | plane offsets (padded to x4)], then uv, mesh, face id.
    const size_t stream_f4 = (size_t)off_tri + 3u * (size_t)n_tri, pd_n = ((size_t)n_tri + 3u) / 4u;
    const size_t all_f4 = stream_f4 + 3u * inst.size() + pd_n;
    auto up16 = [](size_t b) { return (b + 15u) & ~(size_t)15u; };
    const size_t b_all = all_f4 * sizeof(float4), b_uv = up16(3u * (size_t)n_tri * sizeof(float2)), b_u32 = up16((size_t)n_tri * sizeof(uint32_t));
    const size_t b_total = b_all + b_uv + 2u * b_u32 + 16u;
    CU(cudaStreamSynchronize(c->stream));                     // nothing may still read the staging area or the old arrays
    if (b_total > c->h_stage_cap) {
        if (c->h_stage) cudaFreeHost(c->h_stage);
        c->h_stage = nullptr; c->h_stage_cap = 0;
        CU(cudaMallocHost(&c->h_stage, b_total));
        c->h_stage_cap = b_total;
    }
    float4* all = static_cast<float4*>(c->h_stage);
    float2* tri_uv = reinterpret_cast<float2*>(static_cast<uint8_t*>(c->h_stage) + b_all);
    uint32_t* tri_mesh = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(c->h_stage) + b_all + b_uv);
    uint32_t* tri_face = reinterpret_cast<uint32_t*>(static_cast<uint8_t*>(c->h_stage) + b_all + b_uv + b_u32);
    float* tri_pd = reinterpret_cast<float*>(all + stream_f4 + 3u * inst.size());
    if (off_tri) std::memcpy(all, stream.data(), (size_t)off_tri * sizeof(float4));
    for (size_t i = n_tri; i < 4u * pd_n; ++i) tri_pd[i] = 0.0f;
    for (size_t m = 0; m < inst.size(); ++m) {
        MeshInst& mi = inst[m];
        const Src sm = src[m];
        if (mi.tri_count == 0) continue;
        const unsigned n_thr = mi.tri_count < 65536u ? 1u : std::min(16u, std::max(1u, std::thread::hardware_concurrency()));
        std::vector<double> sums(4u * n_thr, 0.0);            // per thread: centroid sums, then the largest squared distance
        std::atomic<bool> bad_index{false};
        auto for_faces = [&](auto&& body) {
            const uint32_t per = (mi.tri_count + n_thr - 1u) / n_thr;
            std::vector<std::thread> pool;
            for (unsigned t = 1; t < n_thr; ++t)
                pool.emplace_back([&, t] { body(std::min(mi.tri_count, t * per), std::min(mi.tri_count, (t + 1u) * per), t); });
            body(0u, std::min(mi.tri_count, per), 0u);
            for (auto& th : pool) th.join();
        };
        for_faces([&](uint32_t f0, uint32_t f1, unsigned t) {
            double cx = 0, cy = 0, cz = 0;
            for (uint32_t f = f0; f < f1; ++f) {
                const float* P[3];
                const size_t tri = (size_t)mi.tri_first + f;
                for (int j = 0; j < 3; ++j) {
                    const uint64_t vi = (uint64_t)sm.vertex_anchor + idx[sm.index_anchor + 3 * f + j];  // kernels/raytracer.cl:93-99
                    if (vi >= d->n_vertices) { bad_index = true; return; }
                    P[j] = vtx + 4 * vi;
                    tri_uv[3 * tri + j] = (uvs && vi < d->n_uvs) ? make_float2(uvs[2 * vi], uvs[2 * vi + 1]) : make_float2(0.0f, 0.0f);
                    cx += P[j][0]; cy += P[j][1]; cz += P[j][2];
                }
                float e1[3], e2[3];
                sub3(P[1], P[0], e1);
                sub3(P[2], P[0], e2);
                const float nx = e1[1] * e2[2] - e1[2] * e2[1], ny = e1[2] * e2[0] - e1[0] * e2[2], nz = e1[0] * e2[1] - e1[1] * e2[0];
                const float inv = 1.0f / std::sqrt(nx * nx + ny * ny + nz * nz);
                float4* T = all + off_tri + 3 * tri;
                T[0] = make_float4(P[0][0], P[0][1], P[0][2], e1[0]);
                T[1] = make_float4(e1[1], e1[2], e2[0], e2[1]);
                T[2] = make_float4(e2[2], nx * inv, ny * inv, nz * inv);
                {   // plane offset for the conservative front-side filter (not parity relevant: it only removes
                    // triangles whose hit parameter would be negative); margin covers float rounding on the device
                    const double nd[3] = {(double)(nx * inv), (double)(ny * inv), (double)(nz * inv)};
                    const double pd = nd[0] * P[0][0] + nd[1] * P[0][1] + nd[2] * P[0][2];
                    tri_pd[tri] = (float)(pd - 4e-6 * (1.0 + std::fabs(pd)));
                }
                tri_mesh[tri] = (uint32_t)m;
                tri_face[tri] = mi.face_id_base + f;
            }
            sums[4 * t] = cx; sums[4 * t + 1] = cy; sums[4 * t + 2] = cz;
        });
        if (bad_index) return fail(RT_ERR_INVALID, "mesh %u: vertex index out of bounds", sm.mesh);
        // conservative bounding sphere of the instance: centroid of the face corners + largest distance, inflated
        double cx = 0, cy = 0, cz = 0;
        for (unsigned t = 0; t < n_thr; ++t) { cx += sums[4 * t]; cy += sums[4 * t + 1]; cz += sums[4 * t + 2]; }
        const double n_corners = 3.0 * mi.tri_count;
        cx /= n_corners; cy /= n_corners; cz /= n_corners;
        for_faces([&](uint32_t f0, uint32_t f1, unsigned t) {
            double r2 = 0;
            for (uint32_t f = f0; f < f1; ++f)
                for (int j = 0; j < 3; ++j) {
                    const float* q = vtx + 4 * ((uint64_t)sm.vertex_anchor + idx[sm.index_anchor + 3 * f + j]);
                    const double dx = q[0] - cx, dy = q[1] - cy, dz = q[2] - cz;
                    r2 = std::fmax(r2, dx * dx + dy * dy + dz * dz);
                }
            sums[4 * t + 3] = r2;
        });
        double r2 = 0;
        for (unsigned t = 0; t < n_thr; ++t) r2 = std::fmax(r2, sums[4 * t + 3]);
        mi.bound[0] = (float)cx; mi.bound[1] = (float)cy; mi.bound[2] = (float)cz;
        mi.bound[3] = (float)(std::sqrt(r2) * 1.001 + 1e-4);
    }
    for (uint32_t i = 0; i < d->n_materials; ++i) {
        float t;
        std::memcpy(&t, &mats[i].type, 4);
        mat.push_back(make_float4(t, mats[i].extra, 0.0f, 0.0f));
        mat.push_back(make_float4(mats[i].color[0], mats[i].color[1], mats[i].color[2], 0.0f));
    }
    if (!inst.empty()) std::memcpy(all + stream_f4, inst.data(), inst.size() * sizeof(MeshInst));
    c->staged_bytes = all_f4 * sizeof(float4) + (kTraceThreads / 32) * 1024u;
    c->n_finite = d->n_spheres + d->n_lenses + n_tri;
    if (c->prm.accel == RT_ACCEL_BRUTE && c->staged_bytes > kMaxStagedBytes)
        return fail(RT_ERR_UNSUPPORTED, "scene needs %zu bytes of shared memory for the brute-force path (max %zu); use RT_ACCEL_AUTO or RT_ACCEL_BVH",
                    c->staged_bytes, kMaxStagedBytes);

    cudaStream_t s = c->stream;
    CU(c->b_stream.upload(all, all_f4, s));  CU(c->b_sph_shade.upload(sph_shade, s)); CU(c->b_sph_mat.upload(sph_mat, s));
    CU(c->b_pla_normal.upload(pla_normal, s)); CU(c->b_pla_mat.upload(pla_mat, s));
    CU(c->b_len_shade.upload(len_shade, s)); CU(c->b_len_mat.upload(len_mat, s));
    CU(c->b_tri_uv.upload(tri_uv, 3u * (size_t)n_tri, s)); CU(c->b_tri_mesh.upload(tri_mesh, n_tri, s));
    CU(c->b_tri_face.upload(tri_face, n_tri, s)); CU(c->b_mat.upload(mat, s));
    CU(cudaStreamSynchronize(s));

    DevScene& sc = c->sc;
    sc.stream = c->b_stream.p; sc.stream_f4 = (uint32_t)stream_f4;
    sc.n_sph = d->n_spheres; sc.n_pla = d->n_planes; sc.n_len = d->n_lenses; sc.n_tri = n_tri; sc.n_mesh = (uint32_t)inst.size();
    sc.off_pla = off_pla; sc.off_len = off_len; sc.off_tri = off_tri;
    sc.mesh = reinterpret_cast<const MeshInst*>(c->b_stream.p + stream_f4);
    sc.sph_shade = c->b_sph_shade.p; sc.sph_mat = c->b_sph_mat.p;
    sc.pla_normal = c->b_pla_normal.p; sc.pla_mat = c->b_pla_mat.p;
    sc.len_shade = c->b_len_shade.p; sc.len_mat = c->b_len_mat.p;
    sc.tri_uv = c->b_tri_uv.p; sc.tri_mesh = c->b_tri_mesh.p; sc.tri_pd = reinterpret_cast<const float*>(c->b_stream.p + stream_f4 + 3 * inst.size());
    sc.mat = c->b_mat.p; sc.n_mat = d->n_materials;
    c->stats.n_primitives = d->n_spheres + d->n_planes + d->n_lenses + n_tri;
    c->stats.bvh_nodes = 0;
    free_lbvh(c->bvh);
    c->max_texture_used = max_tex;
    c->certify_valid = false;
    c->head_valid = false;
    c->have_scene = true;
    return RT_OK;
}

// Pick the intersection path for the current scene and parameters and build the LBVH on demand.
// RT_ACCEL_AUTO: brute force from shared memory while the scene fits and has at most 512 finite primitives.
int ensure_accel(rt_ctx* c) {
    bool want_bvh = c->prm.accel == RT_ACCEL_BVH;
    if (c->prm.accel == RT_ACCEL_AUTO) want_bvh = c->staged_bytes > kMaxStagedBytes || c->n_finite > 512u;
    if (c->prm.accel == RT_ACCEL_BRUTE && c->staged_bytes > kMaxStagedBytes)
        return fail(RT_ERR_UNSUPPORTED, "scene needs %zu bytes of shared memory for the brute-force path (max %zu)", c->staged_bytes, kMaxStagedBytes);
    if (want_bvh && c->n_finite == 0) want_bvh = false;
    if (!want_bvh) {
        if (c->bvh.n_leaves) { free_lbvh(c->bvh); c->stats.bvh_nodes = 0; c->certify_valid = false; c->head_valid = false; }
        return RT_OK;
    }
    if (c->sc.n_tri > 0 && c->prm.mesh_mode != RT_MESH_NEAREST)
        return fail(RT_ERR_UNSUPPORTED, "the LBVH path answers the nearest-front-facing mesh rule only: set mesh_mode = RT_MESH_NEAREST "
                                        "(the reference's first-front-facing-in-face-order rule depends on scan order)");
    if (c->bvh.n_leaves == c->n_finite) return RT_OK;
    BvhBuildInput in{};
    in.stream = c->sc.stream; in.off_len = c->sc.off_len; in.off_tri = c->sc.off_tri;
    in.n_sph = c->sc.n_sph; in.n_len = c->sc.n_len; in.n_tri = c->sc.n_tri; in.tri_face_id = c->b_tri_face.p;
    {   // a failed build leaves no half-built tree behind: the next render call retries (or fails again), never traverses garbage
        const cudaError_t be = getenv("RTB200_TEST_FAIL_BVH_BUILD") ? cudaErrorMemoryAllocation : build_lbvh(in, c->bvh, c->stream);
        if (be != cudaSuccess) {
            free_lbvh(c->bvh);
            c->stats.bvh_nodes = 0; c->certify_valid = false; c->head_valid = false;
            return fail(RT_ERR_CUDA, "build_lbvh: %s", cudaGetErrorString(be));
        }
    }
    c->stats.bvh_nodes = c->bvh.n_leaves > 1 ? c->bvh.n_leaves - 1 : 1;
    c->stats.kernel_launches += 8;
    c->certify_valid = false;
    c->head_valid = false;
    return RT_OK;
}
DevBvh dev_bvh(const rt_ctx* c) {
    DevBvh b{};
    b.nodes = c->bvh.nodes; b.leaves = c->bvh.leaves; b.n_leaves = c->bvh.n_leaves;
    b.root = c->bvh.n_leaves > 1 ? 0 : ~0;
    return b;
}

// Fast arithmetic certifies its primary hits: one Exact nearest-hit query per pixel, recomputed only when the
// scene, the camera or the mesh rule changes.  The fast kernels compare their own primary id with it and fall
// back to the Exact intersection on the (knife-edge) pixels where they differ.
int ensure_certified(rt_ctx* c, const float* cam) {
    if (c->certify_valid && c->certify_mesh_mode == c->prm.mesh_mode && std::memcmp(c->certify_cam, cam, 48) == 0) return RT_OK;
    const size_t n = (size_t)c->W * c->H;
    if (!c->d_certify) CU(cudaMalloc(&c->d_certify, n * sizeof(uint32_t)));
    ProbeLaunch L{};
    L.sc = c->sc; L.tex = c->tex; L.table = c->d_table;
    std::memcpy(L.cam, cam, 48);
    L.W = c->W; L.H = c->H; L.depth = c->prm.max_depth; L.mesh_mode = c->prm.mesh_mode;
    L.n = (uint32_t)n; L.out_id = c->d_certify; L.bvh = dev_bvh(c);
    CU(launch_probe_hits_strict(L, c->stream));
    c->stats.kernel_launches += 1;
    std::memcpy(c->certify_cam, cam, 48);
    c->certify_mesh_mode = c->prm.mesh_mode;
    c->certify_valid = true;
    return RT_OK;
}

// Path heads (rt_device.h, DevHead): one Exact walk per pixel up to its first random-consuming hit, then a
// partition of the slots into stochastic / other.  Rebuilt only when scene, camera, depth, mesh rule or the
// intersection path change.  *rebuilt tells the caller whether this call did the work.
int ensure_head(rt_ctx* c, const float* cam, bool* rebuilt) {
    *rebuilt = false;
    const bool bvh = c->bvh.n_leaves != 0;
    const int region[6] = {c->rx0, c->ry0, c->rx1, c->ry1, (int)c->split_n, (int)c->split_i};
    if (c->head_valid && c->head_mesh_mode == c->prm.mesh_mode && c->head_depth == c->prm.max_depth && c->head_bvh == bvh &&
        std::memcmp(c->head_cam, cam, 48) == 0 && std::memcmp(c->head_region, region, sizeof region) == 0)
        return RT_OK;
    const uint32_t n_slots = (uint32_t)(((c->W + 7) / 8) * ((c->H + 3) / 4)) * 32u;
    if (!c->d_head_rec) {
        CU(cudaMalloc(&c->d_head_rec, (size_t)n_slots * 4 * sizeof(float4)));
        CU(cudaMalloc(&c->d_head_flag, n_slots));
        CU(cudaMalloc(&c->d_head_list, (size_t)n_slots * sizeof(uint32_t)));
        CU(cudaMalloc(&c->d_head_key, (size_t)n_slots * sizeof(uint32_t)));
        CU(cudaMalloc(&c->d_head_key_sorted, (size_t)n_slots * sizeof(uint32_t)));
        CU(cudaMalloc(&c->d_head_iota, (size_t)n_slots * sizeof(uint32_t)));
        CU(fill_iota(c->d_head_iota, n_slots, c->stream));
        size_t bytes = 0;
        CU(sort_heads(c->d_head_key, c->d_head_key_sorted, c->d_head_iota, c->d_head_list, n_slots, nullptr, bytes, c->stream));
        CU(cudaMalloc(&c->d_head_scratch, bytes ? bytes : 1));
        c->head_scratch_bytes = bytes;
    }
    HeadLaunch L{};
    L.sc = c->sc;
    std::memcpy(L.cam, cam, 48);
    L.W = c->W; L.H = c->H; L.depth = c->prm.max_depth; L.mesh_mode = c->prm.mesh_mode; L.bvh = dev_bvh(c);
    L.rx0 = c->rx0; L.ry0 = c->ry0; L.rx1 = c->rx1; L.ry1 = c->ry1; L.split_n = c->split_n; L.split_i = c->split_i;
    L.key_mask = 0x3FFFu;
    if (const char* e = std::getenv("RTB200_HEAD_SORT")) L.key_mask = (uint32_t)std::strtoul(e, nullptr, 0);   // A/B: 0 = slot order
    L.rec = c->d_head_rec; L.flag = c->d_head_flag; L.key = c->d_head_key; L.n_slots = n_slots; L.segments = c->d_head_counts;
    CU(cudaMemsetAsync(c->d_head_counts, 0, 2 * sizeof(unsigned long long), c->stream));
    CU(launch_head(L, c->stream));
    size_t bytes = c->head_scratch_bytes;
    CU(sort_heads(c->d_head_key, c->d_head_key_sorted, c->d_head_iota, c->d_head_list, n_slots, c->d_head_scratch, bytes, c->stream));
    unsigned long long counts[2] = {0, 0};           // segments of the heads, stochastic slots
    CU(cudaMemcpyAsync(counts, c->d_head_counts, sizeof counts, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->head_n_slots = n_slots; c->head_n_stoch = (uint32_t)counts[1]; c->head_segments = counts[0];
    std::memcpy(c->head_cam, cam, 48);
    std::memcpy(c->head_region, region, sizeof region);
    c->head_mesh_mode = c->prm.mesh_mode; c->head_depth = c->prm.max_depth; c->head_bvh = bvh;
    c->head_valid = true;
    c->order_valid = false;
    c->stats.kernel_launches += 4;   // k_head + CUB's radix sort of the 16-bit head keys (histogram, scan, 2 passes; onesweep)
    *rebuilt = true;
    return RT_OK;
}

}  // namespace

// ================================================================================================ API
extern "C" {

static int flush_pending(rt_ctx* c);     // renders the recorded render() / renderAgain() calls (see rt_render)
#define FLUSH(c) do { int fr_ = flush_pending(c); if (fr_ != RT_OK) return fr_; } while (0)

const char* rt_last_error(void) { return g_last_error.c_str(); }
int rt_set_error_message(int code, const char* msg) { return fail(code, "%s", msg ? msg : ""); }   // used by host/host_capi.cpp
const char* rt_version(void) { return "rtb200 0.1 (sm_100a; strict+fast arithmetic; table+counter RNG)"; }

int rt_create(int width, int height, int device, rt_ctx** out) {
    if (!out) return fail(RT_ERR_INVALID, "rt_create: out is null");
    *out = nullptr;
    if (width <= 0 || height <= 0 || width > 65535 || height > 65535)
        return fail(RT_ERR_INVALID, "rt_create: bad size %d x %d", width, height);
    // pixel, slot and list counts are carried as 32-bit integers through the launch structs
    if ((uint64_t)((width + 7) / 8) * (uint64_t)((height + 3) / 4) * 32u >= (1ull << 31))
        return fail(RT_ERR_UNSUPPORTED, "rt_create: %d x %d exceeds the supported 2^31 pixel slots", width, height);
    int n_dev = 0;
    cudaError_t e = cudaGetDeviceCount(&n_dev);
    if (e != cudaSuccess || n_dev == 0)
        return fail(RT_ERR_CUDA, "rt_create: no CUDA device (%s); this library has no CPU fallback",
                    e == cudaSuccess ? "device count is 0" : cudaGetErrorString(e));
    if (device < 0 || device >= n_dev) return fail(RT_ERR_INVALID, "rt_create: device %d out of range (0..%d)", device, n_dev - 1);
    rt_ctx* c = new rt_ctx();
    c->device = device; c->W = width; c->H = height;
    c->rx1 = width; c->ry1 = height;
    Guard g(c);
    const size_t px = (size_t)width * height;
    auto bail = [&](cudaError_t err, const char* what) {
        int r = fail(RT_ERR_CUDA, "rt_create: %s: %s", what, cudaGetErrorString(err));
        rt_destroy(c);
        return r;
    };
    if ((e = cudaStreamCreateWithFlags(&c->own_stream, cudaStreamNonBlocking)) != cudaSuccess) return bail(e, "stream");
    c->stream = c->own_stream;
    if ((e = cudaEventCreate(&c->ev0)) != cudaSuccess) return bail(e, "event");
    if ((e = cudaEventCreate(&c->ev1)) != cudaSuccess) return bail(e, "event");
    if ((e = cudaEventCreate(&c->ev2)) != cudaSuccess) return bail(e, "event");
    if ((e = cudaMalloc(&c->d_image, px * sizeof(float4))) != cudaSuccess) return bail(e, "image buffer");
    // the accumulation buffer carries the flag words of the device-side rank barrier behind its pixels (see PeerBarrier)
    if ((e = cudaMalloc(&c->d_accum_own, px * sizeof(float4) + kPeerFlagWords * sizeof(uint32_t))) != cudaSuccess) return bail(e, "accumulation buffer");
    cudaMemsetAsync(c->d_accum_own + px, 0, kPeerFlagWords * sizeof(uint32_t), c->stream);
    if ((e = cudaMalloc(&c->d_barrier_timeout, sizeof(uint32_t))) != cudaSuccess) return bail(e, "barrier flag");
    cudaMemsetAsync(c->d_barrier_timeout, 0, sizeof(uint32_t), c->stream);
    if ((e = cudaMalloc(&c->d_rays, 6 * sizeof(unsigned long long))) != cudaSuccess) return bail(e, "counters");   // rays, queries, -, queue, head segments, stochastic slots
    c->d_head_counts = c->d_rays + 4;
    c->d_accum = c->d_accum_own;
    cudaMemsetAsync(c->d_image, 0, px * sizeof(float4), c->stream);
    cudaMemsetAsync(c->d_accum, 0, px * sizeof(float4), c->stream);
    // 1x1x1 dummy texture array, as the reference creates for texture-less scenes (src/scene.cpp:188)
    if ((e = cudaMalloc(&c->d_tex, sizeof(float4))) != cudaSuccess) return bail(e, "texture");
    cudaMemsetAsync(c->d_tex, 0, sizeof(float4), c->stream);
    c->tex = DevTexture{c->d_tex, 1, 1, 1};
    if ((e = cudaStreamSynchronize(c->stream)) != cudaSuccess) return bail(e, "init");
    *out = c;
    return RT_OK;
}

int rt_destroy(rt_ctx* c) {
    if (!c) return RT_OK;
    {
        Guard g(c);
        if (c->own_stream) cudaStreamSynchronize(c->own_stream);
        rt_close_peers(c);
        if (c->h_stage) cudaFreeHost(c->h_stage);
        cudaFree(c->d_certify); cudaFree(c->d_work); cudaFree(c->d_barrier_timeout);
        cudaFree(c->d_head_rec); cudaFree(c->d_head_flag); cudaFree(c->d_head_list); cudaFree(c->d_head_key); cudaFree(c->d_head_key_sorted); cudaFree(c->d_head_iota); cudaFree(c->d_head_scratch); cudaFree(c->d_partial);
        cudaFree(c->d_block_cost); cudaFree(c->d_block_sorted); cudaFree(c->d_block_iota); cudaFree(c->d_block_order); cudaFree(c->d_block_scratch);
        cudaFree(c->d_image); cudaFree(c->d_accum_own); cudaFree(c->d_rays); cudaFree(c->d_table); cudaFree(c->d_tex);
        c->b_stream.release(); c->b_sph_shade.release(); c->b_pla_normal.release(); c->b_len_shade.release();
        c->b_mat.release(); c->b_sph_mat.release(); c->b_pla_mat.release(); c->b_len_mat.release();
        c->b_tri_mesh.release(); c->b_tri_uv.release(); c->b_tri_face.release(); free_lbvh(c->bvh);
        if (c->ev0) cudaEventDestroy(c->ev0);
        if (c->ev1) cudaEventDestroy(c->ev1);
        if (c->ev2) cudaEventDestroy(c->ev2);
        if (c->own_stream) cudaStreamDestroy(c->own_stream);
    }
    delete c;
    return RT_OK;
}

int rt_set_stream(rt_ctx* c, void* s) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    FLUSH(c);
    c->stream = s ? (cudaStream_t)s : c->own_stream;
    return RT_OK;
}

int rt_upload_scene(rt_ctx* c, const rt_scene_desc* d) {
    if (!c || !d) return fail(RT_ERR_INVALID, "rt_upload_scene: null argument");
    FLUSH(c);
    Guard g(c);
    c->have_scene = false;
    return convert_scene(c, d);
}

int rt_upload_textures(rt_ctx* c, const float* rgba, int w, int h, int layers) {
    if (!c || !rgba) return fail(RT_ERR_INVALID, "rt_upload_textures: null argument");
    if (w <= 0 || h <= 0 || layers <= 0) return fail(RT_ERR_INVALID, "rt_upload_textures: bad size %d x %d x %d", w, h, layers);
    FLUSH(c);
    Guard g(c);
    const size_t n = (size_t)w * h * layers;
    if (n != (size_t)c->tex.w * c->tex.h * c->tex.layers) {          // new size: allocate first, a failure keeps the old array
        float4* fresh = nullptr;
        CU(cudaMalloc(&fresh, n * sizeof(float4)));
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_tex);
        c->d_tex = fresh;
    }
    c->tex = DevTexture{c->d_tex, w, h, layers};
    CU(cudaMemcpyAsync(c->d_tex, rgba, n * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

// src/raytracer.cpp:67-93 with a caller-chosen seed instead of std::random_device.
int rt_make_random_table(uint32_t seed, float* out) {
    if (!out) return fail(RT_ERR_INVALID, "rt_make_random_table: null output");
    const uint32_t N = RT_RANDOM_TABLE_FLOATS / 4;
    std::mt19937 gen(seed);
    std::uniform_real_distribution<float> uni(0.0f, 1.0f);
    std::normal_distribution<float> nrm(0.0f, 1.0f);
    for (uint32_t i = 0; i < N; ++i) {
        const float gx = nrm(gen);
        const float gy = nrm(gen);
        const float gz = nrm(gen);
        const float radius = std::cbrt(uni(gen));
        const float u = uni(gen);
        const float scale = radius / std::sqrt(gx * gx + gy * gy + gz * gz);
        float* v = out + 3 * (size_t)i;
        v[0] = gx * scale; v[1] = gy * scale; v[2] = gz * scale;
        out[3 * (size_t)N + i] = u;
    }
    return RT_OK;
}

int rt_set_random_table(rt_ctx* c, const float* table, uint32_t n_floats) {
    if (!c || !table) return fail(RT_ERR_INVALID, "rt_set_random_table: null argument");
    if (n_floats != RT_RANDOM_TABLE_FLOATS) return fail(RT_ERR_INVALID, "rt_set_random_table: expected %u floats", RT_RANDOM_TABLE_FLOATS);
    FLUSH(c);
    Guard g(c);
    if (!c->d_table) CU(cudaMalloc(&c->d_table, (size_t)n_floats * sizeof(float)));
    CU(cudaMemcpyAsync(c->d_table, table, (size_t)n_floats * sizeof(float), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return RT_OK;
}

int rt_set_params(rt_ctx* c, const rt_params* p) {
    if (!c || !p) return fail(RT_ERR_INVALID, "rt_set_params: null argument");
    if (p->max_depth == 0) return fail(RT_ERR_INVALID, "max_depth must be >= 1");
    if (p->rng_mode > 2 || p->arith > 1 || p->mesh_mode > 1 || p->accel > 2 || p->path_head > 1) return fail(RT_ERR_INVALID, "rt_set_params: enum out of range");
    FLUSH(c);
    c->prm = *p;
    return RT_OK;
}
int rt_set_region(rt_ctx* c, int x0, int y0, int w, int h) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    FLUSH(c);
    if (w <= 0 || h <= 0) { c->rx0 = c->ry0 = 0; c->rx1 = c->W; c->ry1 = c->H; return RT_OK; }
    if (x0 < 0 || y0 < 0 || x0 > c->W - w || y0 > c->H - h)
        return fail(RT_ERR_INVALID, "rt_set_region: %d x %d at (%d, %d) does not lie inside the %d x %d frame", w, h, x0, y0, c->W, c->H);
    c->rx0 = x0; c->ry0 = y0; c->rx1 = x0 + w; c->ry1 = y0 + h;
    return RT_OK;
}
int rt_set_pixel_split(rt_ctx* c, int n_parts, int part) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    FLUSH(c);
    if (n_parts <= 1) { c->split_n = 1; c->split_i = 0; return RT_OK; }
    if (part < 0 || part >= n_parts) return fail(RT_ERR_INVALID, "rt_set_pixel_split: part %d of %d", part, n_parts);
    c->split_n = (uint32_t)n_parts; c->split_i = (uint32_t)part;
    return RT_OK;
}
int rt_set_work_counting(rt_ctx* c, int on) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    c->count_work = on != 0;
    return RT_OK;
}
int rt_get_work_counts(rt_ctx* c, uint64_t out[8]) {
    if (!c || !out) return fail(RT_ERR_INVALID, "rt_get_work_counts: null argument");
    for (int q = 0; q < 8; ++q) out[q] = c->work_counts[q];
    return RT_OK;
}
int rt_get_params(rt_ctx* c, rt_params* p) {
    if (!c || !p) return fail(RT_ERR_INVALID, "rt_get_params: null argument");
    *p = c->prm;
    return RT_OK;
}

// Sample chunks of the brute-force tail kernel (TraceLaunch::chunks): the fast arm's linear sums are split into items of
// >= 20 samples.  Without a block order (order_blocks) until the launch has ~64 waves of blocks, so that the last wave
// is a few percent of the kernel and not a quarter; with the longest blocks first whole pixels are the better items (the
// sum over more samples varies less between the lanes of a warp) up to ~1024 samples per item and >= 4 waves
// (measured: C2 68.6 Grays/s at 1 chunk, 66.4 at 8; C3, 4096 spp: 42.5 at 1, 57.5 at 4, 55.8 at 32).
// The strict arm keeps one item per pixel: its sums are added in sample order, like the oracle's.
static int plan_chunks(rt_ctx* c, TraceLaunch& L) {
    L.chunks = 1; L.chunk_spp = L.n; L.partial = nullptr;
    if (c->prm.arith != RT_ARITH_FAST || L.accum_mode != RT_ACCUM_SUM || L.head.n_stoch == 0) return RT_OK;
    if (L.bvh.n_leaves) {
        // LBVH kernel: items of kBvhChunkSpp samples (see trace_tail_bvh_body: a few pixels behind glass chains cost hundreds of
        // times the average), at most 32 chunks of a pixel (the chunk sums take chunks x 16 B per stochastic pixel).
        uint64_t chunk_spp = 2;
        if (const char* e = std::getenv("RTB200_BVH_CHUNK_SPP")) chunk_spp = (uint64_t)std::max(1ll, std::atoll(e));
        uint64_t chunks = std::min<uint64_t>((L.n + chunk_spp - 1u) / chunk_spp, 32u);
        chunks = std::min<uint64_t>(chunks, (1ull << 31) / L.head.n_stoch);
        if (chunks < 2) return RT_OK;
        L.chunk_spp = (uint32_t)((L.n + chunks - 1u) / chunks);
        L.chunks = (L.n + L.chunk_spp - 1u) / L.chunk_spp;
        const size_t want = (size_t)L.chunks * L.head.n_stoch;
        if (want > c->partial_cap) {
            CU(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_partial); c->d_partial = nullptr; c->partial_cap = 0;
            CU(cudaMalloc(&c->d_partial, want * sizeof(float4)));
            c->partial_cap = want;
        }
        L.partial = c->d_partial;
        return RT_OK;
    }
    if (!c->sm_count) {
        int dev = 0;
        CU(cudaGetDevice(&dev));
        CU(cudaDeviceGetAttribute(&c->sm_count, cudaDevAttrMultiProcessorCount, dev));
    }
    const uint64_t resident = (uint64_t)c->sm_count * 1024u, n_stoch = L.head.n_stoch;
    uint64_t chunks = (64u * resident + n_stoch - 1u) / n_stoch;
    // at least ~8 waves of blocks (a rank of the pixel split owns 1 / N of the pixels: N = 2 left 5.5 waves of whole pixels)
    if (L.block_order) chunks = std::max<uint64_t>((L.n + 1023u) / 1024u, (8u * resident + n_stoch - 1u) / n_stoch);
    if (const char* e = std::getenv("RTB200_TAIL_CHUNKS")) chunks = (uint64_t)std::atoll(e);
    chunks = std::min<uint64_t>({chunks, L.n / 20u, (1ull << 31) / n_stoch});
    if (chunks < 2) return RT_OK;
    L.chunk_spp = (uint32_t)((L.n + chunks - 1u) / chunks);
    L.chunks = (L.n + L.chunk_spp - 1u) / L.chunk_spp;
    const size_t want = (size_t)L.chunks * n_stoch;
    if (want > c->partial_cap) {
        CU(cudaStreamSynchronize(c->stream));
        cudaFree(c->d_partial); c->d_partial = nullptr; c->partial_cap = 0;
        CU(cudaMalloc(&c->d_partial, want * sizeof(float4)));
        c->partial_cap = want;
    }
    L.partial = c->d_partial;
    return RT_OK;
}

// Longest blocks first (TraceLaunch::block_order, brute-force tail kernel).  The grid is dispatched in block order, and
// what idles the SMs at the end of a launch are expensive blocks that start late; ordering the pixel blocks by
// decreasing cost leaves the cheapest for the end.  The cost of a pixel block is measured, not guessed: the first
// kPilotSamples samples of the call are rendered by a pilot launch that counts the queries of every block, a radix sort
// of the ~10^4 counts gives the order, and the order is kept while the path heads stay valid.  The pilot's samples are
// the call's own first samples (accumulated as usual), so nothing is rendered twice and nothing changes in the image.
static int order_blocks(rt_ctx* c, TraceLaunch& L, int* launches) {
    uint32_t kPilotSamples = 1;   // 1, 2 and 4 samples order the blocks equally well (128-spp calls: 9.97 / 10.01 / 10.06 ms); the shortest pilot costs least
    if (const char* e = std::getenv("RTB200_TAIL_PILOT")) kPilotSamples = (uint32_t)std::max(1, std::atoi(e));
    L.block_cost = nullptr; L.block_order = nullptr;
    const uint32_t pixel_blocks = (L.head.n_stoch + kTraceThreads - 1u) / kTraceThreads;
    if (const char* e = std::getenv("RTB200_TAIL_LPT")) { if (std::atoi(e) == 0) return RT_OK; }
    if (L.bvh.n_leaves || pixel_blocks < 2048u) return RT_OK;                       // under ~2 waves: nothing to order
    if (!c->order_valid) {
        if (L.n < 16u * kPilotSamples) return RT_OK;                                // too short a call to pay for a pilot
        if (pixel_blocks > c->block_cap) {
            CU(cudaStreamSynchronize(c->stream));
            cudaFree(c->d_block_cost); cudaFree(c->d_block_sorted); cudaFree(c->d_block_iota); cudaFree(c->d_block_order); cudaFree(c->d_block_scratch);
            c->d_block_cost = c->d_block_sorted = c->d_block_iota = c->d_block_order = nullptr; c->d_block_scratch = nullptr; c->block_cap = 0;
            const uint32_t cap = (c->head_n_slots + kTraceThreads - 1u) / kTraceThreads;
            CU(cudaMalloc(&c->d_block_cost, cap * sizeof(uint32_t))); CU(cudaMalloc(&c->d_block_sorted, cap * sizeof(uint32_t)));
            CU(cudaMalloc(&c->d_block_iota, cap * sizeof(uint32_t))); CU(cudaMalloc(&c->d_block_order, cap * sizeof(uint32_t)));
            size_t bytes = 0;
            CU(sort_blocks_by_cost(c->d_block_cost, c->d_block_sorted, c->d_block_iota, c->d_block_order, cap, nullptr, bytes, c->stream));
            CU(cudaMalloc(&c->d_block_scratch, bytes ? bytes : 1));
            c->block_scratch_bytes = bytes; c->block_cap = cap;
        }
        CU(cudaMemsetAsync(c->d_block_cost, 0, pixel_blocks * sizeof(uint32_t), c->stream));
        TraceLaunch P = L;
        P.n = kPilotSamples; P.chunks = 1; P.chunk_spp = P.n; P.partial = nullptr; P.block_cost = c->d_block_cost;
        CU(c->prm.arith == RT_ARITH_STRICT ? launch_trace_tail_strict(P, c->stream) : launch_trace_tail_fast(P, c->stream));
        size_t bytes = c->block_scratch_bytes;
        CU(sort_blocks_by_cost(c->d_block_cost, c->d_block_sorted, c->d_block_iota, c->d_block_order, pixel_blocks, c->d_block_scratch, bytes, c->stream));
        c->order_valid = true;
        L.first += kPilotSamples; L.n -= kPilotSamples;
        *launches += 3;
    }
    L.block_order = c->d_block_order;
    return RT_OK;
}

static int check_renderable(rt_ctx* c, const float* cam);
static int do_render(rt_ctx* c, const float* cam, uint32_t first, uint32_t n, uint32_t accum_mode) {
    { int r = check_renderable(c, cam); if (r != RT_OK) return r; }
    if (n == 0) return RT_OK;
    Guard g(c);
    { int r = ensure_accel(c); if (r != RT_OK) return r; }
    TraceLaunch L{};
    L.bvh = dev_bvh(c);
    // waiting lanes that trigger a shading phase of the LBVH kernel: 16 with triangles in the tree, 12 without (64-spp calls on the
    // B200: C4 4.58 / 4.84 / 4.90 Grays/s at 8 / 12 / 16; C5 2.02 / 2.03 / 2.00)
    { const char* e = std::getenv("RTB200_BVH_REGEN"); L.bvh_regen = e ? (uint32_t)std::atoi(e) : 0u; }
    L.sc = c->sc; L.tex = c->tex; L.table = c->d_table;
    if (L.bvh_regen == 0u) L.bvh_regen = c->sc.n_tri ? 16u : 12u;
    std::memcpy(L.cam, cam, 12 * sizeof(float));
    L.W = c->W; L.H = c->H; L.first = first; L.n = n; L.depth = c->prm.max_depth;
    L.rx0 = c->rx0; L.ry0 = c->ry0; L.rx1 = c->rx1; L.ry1 = c->ry1; L.split_n = c->split_n; L.split_i = c->split_i;
    L.rng_mode = c->prm.rng_mode; L.seed = c->prm.seed; L.accum_mode = accum_mode; L.mesh_mode = c->prm.mesh_mode;
    L.target = accum_mode == RT_ACCUM_SUM ? c->d_accum : c->d_image;
    L.rays = c->d_rays;
    { const char* e = std::getenv("RTB200_TAIL_STEAL"); L.steal = (e && std::atoi(e) == 0) ? 0u : 1u; }   // sample stealing, TraceLaunch::steal
    L.queries = c->d_rays + 1;
    L.queue = reinterpret_cast<uint32_t*>(c->d_rays + 3);
    const bool use_head = c->prm.path_head == RT_HEAD_ON;
    if (!use_head && c->prm.arith == RT_ARITH_FAST) {
        int r = ensure_certified(c, cam);
        if (r != RT_OK) return r;
        L.certify_id = c->d_certify;
    }
    CU(cudaMemsetAsync(c->d_rays, 0, 2 * sizeof(unsigned long long), c->stream));
    CU(cudaMemsetAsync(c->d_rays + 3, 0, sizeof(unsigned long long), c->stream));
    CU(cudaEventRecord(c->ev0, c->stream));
    bool head_rebuilt = false;
    int extra_launches = 0;
    if (use_head) {
        int r = ensure_head(c, cam, &head_rebuilt);      // inside the timed region of this call
        if (r != RT_OK) return r;
        L.head.rec = c->d_head_rec; L.head.flag = c->d_head_flag; L.head.list = c->d_head_list;
        L.head.n_slots = c->head_n_slots; L.head.n_stoch = c->head_n_stoch;
        CU(cudaEventRecord(c->ev2, c->stream));
        { int r = order_blocks(c, L, &extra_launches); if (r != RT_OK) return r; }     // may render the first samples
        { int r = plan_chunks(c, L); if (r != RT_OK) return r; }
        const bool counting = c->count_work && L.bvh.n_leaves && c->prm.arith == RT_ARITH_FAST;
        if (counting) {
            if (!c->d_work) CU(cudaMalloc(&c->d_work, 8 * sizeof(unsigned long long)));
            CU(cudaMemsetAsync(c->d_work, 0, 8 * sizeof(unsigned long long), c->stream));
            L.work = c->d_work;
            CU(launch_trace_tail_fast_count(L, c->stream));
            unsigned long long w[8];
            CU(cudaMemcpyAsync(w, c->d_work, sizeof w, cudaMemcpyDeviceToHost, c->stream));
            CU(cudaStreamSynchronize(c->stream));
            for (int q = 0; q < 8; ++q) c->work_counts[q] = w[q];
        } else {
            CU(c->prm.arith == RT_ARITH_STRICT ? launch_trace_tail_strict(L, c->stream) : launch_trace_tail_fast(L, c->stream));
        }
        CU(launch_combine_chunks(L, c->stream));
    } else {
        CU(cudaEventRecord(c->ev2, c->stream));
        CU(c->prm.arith == RT_ARITH_STRICT ? launch_trace_strict(L, c->stream) : launch_trace_fast(L, c->stream));
    }
    CU(cudaEventRecord(c->ev1, c->stream));
    unsigned long long counters[2] = {0, 0};
    CU(cudaMemcpyAsync(counters, c->d_rays, sizeof counters, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    float ms = 0.0f;
    CU(cudaEventElapsedTime(&ms, c->ev0, c->ev1));
    // rays = path segments, what the reference's hitScene is called for: with path heads every sample of a pixel
    // owns the segments of the head although they were traced once.
    const unsigned long long rays = use_head ? (unsigned long long)n * c->head_segments + counters[1] : counters[0];
    c->stats.ms_last = ms;
    CU(cudaEventElapsedTime(&ms, c->ev2, c->ev1));
    c->stats.ms_trace_last = ms;
    c->stats.rays_last = rays;
    c->stats.queries_last = use_head ? counters[1] + (head_rebuilt ? c->head_segments : 0ull) : counters[0];
    c->stats.rays_total += rays;
    if (c->count_work && L.bvh.n_leaves) {       // tail-kernel queries of this call and their plane tests (every query scans the plane list)
        c->work_counts[5] = counters[1];
        c->work_counts[4] = counters[1] * (unsigned long long)c->sc.n_pla;
    }
    c->stats.kernel_launches += (L.chunks > 1u ? 2 : 1) + extra_launches;
    return RT_OK;
}

// The reference finishes its queue after every sample (src/raytracer.cpp:140,161), one launch + one host round trip per
// sample (184 us per 1080p sample through this API in round 1 against 70 us of kernel work).  Nothing can see the image
// between two renderAgain() calls unless it is read, so rt_render / rt_render_again only validate and record the call; the
// recorded samples are rendered by one launch of the batched kernel when the image, the statistics or any state they
// depend on is touched.  Bit-identical to eager execution (tests/test_gpu_parity.py::test_strict_image_bit_exact).
static const uint32_t kMaxPending = 4096;
static int flush_pending(rt_ctx* c) {
    if (!c || c->pend_n == 0) return RT_OK;
    const uint32_t first = c->pend_first, n = c->pend_n;
    c->pend_n = 0;
    return do_render(c, c->pend_cam, first, n, RT_ACCUM_REFERENCE);
}
static int check_renderable(rt_ctx* c, const float* cam) {
    if (!c || !cam) return fail(RT_ERR_INVALID, "render: null argument");
    if (!c->have_scene) return fail(RT_ERR_STATE, "render: no scene uploaded");
    if (c->prm.rng_mode != RT_RNG_COUNTER && !c->d_table) return fail(RT_ERR_STATE, "render: table RNG selected but no random table set");
    if (c->max_texture_used >= (int64_t)c->tex.layers)
        return fail(RT_ERR_STATE, "render: a textured mesh uses texture %lld but only %d layer(s) are uploaded",
                    (long long)c->max_texture_used, c->tex.layers);
    return RT_OK;
}
int rt_render(rt_ctx* c, const float cam[12]) {
    { int r = check_renderable(c, cam); if (r != RT_OK) return r; }
    c->pend_n = 0;                                   // recorded samples of the previous image would be overwritten unseen
    c->stats.rays_total = 0;
    if (std::getenv("RTB200_EAGER")) {
        int r = do_render(c, cam, 0u, 1u, RT_ACCUM_REFERENCE);
        if (r == RT_OK) { c->sample_counter = 0; c->stats.samples_done = 1; }
        return r;
    }
    std::memcpy(c->pend_cam, cam, 48);
    c->pend_first = 0; c->pend_n = 1;
    c->sample_counter = 0; c->stats.samples_done = 1;
    return RT_OK;
}
int rt_render_again(rt_ctx* c, const float cam[12]) {
    { int r = check_renderable(c, cam); if (r != RT_OK) return r; }
    if (std::getenv("RTB200_EAGER")) {
        FLUSH(c);
        int r = do_render(c, cam, c->sample_counter + 1u, 1u, RT_ACCUM_REFERENCE);
        if (r == RT_OK) { c->sample_counter += 1; c->stats.samples_done = c->sample_counter + 1; }
        return r;
    }
    if (c->pend_n && (std::memcmp(c->pend_cam, cam, 48) != 0 || c->pend_n >= kMaxPending)) FLUSH(c);
    if (c->pend_n == 0) { std::memcpy(c->pend_cam, cam, 48); c->pend_first = c->sample_counter + 1u; }
    c->pend_n += 1;
    c->sample_counter += 1; c->stats.samples_done = c->sample_counter + 1;
    return RT_OK;
}
int rt_render_samples(rt_ctx* c, const float cam[12], uint32_t first, uint32_t n, uint32_t accum_mode) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (accum_mode > 1) return fail(RT_ERR_INVALID, "bad accum_mode");
    FLUSH(c);
    if (accum_mode == RT_ACCUM_REFERENCE && first == 0) c->stats.rays_total = 0;
    int r = do_render(c, cam, first, n, accum_mode);
    if (r == RT_OK && n) {
        if (accum_mode == RT_ACCUM_REFERENCE) { c->sample_counter = first + n - 1; c->stats.samples_done = first + n; }
        else c->stats.samples_done += n;
    }
    return r;
}
int rt_clear_accum(rt_ctx* c) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    FLUSH(c);
    Guard g(c);
    CU(cudaMemsetAsync(c->d_accum, 0, (size_t)c->W * c->H * sizeof(float4), c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->stats.samples_done = 0; c->stats.rays_total = 0;
    return RT_OK;
}
int rt_invalidate_caches(rt_ctx* c) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    FLUSH(c);
    c->head_valid = false;
    c->certify_valid = false;
    return RT_OK;
}
int rt_finalize(rt_ctx* c, uint32_t spp) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (spp == 0) return fail(RT_ERR_INVALID, "rt_finalize: spp must be > 0");
    FLUSH(c);
    Guard g(c);
    CU(launch_finalize(c->d_accum, c->finalize_target ? c->finalize_target : c->d_image, c->W * c->H, spp, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->stats.kernel_launches += 1;
    return RT_OK;
}

int rt_read_image(rt_ctx* c, float* rgba) {
    if (!c || !rgba) return fail(RT_ERR_INVALID, "rt_read_image: null argument");
    FLUSH(c);
    Guard g(c);
    CU(cudaMemcpyAsync(rgba, c->d_image, (size_t)c->W * c->H * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return RT_OK;
}
int rt_read_accum(rt_ctx* c, float* rgba) {
    if (!c || !rgba) return fail(RT_ERR_INVALID, "rt_read_accum: null argument");
    FLUSH(c);
    Guard g(c);
    CU(cudaMemcpyAsync(rgba, c->d_accum, (size_t)c->W * c->H * sizeof(float4), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return RT_OK;
}
int rt_restore_image(rt_ctx* c, const float* rgba, uint32_t samples) {
    if (!c || !rgba) return fail(RT_ERR_INVALID, "rt_restore_image: null argument");
    if (samples == 0) return fail(RT_ERR_INVALID, "rt_restore_image: an image holds at least one sample");
    FLUSH(c);
    Guard g(c);
    CU(cudaMemcpyAsync(c->d_image, rgba, (size_t)c->W * c->H * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->sample_counter = samples - 1u;          // rt_render_again renders sample index sample_counter + 1
    c->stats.samples_done = samples;
    return RT_OK;
}
int rt_restore_accum(rt_ctx* c, const float* rgba, uint32_t samples) {
    if (!c || !rgba) return fail(RT_ERR_INVALID, "rt_restore_accum: null argument");
    FLUSH(c);
    Guard g(c);
    CU(cudaMemcpyAsync(c->d_accum, rgba, (size_t)c->W * c->H * sizeof(float4), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->stats.samples_done = samples;
    return RT_OK;
}
int rt_get_stats(rt_ctx* c, rt_stats* s) {
    if (!c || !s) return fail(RT_ERR_INVALID, "rt_get_stats: null argument");
    FLUSH(c);
    *s = c->stats;
    return RT_OK;
}
int rt_accum_device_ptr(rt_ctx* c, void** p, size_t* bytes) {
    if (!c || !p) return fail(RT_ERR_INVALID, "null argument");
    *p = c->d_accum;
    if (bytes) *bytes = (size_t)c->W * c->H * sizeof(float4);
    return RT_OK;
}
int rt_image_device_ptr(rt_ctx* c, void** p, size_t* bytes) {
    if (!c || !p) return fail(RT_ERR_INVALID, "null argument");
    FLUSH(c);
    *p = c->d_image;
    if (bytes) *bytes = (size_t)c->W * c->H * sizeof(float4);
    return RT_OK;
}
int rt_set_finalize_target(rt_ctx* c, void* p, size_t bytes) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (!p) { c->finalize_target = nullptr; return RT_OK; }
    if (bytes < (size_t)c->W * c->H * sizeof(float4)) return fail(RT_ERR_INVALID, "finalize target too small");
    if (((uintptr_t)p & 15u) != 0) return fail(RT_ERR_INVALID, "finalize target must be 16-byte aligned");
    {   // page-locked host memory: the address the DEVICE uses for it is not the host address on every system
        Guard g(c);
        cudaPointerAttributes at{};
        if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return fail(RT_ERR_INVALID, "finalize target: not an address CUDA knows"); }
        if (at.type == cudaMemoryTypeUnregistered || !at.devicePointer)
            return fail(RT_ERR_INVALID, "finalize target is pageable host memory the device cannot reach (page-lock it: cudaHostRegister / cudaHostAlloc, mapped)");
        p = at.devicePointer;
    }
    c->finalize_target = (float4*)p;
    return RT_OK;
}
int rt_set_accum_device_buffer(rt_ctx* c, void* p, size_t bytes) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (!p) { c->d_accum = c->d_accum_own; return RT_OK; }
    if (bytes < (size_t)c->W * c->H * sizeof(float4)) return fail(RT_ERR_INVALID, "accumulation buffer too small");
    if (((uintptr_t)p & 15u) != 0) return fail(RT_ERR_INVALID, "accumulation buffer must be 16-byte aligned");
    c->d_accum = (float4*)p;
    return RT_OK;
}

// ---- fused reduce + finalize over peer memory --------------------------------------------------
int rt_close_peers(rt_ctx* c) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    Guard g(c);
    for (int r = 0; r < kMaxPeers; ++r) {
        if (c->peer_opened[r] && c->peer_accum[r]) cudaIpcCloseMemHandle(c->peer_accum[r]);
        c->peer_accum[r] = nullptr; c->peer_opened[r] = false;
    }
    if (c->peer_image_opened && c->peer_image0) cudaIpcCloseMemHandle(c->peer_image0);
    c->peer_image0 = nullptr; c->peer_image_opened = false;
    c->peer_n = 0; c->peer_flags_ok = false;
    return RT_OK;
}
int rt_ipc_export(rt_ctx* c, void* accum_handle, void* image_handle) {
    if (!c || !accum_handle || !image_handle) return fail(RT_ERR_INVALID, "rt_ipc_export: null argument");
    FLUSH(c);
    static_assert(sizeof(cudaIpcMemHandle_t) == RT_IPC_HANDLE_BYTES, "CUDA IPC handle size");
    if (c->d_accum != c->d_accum_own) return fail(RT_ERR_STATE, "rt_ipc_export: the context accumulates into a caller buffer; only the context's own buffer can be exported");
    Guard g(c);
    cudaIpcMemHandle_t h;
    CU(cudaIpcGetMemHandle(&h, c->d_accum_own));
    std::memcpy(accum_handle, &h, sizeof h);
    CU(cudaIpcGetMemHandle(&h, c->d_image));
    std::memcpy(image_handle, &h, sizeof h);
    return RT_OK;
}
int rt_ipc_open_peers(rt_ctx* c, int n, int rank, const void* accum_handles, const void* image_handle0) {
    if (!c || !accum_handles || !image_handle0) return fail(RT_ERR_INVALID, "rt_ipc_open_peers: null argument");
    if (n < 1 || n > kMaxPeers || rank < 0 || rank >= n) return fail(RT_ERR_INVALID, "rt_ipc_open_peers: bad rank %d of %d (max %d)", rank, n, kMaxPeers);
    rt_close_peers(c);
    Guard g(c);
    for (int r = 0; r < n; ++r) {
        if (r == rank) { c->peer_accum[r] = c->d_accum_own; continue; }
        cudaIpcMemHandle_t h;
        std::memcpy(&h, (const char*)accum_handles + (size_t)r * sizeof h, sizeof h);
        CU(cudaIpcOpenMemHandle(&c->peer_accum[r], h, cudaIpcMemLazyEnablePeerAccess));
        c->peer_opened[r] = true;
    }
    if (rank == 0) c->peer_image0 = c->d_image;
    else {
        cudaIpcMemHandle_t h;
        std::memcpy(&h, image_handle0, sizeof h);
        CU(cudaIpcOpenMemHandle(&c->peer_image0, h, cudaIpcMemLazyEnablePeerAccess));
        c->peer_image_opened = true;
    }
    c->peer_n = n; c->peer_rank = rank;
    c->peer_flags_ok = true;                      // peer_epoch keeps counting: the flag words only ever grow
    return RT_OK;
}
int rt_set_peers_direct(rt_ctx* c, int n, int rank, void* const* accum_ptrs, void* image0) {
    if (!c || !accum_ptrs || !image0) return fail(RT_ERR_INVALID, "rt_set_peers_direct: null argument");
    if (n < 1 || n > kMaxPeers || rank < 0 || rank >= n) return fail(RT_ERR_INVALID, "rt_set_peers_direct: bad rank %d of %d (max %d)", rank, n, kMaxPeers);
    rt_close_peers(c);
    for (int r = 0; r < n; ++r) {
        if (!accum_ptrs[r]) return fail(RT_ERR_INVALID, "rt_set_peers_direct: null buffer for rank %d", r);
        c->peer_accum[r] = accum_ptrs[r];
    }
    c->peer_image0 = image0;
    c->peer_n = n; c->peer_rank = rank;
    return RT_OK;
}
static int finalize_peers(rt_ctx* c, uint32_t spp, bool sync, bool barriers);
int rt_finalize_peers(rt_ctx* c, uint32_t spp) { return finalize_peers(c, spp, true, false); }
int rt_finalize_peers_async(rt_ctx* c, uint32_t spp) { return finalize_peers(c, spp, false, false); }
int rt_finalize_peers_barrier_async(rt_ctx* c, uint32_t spp) { return finalize_peers(c, spp, false, true); }
static int peer_barrier(rt_ctx* c, int phase) {
    PeerBarrier B{};
    const size_t px = (size_t)c->W * c->H;
    for (int r = 0; r < c->peer_n; ++r) B.flags[r] = reinterpret_cast<uint32_t*>((float4*)c->peer_accum[r] + px);
    B.n_ranks = c->peer_n; B.my_rank = c->peer_rank; B.phase = phase; B.epoch = c->peer_epoch;
    B.timeout_cycles = 20000000000ll;                          // ~10 s: a dead rank turns into an error, not a hung device
    B.timed_out = c->d_barrier_timeout;
    CU(launch_peer_barrier(B, c->stream));
    c->stats.kernel_launches += 1;
    return RT_OK;
}
static int finalize_peers(rt_ctx* c, uint32_t spp, bool sync, bool barriers) {
    if (!c) return fail(RT_ERR_INVALID, "null context");
    if (spp == 0) return fail(RT_ERR_INVALID, "rt_finalize_peers: spp must be > 0");
    FLUSH(c);
    if (c->peer_n < 1) return fail(RT_ERR_STATE, "rt_finalize_peers: no peers set (rt_ipc_open_peers / rt_set_peers_direct)");
    Guard g(c);
    PeerFinalize P{};
    for (int r = 0; r < c->peer_n; ++r) P.accum[r] = (const float4*)c->peer_accum[r];
    P.image = c->finalize_target ? c->finalize_target : (float4*)c->peer_image0;
    P.n_ranks = c->peer_n;
    const long long px = (long long)c->W * c->H;
    P.first_pixel = (int)(px * c->peer_rank / c->peer_n);
    P.n_pixels = (int)(px * (c->peer_rank + 1) / c->peer_n) - P.first_pixel;
    P.inv_spp = 1.0f / (float)spp;
    if (barriers) {
        if (!c->peer_flags_ok) return fail(RT_ERR_STATE, "rt_finalize_peers_barrier_async: the device-side barrier needs peers opened with rt_ipc_open_peers");
        c->peer_epoch += 1;
        { int r = peer_barrier(c, 0); if (r != RT_OK) return r; }          // every rank's render is complete
    }
    CU(launch_finalize_peers(P, c->stream));
    if (barriers) { int r = peer_barrier(c, 1); if (r != RT_OK) return r; }  // every slice of rank 0's image is written
    if (sync) CU(cudaStreamSynchronize(c->stream));
    c->stats.kernel_launches += 1;
    return RT_OK;
}
int rt_peer_barrier_status(rt_ctx* c, int* timed_out) {
    if (!c || !timed_out) return fail(RT_ERR_INVALID, "rt_peer_barrier_status: null argument");
    Guard g(c);
    uint32_t v = 0;
    CU(cudaMemcpyAsync(&v, c->d_barrier_timeout, sizeof v, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    *timed_out = (int)v;
    return RT_OK;
}

// ---- probes ------------------------------------------------------------------------------------
extern "C++" {
namespace {
struct Scratch {   // RAII device scratch for the probe calls
    std::vector<void*> ptrs;
    ~Scratch() { for (void* p : ptrs) cudaFree(p); }
    template <class T> cudaError_t alloc(T** p, size_t n) {
        cudaError_t e = cudaMalloc((void**)p, (n ? n : 1) * sizeof(T));
        if (e == cudaSuccess) ptrs.push_back(*p);
        return e;
    }
};
ProbeLaunch base_probe(const rt_ctx* c) {
    ProbeLaunch L{};
    L.sc = c->sc; L.tex = c->tex; L.table = c->d_table;
    L.W = c->W; L.H = c->H; L.depth = c->prm.max_depth;
    L.rng_mode = c->prm.rng_mode; L.seed = c->prm.seed; L.mesh_mode = c->prm.mesh_mode;
    L.bvh = dev_bvh(c);
    return L;
}
}  // namespace
}  // extern "C++"

int rt_read_primary_ids(rt_ctx* c, const float cam[12], uint32_t* ids, float* t) {
    if (!c || !cam) return fail(RT_ERR_INVALID, "rt_read_primary_ids: null argument");
    if (!c->have_scene) return fail(RT_ERR_STATE, "no scene uploaded");
    Guard g(c);
    const size_t n = (size_t)c->W * c->H;
    Scratch s;
    { int r_ = ensure_accel(c); if (r_ != RT_OK) return r_; }
    ProbeLaunch L = base_probe(c);
    std::memcpy(L.cam, cam, 12 * sizeof(float));
    L.n = (uint32_t)n;
    CU(s.alloc(&L.out_id, n));
    CU(s.alloc(&L.out_t, n));
    if (c->prm.arith == RT_ARITH_FAST) {
        int r = ensure_certified(c, cam);
        if (r != RT_OK) return r;
        L.certify_id = c->d_certify;
    }
    CU(c->prm.arith == RT_ARITH_STRICT ? launch_probe_hits_strict(L, c->stream) : launch_probe_hits_fast(L, c->stream));
    if (ids) CU(cudaMemcpyAsync(ids, L.out_id, n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (t) CU(cudaMemcpyAsync(t, L.out_t, n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->stats.kernel_launches += 1;
    return RT_OK;
}

int rt_trace_rays(rt_ctx* c, const float* rays6, uint32_t n, uint32_t* out_id, float* out_t) {
    if (!c || !rays6) return fail(RT_ERR_INVALID, "rt_trace_rays: null argument");
    if (!c->have_scene) return fail(RT_ERR_STATE, "no scene uploaded");
    if (n == 0) return RT_OK;
    Guard g(c);
    Scratch s;
    { int r_ = ensure_accel(c); if (r_ != RT_OK) return r_; }
    ProbeLaunch L = base_probe(c);
    float* d_rays6 = nullptr;
    CU(s.alloc(&d_rays6, (size_t)n * 6));
    CU(cudaMemcpyAsync(d_rays6, rays6, (size_t)n * 24, cudaMemcpyHostToDevice, c->stream));
    L.rays6 = d_rays6; L.n = n;
    CU(s.alloc(&L.out_id, n));
    CU(s.alloc(&L.out_t, n));
    CU(c->prm.arith == RT_ARITH_STRICT ? launch_probe_hits_strict(L, c->stream) : launch_probe_hits_fast(L, c->stream));
    if (out_id) CU(cudaMemcpyAsync(out_id, L.out_id, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    if (out_t) CU(cudaMemcpyAsync(out_t, L.out_t, (size_t)n * 4, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->stats.kernel_launches += 1;
    return RT_OK;
}

// No counterpart in the reference; added: rt_trace_rays for callers that may not
// allocate (CUDA graph capture, serving loops).  Exact arithmetic whatever rt_params.arith says; at most RT_NOALLOC_MAX_RAYS
// rays; calls no cudaMalloc / cudaFree — so it does not BUILD anything either: the intersection path of the scene (brute
// force, or the LBVH) must be the one the parameters ask for already, which any render call or rt_trace_rays establishes.
int rt_trace_rays_noalloc(rt_ctx* c, const float* rays6, uint32_t n, uint32_t* out_id, float* out_t) {
    if (!c || !rays6) return fail(RT_ERR_INVALID, "rt_trace_rays_noalloc: null argument");
    static_assert(RT_NOALLOC_MAX_RAYS == kNoAllocMaxRays, "header and kernel file agree on the size of the static scratch");
    if (n > RT_NOALLOC_MAX_RAYS)
        return fail(RT_ERR_INVALID, "rt_trace_rays_noalloc: %u rays exceed the static scratch of %u rays; split the batch or use rt_trace_rays", n, RT_NOALLOC_MAX_RAYS);
    if (!c->have_scene) return fail(RT_ERR_STATE, "no scene uploaded");
    if (n == 0) return RT_OK;
    bool want_bvh = c->prm.accel == RT_ACCEL_BVH || (c->prm.accel == RT_ACCEL_AUTO && (c->staged_bytes > kMaxStagedBytes || c->n_finite > 512u));
    if (c->n_finite == 0) want_bvh = false;
    if (!want_bvh && c->staged_bytes > kMaxStagedBytes)
        return fail(RT_ERR_UNSUPPORTED, "scene needs %zu bytes of shared memory for the brute-force path (max %zu)", c->staged_bytes, kMaxStagedBytes);
    if (want_bvh && c->sc.n_tri > 0 && c->prm.mesh_mode != RT_MESH_NEAREST)
        return fail(RT_ERR_UNSUPPORTED, "the LBVH path answers the nearest-front-facing mesh rule only: set mesh_mode = RT_MESH_NEAREST");
    if (want_bvh != (c->bvh.n_leaves != 0))
        return fail(RT_ERR_STATE, "rt_trace_rays_noalloc: the intersection structure of this scene is not in the state the parameters ask for (LBVH %s); "
                                  "building or dropping it allocates — call rt_trace_rays or a render call once first", want_bvh ? "not built" : "still present");
    Guard g(c);
    if (!g.ok) return fail(RT_ERR_CUDA, "rt_trace_rays_noalloc: cannot make device %d current", c->device);
    ProbeLaunch L = base_probe(c);
    CU(launch_probe_hits_noalloc(L, rays6, n, out_id, out_t, c->stream));
    c->stats.kernel_launches += 1;
    return RT_OK;
}

int rt_bounce_rays(rt_ctx* c, float* rays6, uint32_t n, const uint32_t* sample, const uint32_t* bounce,
                   const uint32_t* gx, const uint32_t* gy, float* col3, uint8_t* ended) {
    if (!c || !rays6 || !sample || !bounce || !gx || !gy || !col3 || !ended) return fail(RT_ERR_INVALID, "rt_bounce_rays: null argument");
    if (!c->have_scene) return fail(RT_ERR_STATE, "no scene uploaded");
    if (c->prm.rng_mode != RT_RNG_COUNTER && !c->d_table) return fail(RT_ERR_STATE, "table RNG selected but no random table set");
    if (n == 0) return RT_OK;
    Guard g(c);
    Scratch s;
    { int r_ = ensure_accel(c); if (r_ != RT_OK) return r_; }
    ProbeLaunch L = base_probe(c);
    float *d_in = nullptr, *d_out = nullptr, *d_col = nullptr;
    uint32_t *d_s = nullptr, *d_b = nullptr, *d_x = nullptr, *d_y = nullptr;
    uint8_t* d_e = nullptr;
    CU(s.alloc(&d_in, (size_t)n * 6)); CU(s.alloc(&d_out, (size_t)n * 6)); CU(s.alloc(&d_col, (size_t)n * 3));
    CU(s.alloc(&d_s, n)); CU(s.alloc(&d_b, n)); CU(s.alloc(&d_x, n)); CU(s.alloc(&d_y, n)); CU(s.alloc(&d_e, n));
    CU(cudaMemcpyAsync(d_in, rays6, (size_t)n * 24, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_s, sample, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_b, bounce, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_x, gx, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(d_y, gy, (size_t)n * 4, cudaMemcpyHostToDevice, c->stream));
    L.rays6 = d_in; L.rays6_out = d_out; L.n = n; L.sample = d_s; L.bounce = d_b; L.gx = d_x; L.gy = d_y;
    L.col3 = d_col; L.ended = d_e;
    CU(c->prm.arith == RT_ARITH_STRICT ? launch_probe_bounce_strict(L, c->stream) : launch_probe_bounce_fast(L, c->stream));
    CU(cudaMemcpyAsync(rays6, d_out, (size_t)n * 24, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(col3, d_col, (size_t)n * 12, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaMemcpyAsync(ended, d_e, (size_t)n, cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->stats.kernel_launches += 1;
    return RT_OK;
}

int rt_measure_fp32_peak(int device, double* gslots_per_s, double* sm_clock_mhz) {
    if (!gslots_per_s) return fail(RT_ERR_INVALID, "rt_measure_fp32_peak: null output");
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || device < 0 || device >= n_dev)
        return fail(RT_ERR_CUDA, "rt_measure_fp32_peak: no such CUDA device");
    int prev = 0;
    CU(cudaGetDevice(&prev));
    CU(cudaSetDevice(device));
    int sms = 0, khz = 0;
    CU(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device));
    CU(cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, device));
    float* sink = nullptr;
    CU(cudaMalloc(&sink, 4));
    cudaEvent_t e0, e1;
    CU(cudaEventCreate(&e0)); CU(cudaEventCreate(&e1));
    const int blocks = sms * 8, threads = 256, iters = 4096;
    double best = 0.0;
    for (int rep = 0; rep < 5; ++rep) {
        CU(cudaEventRecord(e0, 0));
        CU(launch_fp32_peak(sink, iters, blocks, threads, 0));
        CU(cudaEventRecord(e1, 0));
        CU(cudaEventSynchronize(e1));
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, e0, e1));
        const double slots = (double)blocks * threads * (double)iters * 8.0 * 16.0;
        if (rep > 0) best = std::fmax(best, slots / (ms * 1e-3) / 1e9);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    cudaSetDevice(prev);
    *gslots_per_s = best;
    if (sm_clock_mhz) *sm_clock_mhz = khz / 1000.0;
    return RT_OK;
}

}  // extern "C"
