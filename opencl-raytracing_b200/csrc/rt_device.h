// csrc/rt_device.h — device-side scene layout and launch parameters shared by the host layer and the kernels.
//
// The reference keeps arrays of structs in global memory and chases pointers per primitive
// (kernels/raytracer.cl:74-103).  Here the intersection data is a flat structure-of-float4 "test stream"
// that every block stages into shared memory once and then reads with warp-uniform (broadcast) LDS.128;
// data only needed for the ONE winning primitive of a ray ("shade" arrays) stays in global memory.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "rt_bvh.h"

namespace rtb {

// ---- test stream (float4 units), in hitScene order: spheres, planes, lenses, then mesh instances -------------
//   sphere   : 1 x float4  (c.x, c.y, c.z, r*r)
//   plane    : 2 x float4  (p.xyz, -) (n.xyz, -)
//   lens     : 2 x float4  (p1.xyz, r1*r1) (p2.xyz, r2*r2)
//   triangle : 3 x float4  (A.xyz, e1.x) (e1.y, e1.z, e2.x, e2.y) (e2.z, n.x, n.y, n.z)
//              e1 = B - A, e2 = C - A, n = normalize(cross(e1, e2)) precomputed on the host with the same
//              single-rounded float operations the reference performs per ray (kernels/raytracer.cl:264-265,285)
struct alignas(16) MeshInst {   // one per (model, mesh) pair in hitModel order; read on the device as 3 x uint4
    uint32_t tri_first;    // index into the triangle part of the stream
    uint32_t tri_count;
    uint32_t face_id_base; // reference-global face number of the first triangle (index_anchor / 3)
    uint32_t mat_id;       // the owning model's material (kernels/raytracer.cl:314)
    uint32_t texture_id;   // kernels/raytracer.cl:296
    uint32_t _pad[3];
    float    bound[4];     // conservative bounding sphere (c.xyz, r) of the mesh instance
};

struct DevScene {
    const float4* stream;      // test stream in global memory (staged to shared memory per block)
    uint32_t stream_f4;        // number of float4 in the stream
    uint32_t n_sph, n_pla, n_len, n_tri, n_mesh;
    uint32_t off_pla, off_len, off_tri;   // float4 offsets of each section
    const MeshInst* mesh;      // n_mesh
    // shade arrays (global)
    const float4* sph_shade;   // (c.xyz, r)
    const uint32_t* sph_mat;
    const float4* pla_normal;  // (n.xyz, -)
    const uint32_t* pla_mat;
    const float4* len_shade;   // 2 per lens: (p1.xyz, r1) (p2.xyz, r2)
    const uint32_t* len_mat;
    const float2* tri_uv;      // 3 per triangle (uvA, uvB, uvC)
    const uint32_t* tri_mesh;  // mesh instance of each triangle
    const float* tri_pd;       // per triangle: plane offset n.A minus a safety margin (conservative front-side filter); padded to x4
    const float4* mat;         // 2 per material: (type as int bits, extra, -, -) (color.rgb, -)
    uint32_t n_mat;
};

struct DevTexture {
    const float4* texels;      // [layer][row][col] RGBA f32
    int w, h, layers;
};

// Path heads.  The reference shoots the SAME primary ray for every sample of a pixel (no jitter,
// kernels/raytracer.cl:500-505,517-522), and mirrors, refractive surfaces, lights and misses consume no random
// numbers (kernels/raytracer.cl:454-479), so the path of a pixel is identical for all samples up to its first hit on
// a diffuse / textured / dielectric surface.  That head is computed once per (scene, camera) by k_head_* in Exact
// arithmetic; the render kernels restart every sample from the stored state.
// Slots are pixel tiles in launch order: slot = tile * 32 + lane, tile = 8 x 4 pixels.
//   rec[0 * n_slots + s] = (o.xyz, hit t)     rec[1 * n_slots + s] = (d.xyz, hit bu)
//   rec[2 * n_slots + s] = (out.rgb, hit bv)  rec[3 * n_slots + s] = (bounce index, hit id, hit aux, material id) as bits
//   flag[s]: 0 outside the frame, 1 terminal (rec[2].rgb is the colour of every sample, rec[3].x the segment count),
//            2 stochastic (restart state of the first random-consuming hit).
// list[0, n_stoch) = stochastic slots ordered by (material type, primitive kind[, primitive]) of the head hit, slot order
// inside a class, so that the lanes of a warp of k_trace_tail_* shade the same kind of surface on the first bounce of
// every sample (84 % of the trips on C2); list[n_stoch, n_slots) = all other slots.
struct DevHead {
    const float4* rec;
    const uint8_t* flag;
    const uint32_t* list;
    uint32_t n_slots, n_stoch;
};
constexpr uint8_t kHeadDead = 0, kHeadTerminal = 1, kHeadStochastic = 2;

struct TraceLaunch {
    DevScene sc;
    DevTexture tex;
    const float* table;        // 400 000 floats (table RNG) or nullptr
    float cam[12];             // src/camera.cpp:94-118
    int W, H;
    int rx0, ry0, rx1, ry1;    // render region [rx0, rx1) x [ry0, ry1) of the frame (rt_set_region); pixels outside are untouched
    uint32_t split_n, split_i; // pixel split (rt_set_pixel_split): slot group g (kTraceThreads slots = four 8 x 4 tiles) is rendered iff g % split_n == split_i
    uint32_t first, n;         // samples [first, first + n)
    uint32_t depth;
    uint32_t rng_mode, seed, accum_mode, mesh_mode;
    float4* target;            // image (reference chain) or accumulation buffer (sum)
    unsigned long long* rays;  // device counter
    const uint32_t* certify_id; // fast arithmetic: Exact primary-hit id per pixel (nullptr = no certification)
    DevBvh bvh;                // n_leaves == 0: brute-force path over the staged scene
    uint32_t bvh_regen;        // LBVH kernel: waiting lanes that trigger a shading/regeneration phase (0 = default 8)
    DevHead head;              // path heads of the frame (k_trace_tail_* only)
    unsigned long long* queries; // device counter: nearest-hit queries executed (k_trace_tail_*)
    uint32_t* queue;           // device counter, zero at launch: next entry of head.list to render (k_trace_tail_*_bvh)
    uint32_t tail_blocks;      // persistent blocks of k_trace_tail_*_bvh (set by the launcher)
    // Sample chunks of k_trace_tail_* (brute-force path, linear sum): a work item is (stochastic pixel, chunk of
    // chunk_spp samples); chunk c of list entry j leaves its sum in partial[c * n_stoch + j] and k_combine_chunks adds
    // the chunks of a pixel in chunk order.  chunks <= 1: one item per pixel, summed straight into `target`.
    uint32_t chunks, chunk_spp;
    float4* partial;
    // Sample stealing (k_trace_tail_fast, linear sums): a lane out of samples takes half of the unstarted samples of the busiest
    // lane of its warp.  Reproducible for the same call; the ORDER in which a pixel's samples are added then depends on the other
    // pixels of the warp, so sums of different partitions of a frame (regions, pixel split, block order) agree to rounding, not bit
    // for bit.  RTB200_TAIL_STEAL=0 turns it off.
    uint32_t steal;
    // Longest blocks first (k_trace_tail_*, brute-force path).  A "pixel block" is kTraceThreads consecutive entries of
    // the stochastic list; block_cost[pb] (written when non-null) = queries the block executed; block_order (when
    // non-null) = pixel blocks by decreasing cost: block i of the grid renders chunk i % chunks of pixel block
    // block_order[i / chunks] instead of the blocks in list order.
    uint32_t* block_cost;
    const uint32_t* block_order;
    // Instrumented LBVH kernel (k_trace_tail_fast_bvh_count, untimed): [0] node visits (two box tests each), leaf tests of
    // [1] spheres, [2] lenses, [3] triangles; [6] node visits of flared rays, [7] flared queries.  nullptr in the timed kernels.
    unsigned long long* work;
};

// k_head_*: the sample-independent head of every pixel's path (see rt_kernels.cuh, "path heads").
struct HeadLaunch {
    DevScene sc;
    float cam[12];
    int W, H;
    int rx0, ry0, rx1, ry1;    // render region: slots outside it are flagged kHeadDead
    uint32_t split_n, split_i; // pixel split: slot groups of other parts are flagged kHeadDead
    uint32_t depth, mesh_mode;
    DevBvh bvh;
    float4* rec;               // 4 x n_slots
    uint8_t* flag;             // n_slots
    uint32_t* key;             // n_slots: sort key of the slot (class of its head), stochastic slots first
    uint32_t key_mask;         // bits of the class that take part (0: stochastic slots stay in slot order)
    uint32_t n_slots;
    unsigned long long* segments;   // device counters: [0] sum of head segments over the live pixels, [1] stochastic slots
};

struct ProbeLaunch {           // rt_trace_rays / rt_read_primary_ids / rt_bounce_rays
    DevScene sc;
    DevTexture tex;
    const float* table;
    float cam[12];
    int W, H;
    uint32_t depth, rng_mode, seed, mesh_mode;
    const float* rays6;        // n x 6 (nullptr -> primary rays of the W x H frame)
    float* rays6_out;
    uint32_t n;
    uint32_t* out_id;
    float* out_t;
    const uint32_t *sample, *bounce, *gx, *gy;
    float* col3;
    uint8_t* ended;
    const uint32_t* certify_id; // fast arithmetic: Exact hit id per ray (nullptr = no certification)
    DevBvh bvh;
};

static_assert(sizeof(MeshInst) == 48, "MeshInst is three 16-byte words");

constexpr uint32_t kIdMiss = 0xFFFFFFFFu;
constexpr uint32_t kKindSphere = 1u, kKindPlane = 2u, kKindLens = 3u, kKindTriangle = 4u;
#ifndef RT_TRACE_THREADS
#define RT_TRACE_THREADS 128
#endif
constexpr int kTraceThreads = RT_TRACE_THREADS;          // 4 warps, each an 8 x 4 pixel tile
constexpr size_t kMaxStagedBytes = 160 * 1024;      // staged scene; the kernels add up to 8 KB of per-lane state
constexpr size_t kMaxBlockSmemBytes = 227 * 1024;   // opt-in limit of a block on sm_100

// launchers, one set per arithmetic flavour (rt_kernels_strict.cu / rt_kernels_fast.cu)
cudaError_t launch_trace_strict(const TraceLaunch& p, cudaStream_t s);
cudaError_t launch_trace_fast(const TraceLaunch& p, cudaStream_t s);
cudaError_t launch_probe_hits_strict(const ProbeLaunch& p, cudaStream_t s);
// rt_trace_rays_noalloc: Exact-arithmetic probe whose device scratch is a static device array (rt_kernels_strict.cu)
constexpr uint32_t kNoAllocMaxRays = 65536u;
cudaError_t launch_probe_hits_noalloc(ProbeLaunch p, const float* h_rays6, uint32_t n, uint32_t* h_id, float* h_t, cudaStream_t s);
cudaError_t launch_probe_hits_fast(const ProbeLaunch& p, cudaStream_t s);
cudaError_t launch_probe_bounce_strict(const ProbeLaunch& p, cudaStream_t s);
cudaError_t launch_probe_bounce_fast(const ProbeLaunch& p, cudaStream_t s);
cudaError_t launch_head(const HeadLaunch& p, cudaStream_t s);                 // Exact arithmetic, both arms
cudaError_t launch_trace_tail_strict(const TraceLaunch& p, cudaStream_t s);
cudaError_t launch_trace_tail_fast(const TraceLaunch& p, cudaStream_t s);
cudaError_t launch_trace_tail_fast_count(const TraceLaunch& p, cudaStream_t s);   // LBVH scenes only: fills TraceLaunch::work
cudaError_t launch_combine_chunks(const TraceLaunch& p, cudaStream_t s);
cudaError_t sort_blocks_by_cost(const uint32_t* cost, uint32_t* cost_sorted, uint32_t* iota, uint32_t* order, uint32_t n,
                                void* scratch, size_t& scratch_bytes, cudaStream_t s);
// list = slots ordered by their 16-bit head key (stable CUB radix sort; `slots` = 0..n_slots-1, made by fill_iota)
constexpr uint32_t kHeadKeyBits = 16, kHeadKeyOther = 0xFFFFu;
cudaError_t sort_heads(const uint32_t* key, uint32_t* key_sorted, const uint32_t* slots, uint32_t* list, uint32_t n_slots,
                       void* scratch, size_t& scratch_bytes, cudaStream_t s);
cudaError_t fill_iota(uint32_t* v, uint32_t n, cudaStream_t s);
cudaError_t launch_finalize(const float4* accum, float4* image, int n_pixels, uint32_t spp, cudaStream_t s);
constexpr int kMaxPeers = 16;
struct PeerFinalize {          // k_finalize_peers: sum of n_ranks accumulation buffers -> image of rank 0
    const float4* accum[kMaxPeers];
    float4* image;
    int n_ranks;
    int first_pixel, n_pixels;
    float inv_spp;
};
cudaError_t launch_finalize_peers(const PeerFinalize& p, cudaStream_t s);
// Barrier of the ranks of the sample split on the device: arrival flags in peer memory instead of two NCCL collectives
// around k_finalize_peers.  Every rank owns kPeerFlagWords words behind its accumulation buffer (visible to the peers through
// the same CUDA IPC mapping); phase 0 = "my render is complete", phase 1 = "my slice of the image is written".
constexpr int kPeerFlagWords = 2 * kMaxPeers;
struct PeerBarrier {
    uint32_t* flags[kMaxPeers];   // flags[r] = rank r's flag words (peer pointers; [my_rank] is local)
    int n_ranks, my_rank, phase;
    uint32_t epoch;               // grows by one per barrier round; flags only ever grow
    long long timeout_cycles;     // a rank that never arrives must not hang the device
    uint32_t* timed_out;          // local device word, set when the wait gave up
};
cudaError_t launch_peer_barrier(const PeerBarrier& b, cudaStream_t s);
cudaError_t launch_fp32_peak(float* sink, int iters, int blocks, int threads, cudaStream_t s);

}  // namespace rtb
