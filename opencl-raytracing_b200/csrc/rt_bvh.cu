// csrc/rt_bvh.cu — LBVH construction on the GPU (Lauterbach et al. 2009 / Karras 2012):
//   k_prim_bounds   one thread per primitive: AABB (slightly inflated, conservative) + centroid
//   cub::DeviceReduce  scene bounds of the centroids
//   k_morton        63-bit Morton code of the centroid (21 bits per axis); equal codes are split by sorted position
//   cub::DeviceRadixSort  sort keys, carrying the primitive number
//   k_hierarchy     Karras' binary radix tree: one thread per internal node finds its range and split
//   k_refit         one thread per leaf walks to the root; the second arrival at a node (atomic counter) merges
//   k_emit_nodes    traversal layout: each node stores the boxes of its two children
#include "rt_bvh.h"
#include <cuda_fp16.h>

#include <cfloat>
#include <vector>
#include <cub/cub.cuh>

namespace rtb {
namespace {

struct Box { float lo[3], hi[3]; };

__device__ __forceinline__ void inflate(Box& b) {
    // conservative padding: primitives' exact hits are computed in rounded float arithmetic
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float pad = 1e-5f * fmaxf(fmaxf(fabsf(b.lo[a]), fabsf(b.hi[a])), 1.0f) + 1e-5f * (b.hi[a] - b.lo[a]);
        b.lo[a] -= pad; b.hi[a] += pad;
    }
}

// primitive p in [0, n_sph + n_len + n_tri): spheres, then lenses, then triangles
__global__ void k_prim_bounds(BvhBuildInput in, uint32_t n, Box* boxes, float4* centroids, uint2* leaf_info) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    Box b;
    uint2 info;
    if (p < in.n_sph) {
        const float4 s = in.stream[p];
        const float r = sqrtf(s.w) * 1.0001f;                       // stream holds r*r
        b.lo[0] = s.x - r; b.lo[1] = s.y - r; b.lo[2] = s.z - r; b.hi[0] = s.x + r; b.hi[1] = s.y + r; b.hi[2] = s.z + r;
        info = make_uint2((1u << 28) | p, p);
    } else if (p < in.n_sph + in.n_len) {
        const uint32_t i = p - in.n_sph;
        const float4 l1 = in.stream[in.off_len + 2 * i], l2 = in.stream[in.off_len + 2 * i + 1];
        const float r1 = sqrtf(l1.w) * 1.0001f, r2 = sqrtf(l2.w) * 1.0001f;
        const float c1[3] = {l1.x, l1.y, l1.z}, c2[3] = {l2.x, l2.y, l2.z};
        for (int a = 0; a < 3; ++a) {                               // lens = ball1 ∩ ball2 ⊂ box1 ∩ box2
            b.lo[a] = fmaxf(c1[a] - r1, c2[a] - r2);
            b.hi[a] = fminf(c1[a] + r1, c2[a] + r2);
            if (b.hi[a] < b.lo[a]) { const float m = 0.5f * (b.hi[a] + b.lo[a]); b.lo[a] = b.hi[a] = m; }
        }
        info = make_uint2((3u << 28) | i, i);
    } else {
        const uint32_t t = p - in.n_sph - in.n_len;
        const float4 q0 = in.stream[in.off_tri + 3 * t], q1 = in.stream[in.off_tri + 3 * t + 1], q2 = in.stream[in.off_tri + 3 * t + 2];
        const float A[3] = {q0.x, q0.y, q0.z}, e1[3] = {q0.w, q1.x, q1.y}, e2[3] = {q1.z, q1.w, q2.x};
        for (int a = 0; a < 3; ++a) {
            const float B = A[a] + e1[a], C = A[a] + e2[a];
            b.lo[a] = fminf(A[a], fminf(B, C));
            b.hi[a] = fmaxf(A[a], fmaxf(B, C));
        }
        info = make_uint2((4u << 28) | in.tri_face_id[t], t);
    }
    inflate(b);
    boxes[p] = b;
    centroids[p] = make_float4(0.5f * (b.lo[0] + b.hi[0]), 0.5f * (b.lo[1] + b.hi[1]), 0.5f * (b.lo[2] + b.hi[2]), 0.0f);
    leaf_info[p] = info;
}

struct Bounds3 { float lo[3], hi[3]; };
struct BoundsOp {
    __device__ Bounds3 operator()(const Bounds3& a, const Bounds3& b) const {
        Bounds3 r;
        for (int i = 0; i < 3; ++i) { r.lo[i] = fminf(a.lo[i], b.lo[i]); r.hi[i] = fmaxf(a.hi[i], b.hi[i]); }
        return r;
    }
};
__global__ void k_centroid_bounds_in(const float4* c, uint32_t n, Bounds3* out) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    Bounds3 b;
    b.lo[0] = b.hi[0] = c[p].x; b.lo[1] = b.hi[1] = c[p].y; b.lo[2] = b.hi[2] = c[p].z;
    out[p] = b;
}

// 21 bits -> every third bit of a 63-bit word
__device__ __forceinline__ unsigned long long expand21(unsigned long long v) {
    v &= 0x1FFFFFull;
    v = (v | (v << 32)) & 0x1F00000000FFFFull;
    v = (v | (v << 16)) & 0x1F0000FF0000FFull;
    v = (v | (v << 8)) & 0x100F00F00F00F00Full;
    v = (v | (v << 4)) & 0x10C30C30C30C30C3ull;
    v = (v | (v << 2)) & 0x1249249249249249ull;
    return v;
}
__global__ void k_morton(const float4* centroids, uint32_t n, const Bounds3* scene, unsigned long long* keys, uint32_t* vals) {
    const uint32_t p = blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= n) return;
    const Bounds3 sb = *scene;
    const float c[3] = {centroids[p].x, centroids[p].y, centroids[p].z};
    // 63-bit codes (21 bits per axis): the scene bounds are dominated by a few huge primitives (light sphere, radius
    // 300), so with 10 bits per axis thousands of the small triangles of a mesh shared one cell and the bottom of
    // the tree degenerated into primitive order.
    unsigned long long q[3];
    for (int a = 0; a < 3; ++a) {
        const float ext = fmaxf(sb.hi[a] - sb.lo[a], 1e-20f);
        const double u = fmin(fmax(((double)c[a] - (double)sb.lo[a]) / (double)ext, 0.0), 1.0);
        q[a] = (unsigned long long)fmin(u * 2097152.0, 2097151.0);
    }
    keys[p] = (expand21(q[0]) << 2) | (expand21(q[1]) << 1) | expand21(q[2]);
    vals[p] = p;
}

// common prefix of the (key, position) pairs: equal codes are told apart by their sorted position (Karras 2012 §4)
__device__ __forceinline__ int delta(const unsigned long long* k, int n, int i, int j) {
    if (j < 0 || j >= n) return -1;
    const unsigned long long x = k[i] ^ k[j];
    return x ? __clzll((long long)x) : 64 + __clz(i ^ j);
}

// Karras 2012, "Maximizing Parallelism in the Construction of BVHs, Octrees, and k-d Trees", Fig. 4.
// children >= 0: internal; < 0: leaf ~slot.  parent[] indexed [0, n-1) internal, [n-1, 2n-1) leaves.
__global__ void k_hierarchy(const unsigned long long* keys, int n, int2* children, int* parent) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1) return;
    const int d = (delta(keys, n, i, i + 1) - delta(keys, n, i, i - 1)) >= 0 ? 1 : -1;
    const int dmin = delta(keys, n, i, i - d);
    int lmax = 2;
    while (delta(keys, n, i, i + lmax * d) > dmin) lmax <<= 1;
    int l = 0;
    for (int t = lmax >> 1; t >= 1; t >>= 1)
        if (delta(keys, n, i, i + (l + t) * d) > dmin) l += t;
    const int j = i + l * d;
    const int dnode = delta(keys, n, i, j);
    int s = 0, t = l;
    do {
        t = (t + 1) >> 1;
        if (delta(keys, n, i, i + (s + t) * d) > dnode) s += t;
    } while (t > 1);
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const int left = (lo == gamma) ? ~gamma : gamma;
    const int right = (hi == gamma + 1) ? ~(gamma + 1) : gamma + 1;
    children[i] = make_int2(left, right);
    parent[left < 0 ? (n - 1 + ~left) : left] = i;
    parent[right < 0 ? (n - 1 + ~right) : right] = i;
    if (i == 0) parent[0] = -1;
}

__device__ __forceinline__ Box child_box(int c, const Box* node_box, const Box* prim_box, const uint32_t* sorted) {
    return c < 0 ? prim_box[sorted[~c]] : node_box[c];
}

// curved[node] = 1 when the subtree holds a sphere or a lens: their binary32 tests report hits OUTSIDE the primitive from
// far away (rt_kernels.cuh, slab<true>), so only those subtrees need the distance-proportional box inflation.
// COHERENT: inside k_refit the value was written by another thread of this launch; it is read through L2 (ld.cg),
// because an L1 line fetched earlier for a neighbouring entry may predate the write.
template <bool COHERENT>
__device__ __forceinline__ uint8_t child_curved(int c, const uint8_t* curved, const uint2* leaf_info, const uint32_t* sorted) {
    if (c < 0) return (uint8_t)((leaf_info[sorted[~c]].x >> 28) != 4u);
    return COHERENT ? __ldcg(curved + c) : curved[c];
}
__device__ __forceinline__ Box node_box_coherent(const Box* node_box, int c) {
    Box b;
    const float* f = reinterpret_cast<const float*>(node_box + c);
#pragma unroll
    for (int k = 0; k < 3; ++k) { b.lo[k] = __ldcg(f + k); b.hi[k] = __ldcg(f + 3 + k); }
    return b;
}
__global__ void k_refit(int n, const int2* children, const int* parent, const uint32_t* sorted, const Box* prim_box,
                        Box* node_box, unsigned int* visits, const uint2* leaf_info, uint8_t* curved) {
    const int leaf = blockIdx.x * blockDim.x + threadIdx.x;
    if (leaf >= n) return;
    int node = parent[n - 1 + leaf];
    while (node >= 0) {
        __threadfence();
        if (atomicAdd(&visits[node], 1u) == 0u) return;            // first arrival: the sibling subtree is not done yet
        const int2 ch = children[node];
        const Box a = ch.x < 0 ? prim_box[sorted[~ch.x]] : node_box_coherent(node_box, ch.x);
        const Box b = ch.y < 0 ? prim_box[sorted[~ch.y]] : node_box_coherent(node_box, ch.y);
        Box m;
        for (int k = 0; k < 3; ++k) { m.lo[k] = fminf(a.lo[k], b.lo[k]); m.hi[k] = fmaxf(a.hi[k], b.hi[k]); }
        node_box[node] = m;
        curved[node] = child_curved<true>(ch.x, curved, leaf_info, sorted) | child_curved<true>(ch.y, curved, leaf_info, sorted);
        node = parent[node];
    }
}

__global__ void k_emit_nodes(BvhBuildInput in, int n, const int2* children, const uint32_t* sorted, const Box* prim_box, const Box* node_box,
                             const uint2* leaf_info, const uint8_t* curved, BvhNode* nodes, float4* leaves) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {                                        // leaf record: the primitive's words of the stream + its ids
        const uint2 info = leaf_info[sorted[i]];
        const uint32_t kind = info.x >> 28;
        const float4* src = in.stream + (kind == 1u ? info.y : kind == 3u ? in.off_len + 2u * info.y : in.off_tri + 3u * info.y);
        const uint32_t words = kind == 1u ? 1u : kind == 3u ? 2u : 3u;
        const float4 zero = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        for (uint32_t w = 0; w < 3u; ++w) leaves[4 * (size_t)i + w] = w < words ? src[w] : zero;
        leaves[4 * (size_t)i + 3] = make_float4(__uint_as_float(info.x), __uint_as_float(info.y), 0.0f, 0.0f);
    }
    if (i >= n - 1) return;
    const int2 ch = children[i];
    const Box a = child_box(ch.x, node_box, prim_box, sorted), b = child_box(ch.y, node_box, prim_box, sorted);
    BvhNode nd;
    for (int k = 0; k < 3; ++k) { nd.lo0[k] = a.lo[k]; nd.hi0[k] = a.hi[k]; nd.lo1[k] = b.lo[k]; nd.hi1[k] = b.hi[k]; }
    nd.child0 = ch.x; nd.child1 = ch.y;
    nd.flags = (uint32_t)child_curved<false>(ch.x, curved, leaf_info, sorted) | ((uint32_t)child_curved<false>(ch.y, curved, leaf_info, sorted) << 1);
    const float ra = 0.5005f * fminf(fminf(a.hi[0] - a.lo[0], a.hi[1] - a.lo[1]), a.hi[2] - a.lo[2]) + 1e-6f;
    const float rb = 0.5005f * fminf(fminf(b.hi[0] - b.lo[0], b.hi[1] - b.lo[1]), b.hi[2] - b.lo[2]) + 1e-6f;
    nd.rho = (uint32_t)__half_as_ushort(__float2half_ru(ra)) | ((uint32_t)__half_as_ushort(__float2half_ru(rb)) << 16);   // > 65504: inf, no bound
    nodes[i] = nd;
}

struct Tmp {
    std::vector<void*> p;
    ~Tmp() { for (void* q : p) cudaFree(q); }
    template <class T> cudaError_t get(T** out, size_t n) {
        cudaError_t e = cudaMalloc((void**)out, (n ? n : 1) * sizeof(T));
        if (e == cudaSuccess) p.push_back(*out);
        return e;
    }
};
#define BV(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return e_; } while (0)

}  // namespace

void free_lbvh(BvhBuffers& b) {
    cudaFree(b.nodes); cudaFree(b.leaves);
    b.nodes = nullptr; b.leaves = nullptr; b.n_leaves = 0;
}

cudaError_t build_lbvh(const BvhBuildInput& in, BvhBuffers& out, cudaStream_t st) {
    free_lbvh(out);
    const uint32_t n = in.n_sph + in.n_len + in.n_tri;
    if (n == 0) return cudaSuccess;
    const int T = 256, G = (int)((n + T - 1) / T);
    Tmp tmp;
    Box *prim_box = nullptr, *node_box = nullptr;
    float4* centroids = nullptr;
    uint2* leaf_info = nullptr;
    Bounds3 *cb = nullptr, *scene = nullptr;
    unsigned long long *keys = nullptr, *keys_sorted = nullptr;
    uint32_t *vals = nullptr, *sorted = nullptr;
    int2* children = nullptr;
    int* parent = nullptr;
    unsigned int* visits = nullptr;
    uint8_t* curved = nullptr;
    BV(tmp.get(&prim_box, n)); BV(tmp.get(&centroids, n)); BV(tmp.get(&leaf_info, n)); BV(tmp.get(&cb, n)); BV(tmp.get(&scene, 1));
    BV(tmp.get(&keys, n)); BV(tmp.get(&keys_sorted, n)); BV(tmp.get(&vals, n)); BV(tmp.get(&sorted, n));
    BV(cudaMalloc(&out.leaves, (size_t)n * 4 * sizeof(float4)));
    BV(cudaMalloc(&out.nodes, (size_t)(n > 1 ? n - 1 : 1) * sizeof(BvhNode)));
    // out.n_leaves stays 0 until the last kernel has finished: a failed build must not look like a finished one

    k_prim_bounds<<<G, T, 0, st>>>(in, n, prim_box, centroids, leaf_info);
    k_centroid_bounds_in<<<G, T, 0, st>>>(centroids, n, cb);
    Bounds3 init;
    for (int a = 0; a < 3; ++a) { init.lo[a] = FLT_MAX; init.hi[a] = -FLT_MAX; }
    size_t bytes = 0;
    void* scratch = nullptr;
    BV(cub::DeviceReduce::Reduce(nullptr, bytes, cb, scene, (int)n, BoundsOp(), init, st));
    BV(tmp.get((char**)&scratch, bytes));
    BV(cub::DeviceReduce::Reduce(scratch, bytes, cb, scene, (int)n, BoundsOp(), init, st));
    k_morton<<<G, T, 0, st>>>(centroids, n, scene, keys, vals);
    size_t sbytes = 0;
    void* sscratch = nullptr;
    BV(cub::DeviceRadixSort::SortPairs(nullptr, sbytes, keys, keys_sorted, vals, sorted, (int)n, 0, 64, st));
    BV(tmp.get((char**)&sscratch, sbytes));
    BV(cub::DeviceRadixSort::SortPairs(sscratch, sbytes, keys, keys_sorted, vals, sorted, (int)n, 0, 64, st));
    if (n == 1) {
        k_emit_nodes<<<1, 32, 0, st>>>(in, 1, nullptr, sorted, prim_box, nullptr, leaf_info, nullptr, out.nodes, out.leaves);
        BV(cudaGetLastError());
        BV(cudaStreamSynchronize(st));
        out.n_leaves = n;
        return cudaSuccess;
    }
    BV(tmp.get(&children, n - 1)); BV(tmp.get(&parent, 2 * (size_t)n - 1)); BV(tmp.get(&visits, n - 1)); BV(tmp.get(&node_box, n - 1)); BV(tmp.get(&curved, n - 1));
    BV(cudaMemsetAsync(visits, 0, (size_t)(n - 1) * sizeof(unsigned int), st));
    k_hierarchy<<<G, T, 0, st>>>(keys_sorted, (int)n, children, parent);
    k_refit<<<G, T, 0, st>>>((int)n, children, parent, sorted, prim_box, node_box, visits, leaf_info, curved);
    k_emit_nodes<<<G, T, 0, st>>>(in, (int)n, children, sorted, prim_box, node_box, leaf_info, curved, out.nodes, out.leaves);
    BV(cudaGetLastError());
    BV(cudaStreamSynchronize(st));
    out.n_leaves = n;
    return cudaSuccess;
}

}  // namespace rtb
