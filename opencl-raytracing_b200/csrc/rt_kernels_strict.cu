// csrc/rt_kernels_strict.cu — the bit-exact arm: every kernel instantiated with the Exact arithmetic policy.
// (-fmad=false in the Makefile is belt and braces; the policy's __f*_rn intrinsics are never contracted.)
#include "rt_kernels.cuh"

namespace rtb {
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_trace_strict(const __grid_constant__ TraceLaunch p) {
    extern __shared__ float4 smem[];
    trace_body<Exact, false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_trace_strict_bvh(const __grid_constant__ TraceLaunch p) {
    trace_body_bvh<Exact>(p);
}
// Path heads of the frame (both arithmetic arms use these Exact kernels).
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_head(const __grid_constant__ HeadLaunch p) {
    extern __shared__ float4 smem[];
    head_body<false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_head_bvh(const __grid_constant__ HeadLaunch p) {
    head_body<true>(p, nullptr);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_trace_tail_strict(const __grid_constant__ TraceLaunch p) {
    extern __shared__ float4 smem[];
    trace_tail_body<Exact>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_trace_tail_strict_bvh(const __grid_constant__ TraceLaunch p) {
    trace_tail_bvh_body<Exact, false>(p);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_hits_strict(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    probe_hits_body<Exact, false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_hits_strict_bvh(const __grid_constant__ ProbeLaunch p) {
    probe_hits_body<Exact, true>(p, nullptr);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_bounce_strict(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    probe_bounce_body<Exact, false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_bounce_strict_bvh(const __grid_constant__ ProbeLaunch p) {
    probe_bounce_body<Exact, true>(p, nullptr);
}
cudaError_t launch_trace_strict(const TraceLaunch& p, cudaStream_t s) {
    const size_t tiles = (size_t)((p.W + 7) / 8) * ((p.H + 3) / 4);
    return p.bvh.n_leaves ? launch_staged(k_trace_strict_bvh, p, tiles * 32, s, true) : launch_staged(k_trace_strict, p, tiles * 32, s);
}
cudaError_t launch_head(const HeadLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_staged(k_head_bvh, p, p.n_slots, s, true) : launch_staged(k_head, p, p.n_slots, s);
}
cudaError_t launch_trace_tail_strict(const TraceLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_tail_bvh(k_trace_tail_strict_bvh, p, s)
                          : launch_staged(k_trace_tail_strict, p, tail_work_items(p), s, false, kHeadStateBytes);
}
cudaError_t launch_probe_hits_strict(const ProbeLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_staged(k_probe_hits_strict_bvh, p, p.n, s, true) : launch_staged(k_probe_hits_strict, p, p.n, s);
}
cudaError_t launch_probe_bounce_strict(const ProbeLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_staged(k_probe_bounce_strict_bvh, p, p.n, s, true) : launch_staged(k_probe_bounce_strict, p, p.n, s);
}

// ---- rt_trace_rays_noalloc: the nearest-hit probe for callers that may not allocate -------------------------------------
// (No counterpart in the reference.)  The probe kernel needs device scratch
// for the rays and for its results (rt_trace_rays takes it from cudaMalloc per call); this variant keeps that scratch in a
// file-scope device array of a fixed, modest size: kNoAllocMaxRays rays x (24 B ray + 4 B id + 4 B t) = 2 MiB, held by every
// GPU that loads the library until the process exits.  Every call shares the array and finds the previous call's contents in
// it; nothing has to start at zero (a call only reads the words it has just written), but two calls that overlap on the GPU
// — two streams, two host threads, two contexts on one device — race on it.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocMaxRays * 32u];

extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_hits_strict_noalloc(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    probe_hits_body<Exact, false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_hits_strict_noalloc_bvh(const __grid_constant__ ProbeLaunch p) {
    probe_hits_body<Exact, true>(p, nullptr);
}

// `p` describes the scene; rays and results are host buffers.  No cudaMalloc / cudaFree / allocating library call in here.
cudaError_t launch_probe_hits_noalloc(ProbeLaunch p, const float* h_rays6, uint32_t n, uint32_t* h_id, float* h_t, cudaStream_t s) {
    if (n == 0u) return cudaSuccess;
    if (n > kNoAllocMaxRays) return cudaErrorInvalidValue;
    // In host code `g_noalloc_scratch` is a host-side stand-in for the symbol, not a device address: the address is asked
    // for once per device, with that device current (the caller's Guard made it so).
    static unsigned char* s_base[64] = {nullptr};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!s_base[dev]) {
        void* q = nullptr;
        e = cudaGetSymbolAddress(&q, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        s_base[dev] = static_cast<unsigned char*>(q);
    }
    unsigned char* base = s_base[dev];
    float* d_rays = reinterpret_cast<float*>(base);                                              // n x 6 floats
    uint32_t* d_id = reinterpret_cast<uint32_t*>(base + (size_t)kNoAllocMaxRays * 24u);          // n ids
    float* d_t = reinterpret_cast<float*>(base + (size_t)kNoAllocMaxRays * 28u);                 // n parameters
    if ((e = cudaMemcpyAsync(d_rays, h_rays6, (size_t)n * 24u, cudaMemcpyHostToDevice, s)) != cudaSuccess) return e;
    p.rays6 = d_rays; p.n = n; p.out_id = d_id; p.out_t = d_t; p.certify_id = nullptr;
    e = p.bvh.n_leaves ? launch_staged(k_probe_hits_strict_noalloc_bvh, p, n, s, true) : launch_staged(k_probe_hits_strict_noalloc, p, n, s);
    if (e != cudaSuccess) return e;
    if (h_id && (e = cudaMemcpyAsync(h_id, d_id, (size_t)n * 4u, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return e;
    if (h_t && (e = cudaMemcpyAsync(h_t, d_t, (size_t)n * 4u, cudaMemcpyDeviceToHost, s)) != cudaSuccess) return e;
    return cudaStreamSynchronize(s);
}
}  // namespace rtb
