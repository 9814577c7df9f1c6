// csrc/rt_kernels_fast.cu — the throughput arm: kernels instantiated with the Fused arithmetic policy
// (FMA contraction, MUFU rcp/sqrt/rsqrt), primary hits certified against the Exact policy.
#include "rt_kernels.cuh"
#ifndef RT_MIN_BLOCKS
#define RT_MIN_BLOCKS 7   // 71 registers, no spills, 28 warps per SM: 69.5 vs 68.6 Grays/s at 8 blocks (64 registers, 24 B of spills) on C2
#endif

namespace rtb {
extern "C" __global__ void __launch_bounds__(kTraceThreads, RT_MIN_BLOCKS) k_trace_fast(const __grid_constant__ TraceLaunch p) {
    extern __shared__ float4 smem[];
    trace_body<Fused, false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_trace_fast_bvh(const __grid_constant__ TraceLaunch p) {
    trace_body_bvh<Fused>(p);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads, RT_MIN_BLOCKS) k_trace_tail_fast(const __grid_constant__ TraceLaunch p) {
    extern __shared__ float4 smem[];
    trace_tail_body<Fused>(p, smem);
}
#ifndef RT_BVH_MIN_BLOCKS
#define RT_BVH_MIN_BLOCKS 7
#endif
extern "C" __global__ void __launch_bounds__(kTraceThreads, RT_BVH_MIN_BLOCKS) k_trace_tail_fast_bvh(const __grid_constant__ TraceLaunch p) {
    trace_tail_bvh_body<Fused, false>(p);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads, RT_BVH_MIN_BLOCKS) k_trace_tail_fast_bvh_count(const __grid_constant__ TraceLaunch p) {
    trace_tail_bvh_body<Fused, true>(p);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_hits_fast(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    probe_hits_body<Fused, false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_hits_fast_bvh(const __grid_constant__ ProbeLaunch p) {
    probe_hits_body<Fused, true>(p, nullptr);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_bounce_fast(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    probe_bounce_body<Fused, false>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_bounce_fast_bvh(const __grid_constant__ ProbeLaunch p) {
    probe_bounce_body<Fused, true>(p, nullptr);
}
cudaError_t launch_trace_fast(const TraceLaunch& p, cudaStream_t s) {
    const size_t tiles = (size_t)((p.W + 7) / 8) * ((p.H + 3) / 4);
    return p.bvh.n_leaves ? launch_staged(k_trace_fast_bvh, p, tiles * 32, s, true) : launch_staged(k_trace_fast, p, tiles * 32, s);
}
cudaError_t launch_trace_tail_fast(const TraceLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_tail_bvh(k_trace_tail_fast_bvh, p, s)
                          : launch_staged(k_trace_tail_fast, p, tail_work_items(p), s, false, kHeadStateBytes);
}
cudaError_t launch_trace_tail_fast_count(const TraceLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_tail_bvh(k_trace_tail_fast_bvh_count, p, s) : cudaErrorInvalidValue;
}
cudaError_t launch_probe_hits_fast(const ProbeLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_staged(k_probe_hits_fast_bvh, p, p.n, s, true) : launch_staged(k_probe_hits_fast, p, p.n, s);
}
cudaError_t launch_probe_bounce_fast(const ProbeLaunch& p, cudaStream_t s) {
    return p.bvh.n_leaves ? launch_staged(k_probe_bounce_fast_bvh, p, p.n, s, true) : launch_staged(k_probe_bounce_fast, p, p.n, s);
}
}  // namespace rtb
