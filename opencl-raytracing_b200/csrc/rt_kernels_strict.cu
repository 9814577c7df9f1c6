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
}  // namespace rtb
