// csrc/rt_kernels_strict.cu — the bit-exact arm: every kernel instantiated with the Exact arithmetic policy.
// (-fmad=false in the Makefile is belt and braces; the policy's __f*_rn intrinsics are never contracted.)
#include "rt_kernels.cuh"

namespace rtb {
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_trace_strict(const __grid_constant__ TraceLaunch p) {
    extern __shared__ float4 smem[];
    trace_body<Exact>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_hits_strict(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    probe_hits_body<Exact>(p, smem);
}
extern "C" __global__ void __launch_bounds__(kTraceThreads) k_probe_bounce_strict(const __grid_constant__ ProbeLaunch p) {
    extern __shared__ float4 smem[];
    probe_bounce_body<Exact>(p, smem);
}
cudaError_t launch_trace_strict(const TraceLaunch& p, cudaStream_t s) {
    const size_t tiles = (size_t)((p.W + 7) / 8) * ((p.H + 3) / 4);
    return launch_staged(k_trace_strict, p, tiles * 32, s);
}
cudaError_t launch_probe_hits_strict(const ProbeLaunch& p, cudaStream_t s) { return launch_staged(k_probe_hits_strict, p, p.n, s); }
cudaError_t launch_probe_bounce_strict(const ProbeLaunch& p, cudaStream_t s) { return launch_staged(k_probe_bounce_strict, p, p.n, s); }
}  // namespace rtb
