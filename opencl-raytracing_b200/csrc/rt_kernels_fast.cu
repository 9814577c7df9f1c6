// csrc/rt_kernels_fast.cu — throughput arm.  Build flags (Makefile): -fmad=true -prec-div=false -prec-sqrt=false
#define RT_STRICT 0
#define RT_FLAVOUR fast
#define RT_KERNEL_NAME(n) n##_fast
#include "rt_kernels.cuh"

namespace rtb {
cudaError_t launch_trace_fast(const TraceLaunch& p, cudaStream_t s) {
    const size_t tiles = (size_t)((p.W + 7) / 8) * ((p.H + 3) / 4);
    return fast::launch_staged(fast::k_trace_fast, p, tiles * 32, s);
}
cudaError_t launch_probe_hits_fast(const ProbeLaunch& p, cudaStream_t s) {
    return fast::launch_staged(fast::k_probe_hits_fast, p, p.n, s);
}
cudaError_t launch_probe_bounce_fast(const ProbeLaunch& p, cudaStream_t s) {
    return fast::launch_staged(fast::k_probe_bounce_fast, p, p.n, s);
}
}  // namespace rtb
