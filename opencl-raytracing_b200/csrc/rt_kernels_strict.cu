// csrc/rt_kernels_strict.cu — bit-exact arm.  Build flags (Makefile): -fmad=false -prec-div=true -prec-sqrt=true -ftz=false
#define RT_STRICT 1
#define RT_FLAVOUR strict
#define RT_KERNEL_NAME(n) n##_strict
#include "rt_kernels.cuh"

namespace rtb {
cudaError_t launch_trace_strict(const TraceLaunch& p, cudaStream_t s) {
    const size_t tiles = (size_t)((p.W + 7) / 8) * ((p.H + 3) / 4);
    return strict::launch_staged(strict::k_trace_strict, p, tiles * 32, s);
}
cudaError_t launch_probe_hits_strict(const ProbeLaunch& p, cudaStream_t s) {
    return strict::launch_staged(strict::k_probe_hits_strict, p, p.n, s);
}
cudaError_t launch_probe_bounce_strict(const ProbeLaunch& p, cudaStream_t s) {
    return strict::launch_staged(strict::k_probe_bounce_strict, p, p.n, s);
}
}  // namespace rtb
