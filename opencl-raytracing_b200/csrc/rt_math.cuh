// csrc/rt_math.cuh — arithmetic policies and small vector helpers for the path-tracing kernels.
//
//   Exact : every operation is ONE correctly rounded IEEE binary32 operation in the order written
//           (__fmul_rn/__fadd_rn/... are never contracted into FMAs, whatever the compiler flags), division
//           and square root are the IEEE ones.  This is the arithmetic of the oracle (oracle/rt_oracle.c,
//           oracle/cl_shim.h), so kernels instantiated with it are bit-exact against the reference.
//   Fused : plain operators (the compiler contracts a*b+c into FFMA), MUFU approximations for
//           reciprocal / square root / reciprocal square root.  The throughput arithmetic.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace rtb {

struct V3 { float x, y, z; };
__device__ __forceinline__ V3 mk(float x, float y, float z) { V3 r; r.x = x; r.y = y; r.z = z; return r; }
__device__ __forceinline__ V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ float min_cl(float a, float b) { return (b < a) ? b : a; }   // OpenCL min(x, y)
__device__ __forceinline__ V3 vmin(V3 a, V3 b) { return mk(min_cl(a.x, b.x), min_cl(a.y, b.y), min_cl(a.z, b.z)); }

struct Exact {
    static constexpr bool kExact = true;
    static __device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
    static __device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
    static __device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
    static __device__ __forceinline__ float div(float a, float b) { return __fdiv_rn(a, b); }
    static __device__ __forceinline__ float rcp(float a) { return __fdiv_rn(1.0f, a); }
    static __device__ __forceinline__ float sqrt(float a) { return __fsqrt_rn(a); }
    // a*b + c and a*b - c as two rounded operations
    static __device__ __forceinline__ float mad(float a, float b, float c) { return __fadd_rn(__fmul_rn(a, b), c); }
    static __device__ __forceinline__ float msub(float a, float b, float c) { return __fsub_rn(__fmul_rn(a, b), c); }
    static __device__ __forceinline__ V3 normalize(V3 v);
    static __device__ __forceinline__ V3 vdiv(V3 v, float s) { return mk(__fdiv_rn(v.x, s), __fdiv_rn(v.y, s), __fdiv_rn(v.z, s)); }
    // glibc powf (the oracle) and (float)pow(double) are both correctly rounded in all but ~1e-8 of cases
    static __device__ __forceinline__ float pow5(float x) { return (float)pow((double)x, 5.0); }
};

struct Fused {
    static constexpr bool kExact = false;
    static __device__ __forceinline__ float mul(float a, float b) { return a * b; }
    static __device__ __forceinline__ float add(float a, float b) { return a + b; }
    static __device__ __forceinline__ float sub(float a, float b) { return a - b; }
    static __device__ __forceinline__ float div(float a, float b) { return __fdividef(a, b); }
    static __device__ __forceinline__ float rcp(float a) { return __fdividef(1.0f, a); }
    static __device__ __forceinline__ float sqrt(float a) {
        float r;
        asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
        return r;
    }
    static __device__ __forceinline__ float mad(float a, float b, float c) { return fmaf(a, b, c); }
    static __device__ __forceinline__ float msub(float a, float b, float c) { return fmaf(a, b, -c); }
    static __device__ __forceinline__ V3 normalize(V3 v);
    static __device__ __forceinline__ V3 vdiv(V3 v, float s) { const float r = __fdividef(1.0f, s); return mk(v.x * r, v.y * r, v.z * r); }
    static __device__ __forceinline__ float pow5(float x) { const float x2 = x * x; return x2 * x2 * x; }
};

// Vector helpers in the oracle's operation order: dot = (x*x' + y*y') + z*z', cross as in cl_shim.h.
template <class A> __device__ __forceinline__ V3 vadd(V3 a, V3 b) { return mk(A::add(a.x, b.x), A::add(a.y, b.y), A::add(a.z, b.z)); }
template <class A> __device__ __forceinline__ V3 vsub(V3 a, V3 b) { return mk(A::sub(a.x, b.x), A::sub(a.y, b.y), A::sub(a.z, b.z)); }
template <class A> __device__ __forceinline__ V3 vscale(V3 a, float s) { return mk(A::mul(a.x, s), A::mul(a.y, s), A::mul(a.z, s)); }
template <class A> __device__ __forceinline__ float dot(V3 a, V3 b) {
    return A::mad(a.z, b.z, A::mad(a.y, b.y, A::mul(a.x, b.x)));
}
template <class A> __device__ __forceinline__ V3 cross(V3 a, V3 b) {
    return mk(A::msub(a.y, b.z, A::mul(a.z, b.y)), A::msub(a.z, b.x, A::mul(a.x, b.z)), A::msub(a.x, b.y, A::mul(a.y, b.x)));
}
// a + b*s (component-wise), used for o + d*t
template <class A> __device__ __forceinline__ V3 vmad(V3 b, float s, V3 a) {
    return mk(A::mad(b.x, s, a.x), A::mad(b.y, s, a.y), A::mad(b.z, s, a.z));
}

__device__ __forceinline__ V3 Exact::normalize(V3 v) {
    const float inv = __fdiv_rn(1.0f, __fsqrt_rn(dot<Exact>(v, v)));
    return vscale<Exact>(v, inv);
}
__device__ __forceinline__ V3 Fused::normalize(V3 v) { return vscale<Fused>(v, rsqrtf(dot<Fused>(v, v))); }

}  // namespace rtb
