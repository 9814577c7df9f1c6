// oracle/cl_shim.h — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// A minimal OpenCL-C 1.2 compatibility layer so that the reference's kernel file
// (/root/reference/kernels/raytracer.cl) can be compiled, text otherwise untouched,
// as a C++17 translation unit with g++.  The kernel text is read from the reference
// tree at build time (see oracle/Makefile) and is never copied into this repository.
//
// What the shim pins down (OpenCL leaves these implementation-defined; the choices
// here ARE the oracle's arithmetic and are mirrored by oracle/rt_oracle.c and by the
// `strict` CUDA build):
//   dot(a,b)      = a.x*b.x + a.y*b.y + a.z*b.z          (left to right, unfused)
//   cross(a,b)    = (a.y*b.z - a.z*b.y, a.z*b.x - a.x*b.z, a.x*b.y - a.y*b.x)
//   normalize(v)  = v * (1.0f / sqrtf(dot(v,v)))
//   mix(x,y,a)    = x + (y - x) * a                       (OpenCL 1.2 spec 6.12.4)
//   sign(x)       = 1 / -1 / +-0 / 0 for NaN              (spec 6.12.4)
//   pow(float,int)= powf(x, (float)n)
//   vector / s    = component-wise IEEE division
//   read_imagef   : nearest (int coords) for image2d_t; OpenCL 1.2 spec 8.2 linear
//                   filter with normalized coords for image2d_array_t, taps clamped
//                   to the edge (CLK_ADDRESS_NONE leaves out-of-range taps undefined).
#pragma once
#include <cmath>
#include <cstddef>
#include <cstdint>

typedef unsigned int uint;

struct float2 {
    float x, y;
    float2() {}
    float2(float x_, float y_) : x(x_), y(y_) {}
};

struct alignas(16) float3 {
    float x, y, z, _pad;
    float3() {}
    explicit float3(float s) : x(s), y(s), z(s), _pad(0.0f) {}
    float3(float x_, float y_, float z_) : x(x_), y(y_), z(z_), _pad(0.0f) {}
    float3& operator*=(float s) { x *= s; y *= s; z *= s; return *this; }
};

struct alignas(16) float4 {
    float x, y, z, w;
    float4() {}
    float4(float x_, float y_, float z_, float w_) : x(x_), y(y_), z(z_), w(w_) {}
    float4(const float3& v, float w_) : x(v.x), y(v.y), z(v.z), w(w_) {}
    float4(const float2& v, float z_, float w_) : x(v.x), y(v.y), z(z_), w(w_) {}
    float3 xyz() const { return float3(x, y, z); }
};

struct int2 {
    int x, y;
    int2() {}
    int2(int x_, int y_) : x(x_), y(y_) {}
};

static_assert(sizeof(float3) == 16 && alignof(float3) == 16, "cl float3 is 16 bytes");
static_assert(sizeof(float2) == 8, "cl float2 is 8 bytes");

// ---- operators -------------------------------------------------------------------
inline float3 operator+(const float3& a, const float3& b) { return float3(a.x + b.x, a.y + b.y, a.z + b.z); }
inline float3 operator-(const float3& a, const float3& b) { return float3(a.x - b.x, a.y - b.y, a.z - b.z); }
inline float3 operator-(const float3& a) { return float3(-a.x, -a.y, -a.z); }
inline float3 operator*(const float3& a, const float3& b) { return float3(a.x * b.x, a.y * b.y, a.z * b.z); }
inline float3 operator*(const float3& a, float s) { return float3(a.x * s, a.y * s, a.z * s); }
inline float3 operator*(float s, const float3& a) { return float3(s * a.x, s * a.y, s * a.z); }
inline float3 operator/(const float3& a, float s) { return float3(a.x / s, a.y / s, a.z / s); }
inline float2 operator+(const float2& a, const float2& b) { return float2(a.x + b.x, a.y + b.y); }
inline float2 operator*(const float2& a, float s) { return float2(a.x * s, a.y * s); }

// ---- builtins --------------------------------------------------------------------
inline float sqrt(float x) { return sqrtf(x); }
inline float pow(float x, int n) { return powf(x, (float)n); }
inline float dot(const float3& a, const float3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline float3 cross(const float3& a, const float3& b) {
    return float3(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
inline float3 normalize(const float3& v) {
    float inv = 1.0f / sqrtf(dot(v, v));
    return v * inv;
}
inline float3 sqrt(const float3& v) { return float3(sqrtf(v.x), sqrtf(v.y), sqrtf(v.z)); }
inline float cl_fmin(float a, float b) { return (b < a) ? b : a; }
inline float3 min(const float3& a, const float3& b) {
    return float3(cl_fmin(a.x, b.x), cl_fmin(a.y, b.y), cl_fmin(a.z, b.z));
}
inline float3 mix(const float3& x, const float3& y, float a) { return x + (y - x) * a; }
inline float sign(float x) {
    if (x != x) return 0.0f;
    if (x > 0.0f) return 1.0f;
    if (x < 0.0f) return -1.0f;
    return x;  // +-0
}

// ---- address spaces / qualifiers -------------------------------------------------
#define __kernel
#define __global
#define __constant static const
#define __read_only
#define __write_only

// ---- work-item functions ---------------------------------------------------------
extern thread_local size_t cl_shim_gid[2];
inline size_t get_global_id(uint d) { return cl_shim_gid[d]; }

// ---- images ----------------------------------------------------------------------
typedef uint sampler_t;
enum {
    CLK_NORMALIZED_COORDS_FALSE = 0, CLK_NORMALIZED_COORDS_TRUE = 1,
    CLK_ADDRESS_NONE = 0,
    CLK_FILTER_NEAREST = 0, CLK_FILTER_LINEAR = 16
};

struct image2d_t {        // RGBA float, row-major, row 0 first
    float* data; int width, height;
};
struct image2d_array_t {  // RGBA float, [layer][row][col]
    const float* data; int width, height, layers;
};

inline int get_image_width(const image2d_t& im) { return im.width; }
inline int get_image_height(const image2d_t& im) { return im.height; }

inline float4 read_imagef(const image2d_t& im, sampler_t, int2 c) {
    const float* p = im.data + 4 * ((size_t)c.y * im.width + c.x);
    return float4(p[0], p[1], p[2], p[3]);
}
inline void write_imagef(const image2d_t& im, int2 c, const float4& v) {
    float* p = im.data + 4 * ((size_t)c.y * im.width + c.x);
    p[0] = v.x; p[1] = v.y; p[2] = v.z; p[3] = v.w;
}

inline float4 cl_shim_texel(const image2d_array_t& im, int i, int j, int l) {
    if (i < 0) i = 0; if (i > im.width - 1) i = im.width - 1;
    if (j < 0) j = 0; if (j > im.height - 1) j = im.height - 1;
    const float* p = im.data + 4 * (((size_t)l * im.height + j) * im.width + i);
    return float4(p[0], p[1], p[2], p[3]);
}
// OpenCL 1.2 spec section 8.2, CLK_FILTER_LINEAR on a 2D image array, normalized coordinates.
inline float4 read_imagef(const image2d_array_t& im, sampler_t, const float4& c) {
    float u = c.x * (float)im.width, v = c.y * (float)im.height;
    float fu = floorf(u - 0.5f), fv = floorf(v - 0.5f);
    float a = (u - 0.5f) - fu, b = (v - 0.5f) - fv;
    int i0 = (int)fu, j0 = (int)fv;
    int l = (int)rintf(c.z);
    if (l < 0) l = 0; if (l > im.layers - 1) l = im.layers - 1;
    float4 t00 = cl_shim_texel(im, i0, j0, l), t10 = cl_shim_texel(im, i0 + 1, j0, l);
    float4 t01 = cl_shim_texel(im, i0, j0 + 1, l), t11 = cl_shim_texel(im, i0 + 1, j0 + 1, l);
    float w00 = (1.0f - a) * (1.0f - b), w10 = a * (1.0f - b), w01 = (1.0f - a) * b, w11 = a * b;
    return float4(w00 * t00.x + w10 * t10.x + w01 * t01.x + w11 * t11.x,
                  w00 * t00.y + w10 * t10.y + w01 * t01.y + w11 * t11.y,
                  w00 * t00.z + w10 * t10.z + w01 * t01.z + w11 * t11.z,
                  w00 * t00.w + w10 * t10.w + w01 * t01.w + w11 * t11.w);
}
