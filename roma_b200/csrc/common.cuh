// Shared helpers for the roma_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdarg.h>
#include "../../include/roma_b200.h"

#ifndef __CUDA_ARCH_LIST__
#endif

namespace roma {

// ---- error reporting -------------------------------------------------------------------
void set_error(const char* fmt, ...);
int  fail(int code, const char* fmt, ...);
int  check_launch(const char* what, int launches = 1);   // cudaGetLastError -> positive code + message; counts launches

// Per-device one-time set-up (cudaFuncSetAttribute opt-ins are per device / context, not per process).
// `if (DeviceOnce::pending(flags)) { ...set up...; DeviceOnce::done(flags); }` with `static DeviceOnce::Flags flags;`
struct DeviceOnce {
    struct Flags { bool set[64] = {}; };
    static int current() { int d = 0; return (cudaGetDevice(&d) == cudaSuccess && d >= 0 && d < 64) ? d : -1; }
    static bool pending(const Flags& f) { const int d = current(); return d < 0 || !f.set[d]; }
    static void done(Flags& f) { const int d = current(); if (d >= 0) f.set[d] = true; }
};

#define ROMA_REQUIRE(cond, code, ...) \
    do { if (!(cond)) return ::roma::fail((code), __VA_ARGS__); } while (0)

int sm_count();   // cached per process (current device)

static inline int cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---- element type traits ----------------------------------------------------------------
template <typename T> struct Elem;
template <> struct Elem<float> {
    static constexpr int kVec = 4;   // elements per 16-byte vector
    __device__ static __forceinline__ float to_f(float v) { return v; }
    __device__ static __forceinline__ float from_f(float v) { return v; }
};
template <> struct Elem<__half> {
    static constexpr int kVec = 8;
    __device__ static __forceinline__ float to_f(__half v) { return __half2float(v); }
    __device__ static __forceinline__ __half from_f(float v) { return __float2half_rn(v); }
};
template <> struct Elem<__nv_bfloat16> {
    static constexpr int kVec = 8;
    __device__ static __forceinline__ float to_f(__nv_bfloat16 v) { return __bfloat162float(v); }
    __device__ static __forceinline__ __nv_bfloat16 from_f(float v) { return __float2bfloat16_rn(v); }
};

// unpack a 16-byte vector into floats
template <typename T> __device__ __forceinline__ void unpack16(const uint4& v, float* out);
template <> __device__ __forceinline__ void unpack16<float>(const uint4& v, float* out) {
    out[0] = __uint_as_float(v.x); out[1] = __uint_as_float(v.y);
    out[2] = __uint_as_float(v.z); out[3] = __uint_as_float(v.w);
}
template <> __device__ __forceinline__ void unpack16<__half>(const uint4& v, float* out) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) { float2 f = __half22float2(h[i]); out[2 * i] = f.x; out[2 * i + 1] = f.y; }
}
template <> __device__ __forceinline__ void unpack16<__nv_bfloat16>(const uint4& v, float* out) {
    // bf16 -> fp32 is a 16-bit shift
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        out[2 * i]     = __uint_as_float(w[i] << 16);
        out[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
}
template <typename T> __device__ __forceinline__ uint4 pack16(const float* in);
template <> __device__ __forceinline__ uint4 pack16<float>(const float* in) {
    return make_uint4(__float_as_uint(in[0]), __float_as_uint(in[1]), __float_as_uint(in[2]), __float_as_uint(in[3]));
}
template <> __device__ __forceinline__ uint4 pack16<__half>(const float* in) {
    uint4 v; __half2* h = reinterpret_cast<__half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2half2_rn(in[2 * i], in[2 * i + 1]);
    return v;
}
template <> __device__ __forceinline__ uint4 pack16<__nv_bfloat16>(const float* in) {
    uint4 v; __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(in[2 * i], in[2 * i + 1]);
    return v;
}

__device__ __forceinline__ uint4 ldg16(const void* p) {
    return __ldg(reinterpret_cast<const uint4*>(p));
}

// grid_sample un-normalisation, align_corners=False (ATen grid_sampler_unnormalize)
__device__ __forceinline__ float unnormalize(float g, int size) {
    return ((g + 1.f) * (float)size - 1.f) * 0.5f;
}
// pixel-centre coordinate of index i out of n: linspace(-1+1/n, 1-1/n, n)[i]
__device__ __forceinline__ float pixel_center(int i, int n) {
    return (float)(2 * i + 1) / (float)n - 1.f;
}

// ATen upsample_bilinear2d source index (align_corners=False): src=(dst+0.5)*scale-0.5, clamped at 0
__device__ __forceinline__ void bilinear_src(int dst, float scale, int in_size, int& i0, int& i1, float& l1) {
    float s = ((float)dst + 0.5f) * scale - 0.5f;
    s = s < 0.f ? 0.f : s;
    i0 = (int)s;
    if (i0 > in_size - 1) i0 = in_size - 1;
    i1 = i0 + (i0 < in_size - 1 ? 1 : 0);
    l1 = s - (float)i0;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

}  // namespace roma
