// (f-2) ConvRefiner body: depthwise 5x5 conv + folded BatchNorm + ReLU in one pass, channels-last.
//
// Replaces (reference): the conv1 -> norm -> relu triple of every ConvRefiner block
//   romatch/models/matcher.py:77-103 (create_block: depthwise kernel_size=5, padding=2, BatchNorm2d, ReLU)
// as it runs at inference (BatchNorm in eval mode is an affine map; the host folds it into the weights/bias).
// In the reference this is four ATen launches per block (cudnn depthwise conv, bias add_, batch_norm, relu_),
// 81 blocks per 560->864 match; here it is one launch per block and one read + one write of the activation.
//
// Layout: x, out (B, h, w, Cp) channels-last, Cp even; weights fp32 (25, Cp) tap-major so that one tap of
// consecutive channels is contiguous; bias fp32 (Cp).  fp32 accumulation.
// Mapping: a thread owns one channel PAIR (one 32-bit load for 16-bit types), a run of 4 output columns and a
// strip of RY output rows.  It keeps the 5 x 8 input window (packed) in registers and slides it down the strip:
// per output row it loads 8 new values and does 4 x 25 x 2 FMAs — the kernel is FP32-FMA bound (25 MACs per
// output element, ~0.44 TMAC per 8-pair step), not HBM bound; lanes run over channel pairs, so a warp reads
// 128 contiguous bytes per tap.
#include "common.cuh"

namespace roma {

template <typename T> struct Pair;
template <> struct Pair<float> {
    using Raw = float2;
    __device__ static __forceinline__ Raw zero() { return make_float2(0.f, 0.f); }
    __device__ static __forceinline__ Raw load(const float* p) { return *reinterpret_cast<const float2*>(p); }
    __device__ static __forceinline__ float2 unpack(Raw r) { return r; }
    __device__ static __forceinline__ void store(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
};
template <> struct Pair<__half> {
    using Raw = uint32_t;
    __device__ static __forceinline__ Raw zero() { return 0u; }
    __device__ static __forceinline__ Raw load(const __half* p) { return *reinterpret_cast<const uint32_t*>(p); }
    __device__ static __forceinline__ float2 unpack(Raw r) { return __half22float2(*reinterpret_cast<const __half2*>(&r)); }
    __device__ static __forceinline__ void store(__half* p, float a, float b) { *reinterpret_cast<__half2*>(p) = __floats2half2_rn(a, b); }
};
template <> struct Pair<__nv_bfloat16> {
    using Raw = uint32_t;
    __device__ static __forceinline__ Raw zero() { return 0u; }
    __device__ static __forceinline__ Raw load(const __nv_bfloat16* p) { return *reinterpret_cast<const uint32_t*>(p); }
    __device__ static __forceinline__ float2 unpack(Raw r) { return make_float2(__uint_as_float(r << 16), __uint_as_float(r & 0xffff0000u)); }
    __device__ static __forceinline__ void store(__nv_bfloat16* p, float a, float b) { *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b); }
};

constexpr int kDwRY = 8;      // output rows per thread strip
constexpr int kDwTW = 4;      // output columns per thread

template <typename T>
__global__ void __launch_bounds__(128, 4)
dwconv5_kernel(const T* __restrict__ x, const float* __restrict__ wgt, const float* __restrict__ bias, T* __restrict__ out,
               int B, int h, int w, int Cp, int relu, long long nthreads) {
    using P = Pair<T>;
    const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (gid >= nthreads) return;
    const int npairs = Cp >> 1, runs = (w + kDwTW - 1) / kDwTW, strips = (h + kDwRY - 1) / kDwRY;
    const int pair = (int)(gid % npairs);
    long long t = gid / npairs;
    const int run = (int)(t % runs); t /= runs;
    const int strip = (int)(t % strips);
    const int b = (int)(t / strips);
    const int c = pair * 2, x0 = run * kDwTW, y0 = strip * kDwRY;

    float2 wv[25];
#pragma unroll
    for (int k = 0; k < 25; ++k) wv[k] = __ldg(reinterpret_cast<const float2*>(wgt + (long long)k * Cp + c));
    const float2 bv = __ldg(reinterpret_cast<const float2*>(bias + c));

    const T* xb = x + (long long)b * h * w * Cp + c;
    T* ob = out + (long long)b * h * w * Cp + c;
    auto load_row = [&](int y, typename P::Raw* dst) {
        const bool yok = y >= 0 && y < h;
#pragma unroll
        for (int j = 0; j < kDwTW + 4; ++j) {
            const int xx = x0 - 2 + j;
            dst[j] = (yok && xx >= 0 && xx < w) ? P::load(xb + ((long long)y * w + xx) * Cp) : P::zero();
        }
    };
    typename P::Raw win[5][kDwTW + 4];
#pragma unroll
    for (int i = 0; i < 4; ++i) load_row(y0 - 2 + i, win[i]);

#pragma unroll 1
    for (int yy = 0; yy < kDwRY; ++yy) {
        const int y = y0 + yy;
        if (y >= h) break;
        load_row(y + 2, win[4]);
        float2 acc[kDwTW];
#pragma unroll
        for (int ox = 0; ox < kDwTW; ++ox) acc[ox] = bv;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
#pragma unroll
            for (int j = 0; j < kDwTW + 4; ++j) {
                const float2 v = P::unpack(win[i][j]);
#pragma unroll
                for (int ox = 0; ox < kDwTW; ++ox) {
                    const int dx = j - ox;
                    if (dx >= 0 && dx < 5) {
                        acc[ox].x = fmaf(v.x, wv[i * 5 + dx].x, acc[ox].x);
                        acc[ox].y = fmaf(v.y, wv[i * 5 + dx].y, acc[ox].y);
                    }
                }
            }
        }
#pragma unroll
        for (int ox = 0; ox < kDwTW; ++ox) {
            if (x0 + ox < w) {
                float a = acc[ox].x, bb = acc[ox].y;
                if (relu) { a = fmaxf(a, 0.f); bb = fmaxf(bb, 0.f); }
                P::store(ob + ((long long)y * w + x0 + ox) * Cp, a, bb);
            }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < kDwTW + 4; ++j) win[i][j] = win[i + 1][j];
    }
}

}  // namespace roma

extern "C" int roma_dwconv5x5_bias_act(const void* x, const float* weight, const float* bias, void* out,
                                       int B, int h, int w, int C, int relu, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && weight && bias && out, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: null pointer");
    ROMA_REQUIRE(B > 0 && h > 0 && w > 0 && C > 0, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: bad sizes");
    ROMA_REQUIRE(C % 2 == 0, ROMA_ERR_UNSUPPORTED, "roma_dwconv5x5_bias_act: channel count %d must be even (pad it)", C);
    ROMA_REQUIRE(x != out, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: in-place operation is not supported");
    const uintptr_t al = dtype == ROMA_F32 ? 7u : 3u;
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & al) == 0 &&
                 ((reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(bias)) & 7u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_dwconv5x5_bias_act: misaligned pointer");
    const long long n = (long long)B * ((h + kDwRY - 1) / kDwRY) * ((w + kDwTW - 1) / kDwTW) * (C / 2);
    const int grid = cdiv(n, 128);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (dtype) {
        case ROMA_F32:  dwconv5_kernel<float><<<grid, 128, 0, st>>>((const float*)x, weight, bias, (float*)out, B, h, w, C, relu, n); break;
        case ROMA_F16:  dwconv5_kernel<__half><<<grid, 128, 0, st>>>((const __half*)x, weight, bias, (__half*)out, B, h, w, C, relu, n); break;
        case ROMA_BF16: dwconv5_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>((const __nv_bfloat16*)x, weight, bias, (__nv_bfloat16*)out, B, h, w, C, relu, n); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: unknown dtype %d", dtype);
    }
    return check_launch("roma_dwconv5x5_bias_act");
}
