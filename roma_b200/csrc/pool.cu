// Encoder-side streaming kernel: 2x2 / stride-2 max pooling on channels-last feature maps.
//
// Replaces (reference): the nn.MaxPool2d(2, 2) layers of torchvision's vgg19_bn.features as RoMa walks them in
//   romatch/models/encoders.py:57-66 (VGG19.forward keeps the pre-pool activation of every stage as the feature
//   of that scale and pools it for the next stage).  The ATen NHWC kernel runs at ~25 % of HBM speed on these
//   shapes; this one moves 16 bytes per thread and load (8 bf16 / fp16 or 4 fp32 channels) and is a pure stream:
//   algorithmic bytes = B*h*w*C*sizeof(T) read + a quarter of that written.
// Floor semantics (MaxPool2d default): an odd trailing row / column is dropped.  max() of finite values is exact
// in every dtype, so results are bit-identical to ATen's.
#include "common.cuh"

namespace roma {

template <typename T> __device__ __forceinline__ uint4 max16(uint4 a, uint4 b);
template <> __device__ __forceinline__ uint4 max16<float>(uint4 a, uint4 b) {
    return make_uint4(__float_as_uint(fmaxf(__uint_as_float(a.x), __uint_as_float(b.x))),
                      __float_as_uint(fmaxf(__uint_as_float(a.y), __uint_as_float(b.y))),
                      __float_as_uint(fmaxf(__uint_as_float(a.z), __uint_as_float(b.z))),
                      __float_as_uint(fmaxf(__uint_as_float(a.w), __uint_as_float(b.w))));
}
template <> __device__ __forceinline__ uint4 max16<__half>(uint4 a, uint4 b) {
    uint4 r;
    const __half2* x = reinterpret_cast<const __half2*>(&a);
    const __half2* y = reinterpret_cast<const __half2*>(&b);
    __half2* o = reinterpret_cast<__half2*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = __hmax2(x[i], y[i]);
    return r;
}
template <> __device__ __forceinline__ uint4 max16<__nv_bfloat16>(uint4 a, uint4 b) {
    uint4 r;
    const __nv_bfloat162* x = reinterpret_cast<const __nv_bfloat162*>(&a);
    const __nv_bfloat162* y = reinterpret_cast<const __nv_bfloat162*>(&b);
    __nv_bfloat162* o = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; ++i) o[i] = __hmax2(x[i], y[i]);
    return r;
}

template <typename T>
__global__ void __launch_bounds__(256)
maxpool2_kernel(const T* __restrict__ x, T* __restrict__ out, int h, int w, int C, int ho, int wo, long long nvec_out) {
    constexpr int kVec = Elem<T>::kVec;
    const int vpp = C / kVec;                                   // 16-byte vectors per pixel
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec_out; i += stride) {
        const int v = (int)(i % vpp);
        long long t = i / vpp;
        const int xo = (int)(t % wo); t /= wo;
        const int yo = (int)(t % ho);
        const long long b = t / ho;
        const T* p = x + ((b * h + 2 * yo) * w + 2 * xo) * C + (long long)v * kVec;
        const uint4 a = ldg16(p), bq = ldg16(p + C);
        const uint4 c = ldg16(p + (long long)w * C), d = ldg16(p + (long long)w * C + C);
        *reinterpret_cast<uint4*>(out + i * kVec) = max16<T>(max16<T>(a, bq), max16<T>(c, d));
    }
}

}  // namespace roma

extern "C" int roma_maxpool2x2_nhwc(const void* x, void* out, int B, int h, int w, int C, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && out, ROMA_ERR_BAD_ARG, "roma_maxpool2x2_nhwc: null pointer");
    ROMA_REQUIRE(B > 0 && h >= 2 && w >= 2 && C > 0, ROMA_ERR_BAD_ARG, "roma_maxpool2x2_nhwc: bad sizes");
    const int kvec = dtype == ROMA_F32 ? 4 : 8;
    ROMA_REQUIRE(C % kvec == 0, ROMA_ERR_UNSUPPORTED, "roma_maxpool2x2_nhwc: C = %d must be a multiple of %d", C, kvec);
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0, ROMA_ERR_ALIGNMENT,
                 "roma_maxpool2x2_nhwc: x and out must be 16-byte aligned");
    const int ho = h / 2, wo = w / 2;
    const long long nvec = (long long)B * ho * wo * (C / kvec);
    long long blocks = (nvec + 255) / 256;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (dtype) {
        case ROMA_F32:  maxpool2_kernel<float><<<(unsigned)blocks, 256, 0, st>>>((const float*)x, (float*)out, h, w, C, ho, wo, nvec); break;
        case ROMA_F16:  maxpool2_kernel<__half><<<(unsigned)blocks, 256, 0, st>>>((const __half*)x, (__half*)out, h, w, C, ho, wo, nvec); break;
        case ROMA_BF16: maxpool2_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>((const __nv_bfloat16*)x, (__nv_bfloat16*)out, h, w, C, ho, wo, nvec); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_maxpool2x2_nhwc: unknown dtype %d", dtype);
    }
    return check_launch("roma_maxpool2x2_nhwc");
}
