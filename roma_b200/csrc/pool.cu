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

// I: index type — unsigned int whenever every element offset of x fits 32 bits (all encoder shapes): the three div / mod pairs of
// the output decode are 64-bit software routines otherwise (~300 instructions per 16-byte vector against 4 loads and 1 store).
template <typename T, typename I>
__global__ void __launch_bounds__(256)
maxpool2_kernel(const T* __restrict__ x, T* __restrict__ out, int h, int w, int C, int ho, int wo, long long nvec_out) {
    constexpr int kVec = Elem<T>::kVec;
    const I vpp = (I)(C / kVec);                                // 16-byte vectors per pixel
    const I stride = (I)gridDim.x * blockDim.x, n = (I)nvec_out;
    for (I i = (I)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        const I v = i % vpp;
        I t = i / vpp;
        const I xo = t % (I)wo; t /= (I)wo;
        const I yo = t % (I)ho;
        const I b = t / (I)ho;
        const T* p = x + ((b * (I)h + 2 * yo) * (I)w + 2 * xo) * (I)C + v * kVec;
        const uint4 a = ldg16(p), bq = ldg16(p + C);
        const uint4 c = ldg16(p + (I)w * (I)C), d = ldg16(p + (I)w * (I)C + C);
        *reinterpret_cast<uint4*>(out + i * kVec) = max16<T>(max16<T>(a, bq), max16<T>(c, d));
    }
}

template <typename T>
static void launch_maxpool2(const void* x, void* out, int B, int h, int w, int C, int ho, int wo, long long nvec, unsigned blocks, cudaStream_t st) {
    const bool small = (long long)B * h * w * C < (1ll << 32) && nvec + (long long)blocks * 256 < (1ll << 32);
    if (small) maxpool2_kernel<T, unsigned int><<<blocks, 256, 0, st>>>((const T*)x, (T*)out, h, w, C, ho, wo, nvec);
    else maxpool2_kernel<T, long long><<<blocks, 256, 0, st>>>((const T*)x, (T*)out, h, w, C, ho, wo, nvec);
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
        case ROMA_F32:  launch_maxpool2<float>(x, out, B, h, w, C, ho, wo, nvec, (unsigned)blocks, st); break;
        case ROMA_F16:  launch_maxpool2<__half>(x, out, B, h, w, C, ho, wo, nvec, (unsigned)blocks, st); break;
        case ROMA_BF16: launch_maxpool2<__nv_bfloat16>(x, out, B, h, w, C, ho, wo, nvec, (unsigned)blocks, st); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_maxpool2x2_nhwc: unknown dtype %d", dtype);
    }
    return check_launch("roma_maxpool2x2_nhwc");
}
