// Encoder-side streaming kernel: exact (erf) GELU over a flat tensor.
//
// Replaces (reference): nn.GELU() between fc1 and fc2 of every transformer MLP
//   (romatch/models/transformer/layers/mlp.py:23, 31, 37; 24 DINOv2 blocks on (B * 1601, 4096) + 5 embedding-decoder
//   blocks per 560 -> 864 match).  ATen evaluates 0.5 * x * (1 + erf(x / sqrt 2)) in fp32 and runs at ~2.9 TB/s on
//   (25616, 4096) bf16 (146 us); the op is a pure stream: algorithmic bytes = 2 * n * sizeof(T).
// Arithmetic: GELU(x) = x * Phi(x), Phi(-|x|) = 0.5 * erfc(|x| / sqrt 2) from the Chebyshev-fitted form
//   erfc(z) = t * exp(-z^2 + P(t)), t = 1 / (1 + z / 2)  (fractional error < 1.2e-7 for all z >= 0), evaluated with one
//   MUFU.RCP, one MUFU.EX2 and ten FMAs (packed FFMA2 over element pairs); x >= 0 uses x - x * Phi(-x).  Unlike 0.5 * x * (1 + erf) this keeps full RELATIVE
//   accuracy in the negative tail (no cancellation), so 16-bit outputs are the correctly rounded exact GELU; they differ
//   from ATen's by one unit in the last place in ~2 % (bf16) / ~3 % (fp16) of the elements, all with |y| < 0.05.
#include "common.cuh"

namespace roma {

__device__ __forceinline__ float rcp_approx(float v) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }
__device__ __forceinline__ float ex2_approx(float v) { float r; asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(v)); return r; }

// two elements at a time: the polynomial and the products run as packed FFMA2 / FMUL2 (one issue slot for two lanes'
// worth of work - the kernel is issue-bound, not FMA-pipe-bound), the two MUFU ops and the final select per element
__device__ __forceinline__ float2 gelu_erf2(float2 x) {
    constexpr float L = 1.4426950408889634f;                          // log2(e): exp(u) = ex2(L * u)
    const float2 z = __fmul2_rn(make_float2(fabsf(x.x), fabsf(x.y)), make_float2(0.70710678118654752f, 0.70710678118654752f));
    const float2 den = __ffma2_rn(make_float2(0.5f, 0.5f), z, make_float2(1.0f, 1.0f));
    const float2 t = make_float2(rcp_approx(den.x), rcp_approx(den.y));
#define ROMA_C2(v) make_float2((v), (v))
    float2 p = ROMA_C2(0.17087277f * L);
    p = __ffma2_rn(p, t, ROMA_C2(-0.82215223f * L));
    p = __ffma2_rn(p, t, ROMA_C2(1.48851587f * L));
    p = __ffma2_rn(p, t, ROMA_C2(-1.13520398f * L));
    p = __ffma2_rn(p, t, ROMA_C2(0.27886807f * L));
    p = __ffma2_rn(p, t, ROMA_C2(-0.18628806f * L));
    p = __ffma2_rn(p, t, ROMA_C2(0.09678418f * L));
    p = __ffma2_rn(p, t, ROMA_C2(0.37409196f * L));
    p = __ffma2_rn(p, t, ROMA_C2(1.00002368f * L));
    p = __ffma2_rn(p, t, ROMA_C2(-1.26551223f * L - 1.0f));           // - 1: the factor 0.5 of Phi
    const float2 u = __ffma2_rn(__fmul2_rn(z, ROMA_C2(-L)), z, p);
#undef ROMA_C2
    const float2 e = make_float2(ex2_approx(u.x), ex2_approx(u.y));
    const float2 xq = __fmul2_rn(x, __fmul2_rn(t, e));                // x * Phi(-|x|)
    return make_float2(x.x >= 0.f ? x.x - xq.x : xq.x, x.y >= 0.f ? x.y - xq.y : xq.y);
}

template <typename T>
__global__ void __launch_bounds__(256)
gelu_kernel(const T* __restrict__ x, T* __restrict__ out, long long nvec) {
    constexpr int kVec = Elem<T>::kVec;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += stride) {
        float v[kVec];
        unpack16<T>(ldg16(x + i * kVec), v);
#pragma unroll
        for (int k = 0; k < kVec; k += 2) {
            const float2 r = gelu_erf2(make_float2(v[k], v[k + 1]));
            v[k] = r.x; v[k + 1] = r.y;
        }
        *reinterpret_cast<uint4*>(out + i * kVec) = pack16<T>(v);
    }
}

template <typename T>
static void launch_gelu(const void* x, void* out, long long n, cudaStream_t st) {
    const long long nvec = n / Elem<T>::kVec;
    long long blocks = (nvec + 255) / 256;
    const long long cap = (long long)sm_count() * 32;
    if (blocks > cap) blocks = cap;
    gelu_kernel<T><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, (T*)out, nvec);
}

}  // namespace roma

extern "C" int roma_gelu(const void* x, void* out, long long n, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && out, ROMA_ERR_BAD_ARG, "roma_gelu: null pointer");
    ROMA_REQUIRE(n > 0, ROMA_ERR_BAD_ARG, "roma_gelu: bad size");
    const int kvec = dtype == ROMA_F32 ? 4 : 8;
    ROMA_REQUIRE(n % kvec == 0, ROMA_ERR_UNSUPPORTED, "roma_gelu: n = %lld must be a multiple of %d", n, kvec);
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0, ROMA_ERR_ALIGNMENT,
                 "roma_gelu: pointers must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (dtype) {
        case ROMA_F32:  launch_gelu<float>(x, out, n, st); break;
        case ROMA_F16:  launch_gelu<__half>(x, out, n, st); break;
        case ROMA_BF16: launch_gelu<__nv_bfloat16>(x, out, n, st); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_gelu: unknown dtype %d", dtype);
    }
    return check_launch("roma_gelu");
}
