// Encoder-side streaming kernel: LayerNorm over the last dimension of a (rows, D) token matrix.
//
// Replaces (reference): the nn.LayerNorm calls of the DINOv2 ViT-L/14 blocks and of the embedding decoder
//   (romatch/models/transformer/layers/block.py:88-101 norm1 / norm2, dinov2.py:236 final norm), 59 launches per
//   560->864 match.  ATen's kernel reaches ~25 % of HBM speed on (25616, 1024) bf16; here one warp owns one row,
//   keeps it in registers (NV 16-byte vectors per lane), reduces mean and the centred second moment with
//   shuffles in fp32 and writes the row once: algorithmic bytes = 2 * rows * D * sizeof(T).
// Statistics in fp32 (two-pass, as accurate as ATen's Welford), affine parameters in the storage dtype.
#include "common.cuh"

namespace roma {

template <typename T, int NV>
__global__ void __launch_bounds__(256)
layernorm_kernel(const T* __restrict__ x, const float* __restrict__ shift, const T* __restrict__ gamma, const T* __restrict__ beta,
                 T* __restrict__ out, long long rows, int D, float eps) {
    constexpr int kVec = Elem<T>::kVec;
    const int lane = threadIdx.x & 31;
    const int nvec = D / kVec;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows; r += warps) {
        const T* xr = x + r * D;
        float v[NV][kVec];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (vi < nvec) {
                unpack16<T>(ldg16(xr + vi * kVec), v[i]);
                if (shift) {                     // residual stream stored without its accumulated bias vector (see header)
#pragma unroll
                    for (int k = 0; k < kVec; ++k) v[i][k] += __ldg(shift + vi * kVec + k);
                }
#pragma unroll
                for (int k = 0; k < kVec; ++k) s += v[i][k];
            } else {
#pragma unroll
                for (int k = 0; k < kVec; ++k) v[i][k] = 0.f;
            }
        }
        const float mean = warp_sum(s) / (float)D;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            if (lane + 32 * i < nvec) {
#pragma unroll
                for (int k = 0; k < kVec; ++k) { const float d = v[i][k] - mean; q = fmaf(d, d, q); }
            }
        }
        const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
        T* orow = out + r * D;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (vi < nvec) {
                float g[kVec], b[kVec], o[kVec];
                unpack16<T>(ldg16(gamma + vi * kVec), g);
                unpack16<T>(ldg16(beta + vi * kVec), b);
#pragma unroll
                for (int k = 0; k < kVec; ++k) o[k] = fmaf((v[i][k] - mean) * rstd, g[k], b[k]);
                *reinterpret_cast<uint4*>(orow + vi * kVec) = pack16<T>(o);
            }
        }
    }
}

template <typename T>
static int launch_layernorm(const void* x, const float* shift, const void* g, const void* b, void* out, long long rows, int D, float eps, cudaStream_t st) {
    constexpr int kVec = Elem<T>::kVec;
    const int nv = cdiv(D / kVec, 32);
    long long blocks = (rows + 7) / 8;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
#define ROMA_LN(NVC) layernorm_kernel<T, NVC><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, shift, (const T*)g, (const T*)b, (T*)out, rows, D, eps)
    if (nv <= 1) ROMA_LN(1);
    else if (nv <= 2) ROMA_LN(2);
    else if (nv <= 4) ROMA_LN(4);
    else if (nv <= 8) ROMA_LN(8);
    else return fail(ROMA_ERR_UNSUPPORTED, "roma_layernorm: D = %d too wide (max %d)", D, 8 * 32 * kVec);
#undef ROMA_LN
    return 0;
}

}  // namespace roma

static int layernorm_entry(const char* who, const void* x, const float* shift, const void* gamma, const void* beta, void* out,
                           long long rows, int D, float eps, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && gamma && beta && out, ROMA_ERR_BAD_ARG, "%s: null pointer", who);
    ROMA_REQUIRE(rows > 0 && D > 0, ROMA_ERR_BAD_ARG, "%s: bad sizes", who);
    const int kvec = dtype == ROMA_F32 ? 4 : 8;
    ROMA_REQUIRE(D % kvec == 0, ROMA_ERR_UNSUPPORTED, "%s: D = %d must be a multiple of %d", who, D, kvec);
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(gamma) |
                   reinterpret_cast<uintptr_t>(beta)) & 15u) == 0, ROMA_ERR_ALIGNMENT, "%s: pointers must be 16-byte aligned", who);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int rc;
    switch (dtype) {
        case ROMA_F32:  rc = launch_layernorm<float>(x, shift, gamma, beta, out, rows, D, eps, st); break;
        case ROMA_F16:  rc = launch_layernorm<__half>(x, shift, gamma, beta, out, rows, D, eps, st); break;
        case ROMA_BF16: rc = launch_layernorm<__nv_bfloat16>(x, shift, gamma, beta, out, rows, D, eps, st); break;
        default: return fail(ROMA_ERR_BAD_ARG, "%s: unknown dtype %d", who, dtype);
    }
    if (rc) return rc;
    return check_launch(who);
}

extern "C" int roma_layernorm(const void* x, const void* gamma, const void* beta, void* out,
                              long long rows, int D, float eps, int dtype, void* stream) {
    return layernorm_entry("roma_layernorm", x, nullptr, gamma, beta, out, rows, D, eps, dtype, stream);
}

// LayerNorm(x + shift), shift fp32 (D): the residual stream of a pre-norm transformer can be kept WITHOUT the bias
// vectors of its residual branches (x' = x - sum of the biases so far), so that every branch is one GEMM accumulating
// in place (beta = 1, no separate residual add and no bias pass); the accumulated bias re-enters here, in fp32, in
// front of the statistics.
extern "C" int roma_layernorm_shift(const void* x, const float* shift, const void* gamma, const void* beta, void* out,
                                    long long rows, int D, float eps, int dtype, void* stream) {
    ROMA_REQUIRE(shift, ROMA_ERR_BAD_ARG, "roma_layernorm_shift: null shift");
    return layernorm_entry("roma_layernorm_shift", x, shift, gamma, beta, out, rows, D, eps, dtype, stream);
}
