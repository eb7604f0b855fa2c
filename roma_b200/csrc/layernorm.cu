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

// kVec floats -> kVec elements of TO, one store (16 bytes when the element sizes match, 8 when fp32 -> 16-bit)
template <typename TO, int N> __device__ __forceinline__ void store_row_vec(TO* dst, const float* o) {
    if constexpr (sizeof(TO) * N == 16) {
        *reinterpret_cast<uint4*>(dst) = pack16<TO>(o);
    } else {
        static_assert(sizeof(TO) == 2 && N == 4, "fp32 rows to a 16-bit output");
        TO t[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) t[k] = Elem<TO>::from_f(o[k]);
        *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(t);
    }
}

// TO = T, or a 16-bit type for fp32 rows: the cast autocast would insert in front of the Linear that consumes the
// normalised tokens (fp32 LayerNorm output rounded once to the 16-bit dtype — the same bits, one pass less).
// FULL: D == 32 * NV * kVec — every lane owns NV whole vectors, so the row's loads carry no predicate and are all issued
// before the first use (with the `vi < nvec` tests each vector was its own basic block: load, wait, use, next load).
template <typename T, int NV, typename TO = T, bool FULL = false>
__global__ void __launch_bounds__(256)
layernorm_kernel(const T* __restrict__ x, const float* __restrict__ shift, const T* __restrict__ gamma, const T* __restrict__ beta,
                 TO* __restrict__ out, long long rows, int D, float eps) {
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
            if (FULL || vi < nvec) {
                unpack16<T>(ldg16(xr + vi * kVec), v[i]);
            } else {
#pragma unroll
                for (int k = 0; k < kVec; ++k) v[i][k] = 0.f;
            }
        }
        if (shift) {                             // residual stream stored without its accumulated bias vector (see header)
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int vi = lane + 32 * i;
                if (FULL || vi < nvec) {
#pragma unroll
                    for (int k = 0; k < kVec; ++k) v[i][k] += __ldg(shift + vi * kVec + k);
                }
            }
        }
#pragma unroll
        for (int i = 0; i < NV; ++i) {
#pragma unroll
            for (int k = 0; k < kVec; ++k) s += v[i][k];     // lanes beyond the row hold zeros
        }
        const float mean = warp_sum(s) / (float)D;
        float q = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            if (FULL || lane + 32 * i < nvec) {
#pragma unroll
                for (int k = 0; k < kVec; ++k) { const float d = v[i][k] - mean; q = fmaf(d, d, q); }
            }
        }
        const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
        TO* orow = out + r * D;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (FULL || vi < nvec) {
                float g[kVec], b[kVec], o[kVec];
                unpack16<T>(ldg16(gamma + vi * kVec), g);
                unpack16<T>(ldg16(beta + vi * kVec), b);
#pragma unroll
                for (int k = 0; k < kVec; ++k) o[k] = fmaf((v[i][k] - mean) * rstd, g[k], b[k]);
                store_row_vec<TO, kVec>(orow + vi * kVec, o);
            }
        }
    }
}

// s = x + y (rounded to T, as ATen's add of two T tensors) written to `sum` (may alias x) and LayerNorm(s) written to
// `out`: the residual add of a pre-norm transformer block together with the norm that follows it — one read of x and
// y, one write of each result, instead of add (2 reads, 1 write) + LayerNorm (1 read, 1 write).  Same bits as the pair.
template <typename T, int NV, bool FULL = false>
__global__ void __launch_bounds__(256)
add_layernorm_kernel(const T* __restrict__ x, const T* __restrict__ y, const T* __restrict__ gamma, const T* __restrict__ beta,
                     T* __restrict__ sum, T* __restrict__ out, long long rows, int D, float eps) {
    constexpr int kVec = Elem<T>::kVec;
    const int lane = threadIdx.x & 31;
    const int nvec = D / kVec;
    const long long warps = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long r = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; r < rows; r += warps) {
        const T* xr = x + r * D;
        const T* yr = y + r * D;
        T* sr = sum + r * D;
        float v[NV][kVec];
        float s = 0.f;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (FULL || vi < nvec) {
                float a[kVec], b[kVec];
                unpack16<T>(ldg16(xr + vi * kVec), a);
                unpack16<T>(ldg16(yr + vi * kVec), b);
#pragma unroll
                for (int k = 0; k < kVec; ++k) v[i][k] = Elem<T>::to_f(Elem<T>::from_f(a[k] + b[k]));   // the stored (rounded) sum
                *reinterpret_cast<uint4*>(sr + vi * kVec) = pack16<T>(v[i]);
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
            if (FULL || lane + 32 * i < nvec) {
#pragma unroll
                for (int k = 0; k < kVec; ++k) { const float d = v[i][k] - mean; q = fmaf(d, d, q); }
            }
        }
        const float rstd = rsqrtf(warp_sum(q) / (float)D + eps);
        T* orow = out + r * D;
#pragma unroll
        for (int i = 0; i < NV; ++i) {
            const int vi = lane + 32 * i;
            if (FULL || vi < nvec) {
                float g[kVec], b[kVec], o[kVec];
                unpack16<T>(ldg16(gamma + vi * kVec), g);
                unpack16<T>(ldg16(beta + vi * kVec), b);
#pragma unroll
                for (int k = 0; k < kVec; ++k) o[k] = fmaf((v[i][k] - mean) * rstd, g[k], b[k]);
                store_row_vec<T, kVec>(orow + vi * kVec, o);
            }
        }
    }
}

template <typename T>
static int launch_add_layernorm(const void* x, const void* y, const void* g, const void* b, void* sum, void* out, long long rows, int D,
                                float eps, cudaStream_t st) {
    constexpr int kVec = Elem<T>::kVec;
    const int nv = cdiv(D / kVec, 32);
    long long blocks = (rows + 7) / 8;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
#define ROMA_ALN(NVC) do { if (D == 32 * NVC * kVec) add_layernorm_kernel<T, NVC, true><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, (const T*)y, (const T*)g, (const T*)b, (T*)sum, (T*)out, rows, D, eps); \
                           else add_layernorm_kernel<T, NVC, false><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, (const T*)y, (const T*)g, (const T*)b, (T*)sum, (T*)out, rows, D, eps); } while (0)
    if (nv <= 1) ROMA_ALN(1);
    else if (nv <= 2) ROMA_ALN(2);
    else if (nv <= 4) ROMA_ALN(4);
    else if (nv <= 8) ROMA_ALN(8);
    else return fail(ROMA_ERR_UNSUPPORTED, "roma_add_layernorm: D = %d too wide (max %d)", D, 8 * 32 * kVec);
#undef ROMA_ALN
    return 0;
}

template <typename T, typename TO = T>
static int launch_layernorm(const void* x, const float* shift, const void* g, const void* b, void* out, long long rows, int D, float eps, cudaStream_t st) {
    constexpr int kVec = Elem<T>::kVec;
    const int nv = cdiv(D / kVec, 32);
    long long blocks = (rows + 7) / 8;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
#define ROMA_LN(NVC) do { if (D == 32 * NVC * kVec) layernorm_kernel<T, NVC, TO, true><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, shift, (const T*)g, (const T*)b, (TO*)out, rows, D, eps); \
                          else layernorm_kernel<T, NVC, TO, false><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, shift, (const T*)g, (const T*)b, (TO*)out, rows, D, eps); } while (0)
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

// LayerNorm of fp32 rows written in a 16-bit dtype (out_dtype = ROMA_F16 / ROMA_BF16): what autocast computes as an fp32
// layer_norm followed by a cast in front of the consuming Linear (transformer/layers/block.py:88-101 under the autocast of
// transformer/__init__.py), in one pass.  gamma, beta fp32.
extern "C" int roma_layernorm_cast(const void* x, const void* gamma, const void* beta, void* out,
                                   long long rows, int D, float eps, int out_dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && gamma && beta && out, ROMA_ERR_BAD_ARG, "roma_layernorm_cast: null pointer");
    ROMA_REQUIRE(rows > 0 && D > 0 && D % 4 == 0, ROMA_ERR_BAD_ARG, "roma_layernorm_cast: bad sizes (D %% 4 == 0)");
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta)) & 15u) == 0 &&
                 (reinterpret_cast<uintptr_t>(out) & 7u) == 0, ROMA_ERR_ALIGNMENT, "roma_layernorm_cast: inputs 16-byte, output 8-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int rc;
    switch (out_dtype) {
        case ROMA_F16:  rc = launch_layernorm<float, __half>(x, nullptr, gamma, beta, out, rows, D, eps, st); break;
        case ROMA_BF16: rc = launch_layernorm<float, __nv_bfloat16>(x, nullptr, gamma, beta, out, rows, D, eps, st); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_layernorm_cast: out_dtype must be ROMA_F16 or ROMA_BF16");
    }
    if (rc) return rc;
    return check_launch("roma_layernorm_cast");
}

// out = x + y with x fp32 and y 16-bit (the residual add of an fp32 token stream and an autocast branch), fp32 result:
// ATen runs this mixed-dtype add unvectorised (85 us for 26 M elements); here 16-byte loads of x and 8-byte loads of y.
namespace roma {
template <typename TY>
__global__ void __launch_bounds__(256) residual_add_kernel(const float* __restrict__ x, const TY* __restrict__ y, float* __restrict__ out, long long n4) {
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
        const float4 a = reinterpret_cast<const float4*>(x)[i];
        const uint2 raw = reinterpret_cast<const uint2*>(y)[i];
        const TY* t = reinterpret_cast<const TY*>(&raw);
        reinterpret_cast<float4*>(out)[i] = make_float4(a.x + Elem<TY>::to_f(t[0]), a.y + Elem<TY>::to_f(t[1]),
                                                       a.z + Elem<TY>::to_f(t[2]), a.w + Elem<TY>::to_f(t[3]));
    }
}
}  // namespace roma

extern "C" int roma_residual_add(const float* x, const void* y, float* out, long long n, int y_dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && y && out && n > 0 && n % 4 == 0, ROMA_ERR_BAD_ARG, "roma_residual_add: bad arguments (n %% 4 == 0)");
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0 && (reinterpret_cast<uintptr_t>(y) & 7u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_residual_add: x / out 16-byte, y 8-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const long long n4 = n / 4;
    long long blocks = (n4 + 255) / 256;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (y_dtype == ROMA_F16) residual_add_kernel<__half><<<(unsigned)blocks, 256, 0, st>>>(x, (const __half*)y, out, n4);
    else if (y_dtype == ROMA_BF16) residual_add_kernel<__nv_bfloat16><<<(unsigned)blocks, 256, 0, st>>>(x, (const __nv_bfloat16*)y, out, n4);
    else return fail(ROMA_ERR_BAD_ARG, "roma_residual_add: y_dtype must be ROMA_F16 or ROMA_BF16");
    return check_launch("roma_residual_add");
}

// sum = x + y and out = LayerNorm(sum) in one pass (x, y, sum, out, gamma, beta of one dtype; sum may alias x).
extern "C" int roma_add_layernorm(const void* x, const void* y, const void* gamma, const void* beta, void* sum, void* out,
                                  long long rows, int D, float eps, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && y && gamma && beta && sum && out, ROMA_ERR_BAD_ARG, "roma_add_layernorm: null pointer");
    ROMA_REQUIRE(rows > 0 && D > 0, ROMA_ERR_BAD_ARG, "roma_add_layernorm: bad sizes");
    const int kvec = dtype == ROMA_F32 ? 4 : 8;
    ROMA_REQUIRE(D % kvec == 0, ROMA_ERR_UNSUPPORTED, "roma_add_layernorm: D = %d must be a multiple of %d", D, kvec);
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y) | reinterpret_cast<uintptr_t>(sum) |
                   reinterpret_cast<uintptr_t>(out) | reinterpret_cast<uintptr_t>(gamma) | reinterpret_cast<uintptr_t>(beta)) & 15u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_add_layernorm: pointers must be 16-byte aligned");
    ROMA_REQUIRE(out != x && out != y && out != sum, ROMA_ERR_BAD_ARG, "roma_add_layernorm: out must not alias an input or sum");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int rc;
    switch (dtype) {
        case ROMA_F32:  rc = launch_add_layernorm<float>(x, y, gamma, beta, sum, out, rows, D, eps, st); break;
        case ROMA_F16:  rc = launch_add_layernorm<__half>(x, y, gamma, beta, sum, out, rows, D, eps, st); break;
        case ROMA_BF16: rc = launch_add_layernorm<__nv_bfloat16>(x, y, gamma, beta, sum, out, rows, D, eps, st); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_add_layernorm: unknown dtype %d", dtype);
    }
    if (rc) return rc;
    return check_launch("roma_add_layernorm");
}
