// Testing aids for K3b (tests/test_kde_exact_gpu.py, tools/explore_kde_exact.py): the squared distance of ATen's
// _euclidean_dist — d2(i, j) = half([-2a_i, |a_i|^2, 1] . [a_j, 1, |a_j|^2]) — for ALL pairs of a small point set, by each
// candidate arithmetic form, so that the form cuBLAS's half GEMM uses on this GPU can be identified bit for bit
// against `x1_.matmul(x2_.mT)` (romatch/utils/kde.py:9-11 -> torch.cdist -> _euclidean_dist).
#include "common.cuh"

namespace roma {

__device__ __forceinline__ float rhd(float v) { return __half2float(__float2half_rn(v)); }

__device__ __forceinline__ uint32_t pack_h2d(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

// element k (0..7) of the 6-wide augmented vectors, zero padded: kind 0 = x1_ = [-2a, |a|^2, 1], kind 1 = x2_ = [a, 1, |a|^2]
__device__ __forceinline__ float aug(const float* __restrict__ x, int p, int N, int k, int kind) {
    if (p >= N || k >= 6) return 0.f;
    const float4 v4 = reinterpret_cast<const float4*>(x)[p];
    const float v[4] = {rhd(v4.x), rhd(v4.y), rhd(v4.z), rhd(v4.w)};
    const float n = rhd(rhd(v[0] * v[0]) + rhd(v[1] * v[1]) + rhd(v[2] * v[2]) + rhd(v[3] * v[3]));
    if (kind == 0) return k < 4 ? -2.f * v[k] : (k == 4 ? n : 1.f);
    return k < 4 ? v[k] : (k == 4 ? 1.f : n);
}

// variant 0: sequential fp32 chain (the order of kde_partial_kernel<true>)
// variant 1: mma.sync m16n8k8  f16, A rows = x1_[i], B cols = x2_[j]
// variant 2: mma.sync m16n8k16 f16, same operands, K zero padded to 16
// variant 3: mma.sync m16n8k16 f16, roles exchanged (A rows = x2_[j], B cols = x1_[i]) — the operand order cuBLAS sees
//            for a row-major x1_ @ x2_^T
// variant 4: as 1 with roles exchanged
// out[i * N + j] = half bits
template <int VAR>
__global__ void kde_debug_d2_kernel(const float* __restrict__ x, int N, unsigned short* __restrict__ out) {
    const int lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
    const int m0 = (blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5)) * 16;
    if (m0 >= N) return;
    if (VAR == 0) {
        for (int n0 = 0; n0 < N; n0 += 8) {
            for (int e = lane; e < 128; e += 32) {
                const int i = m0 + (e >> 3), j = n0 + (e & 7);
                if (i >= N || j >= N) continue;
                float d2 = aug(x, i, N, 0, 0) * aug(x, j, N, 0, 1);
                d2 = fmaf(aug(x, i, N, 1, 0), aug(x, j, N, 1, 1), d2);
                d2 = fmaf(aug(x, i, N, 2, 0), aug(x, j, N, 2, 1), d2);
                d2 = fmaf(aug(x, i, N, 3, 0), aug(x, j, N, 3, 1), d2);
                d2 = d2 + aug(x, i, N, 4, 0);
                d2 = d2 + aug(x, j, N, 5, 1);
                out[(long long)i * N + j] = __half_as_ushort(__float2half_rn(d2));
            }
        }
        return;
    }
    constexpr bool kSwap = (VAR == 3 || VAR == 4);
    constexpr bool kK16 = (VAR == 2 || VAR == 3);
    const int akind = kSwap ? 1 : 0, bkind = kSwap ? 0 : 1;
    // A fragment: rows m0 + g (a0) and m0 + g + 8 (a1), k = 2 t4, 2 t4 + 1
    const uint32_t a0 = pack_h2d(aug(x, m0 + g, N, 2 * t4, akind), aug(x, m0 + g, N, 2 * t4 + 1, akind));
    const uint32_t a1 = pack_h2d(aug(x, m0 + g + 8, N, 2 * t4, akind), aug(x, m0 + g + 8, N, 2 * t4 + 1, akind));
    for (int n0 = 0; n0 < N; n0 += 8) {
        const uint32_t b0 = pack_h2d(aug(x, n0 + g, N, 2 * t4, bkind), aug(x, n0 + g, N, 2 * t4 + 1, bkind));
        float d[4];
        if (kK16) {
            asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%10, %10, %10, %10};"
                         : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3])
                         : "r"(a0), "r"(a1), "r"(0u), "r"(0u), "r"(b0), "r"(0u), "f"(0.f));
        } else {
            asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%7, %7, %7, %7};"
                         : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3]) : "r"(a0), "r"(a1), "r"(b0), "f"(0.f));
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const int m = m0 + g + 8 * (e >> 1), n = n0 + 2 * t4 + (e & 1);
            if (m >= N || n >= N) continue;
            const unsigned short bits = __half_as_ushort(__float2half_rn(d[e]));
            if (kSwap) out[(long long)n * N + m] = bits;      // M index is j, N index is i
            else       out[(long long)m * N + n] = bits;
        }
    }
}

// the table of the half chain after d2: clamp_min(0), sqrt, **2, neg, * (1 / (2 std^2)), exp — each rounded to half
__global__ void kde_debug_lut_kernel(float* __restrict__ lut, float inv_two_var) {
    const int bits = blockIdx.x * blockDim.x + threadIdx.x;
    if (bits > 0x7C00) return;
    const float d2 = __half2float(__ushort_as_half((unsigned short)bits));
    const float d = rhd(sqrtf(fmaxf(d2, 0.f)));
    const float s = rhd(d * d);
    const float a = rhd(-s * inv_two_var);
    lut[bits] = rhd(expf(a));
}

}  // namespace roma

extern "C" int roma_kde_debug_d2(const float* x, int N, int variant, void* out_half_bits, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && out_half_bits && N > 0 && N <= 16384, ROMA_ERR_BAD_ARG, "roma_kde_debug_d2: bad arguments (N <= 16384)");
    ROMA_REQUIRE(variant >= 0 && variant <= 4, ROMA_ERR_BAD_ARG, "roma_kde_debug_d2: variant %d", variant);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    unsigned short* out = reinterpret_cast<unsigned short*>(out_half_bits);
    const int warps = 4, grid = cdiv(cdiv(N, 16), warps);
    switch (variant) {
        case 0: kde_debug_d2_kernel<0><<<grid, warps * 32, 0, st>>>(x, N, out); break;
        case 1: kde_debug_d2_kernel<1><<<grid, warps * 32, 0, st>>>(x, N, out); break;
        case 2: kde_debug_d2_kernel<2><<<grid, warps * 32, 0, st>>>(x, N, out); break;
        case 3: kde_debug_d2_kernel<3><<<grid, warps * 32, 0, st>>>(x, N, out); break;
        default: kde_debug_d2_kernel<4><<<grid, warps * 32, 0, st>>>(x, N, out); break;
    }
    return check_launch("roma_kde_debug_d2");
}

extern "C" int roma_kde_debug_lut(float* lut, double std, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(lut && std > 0.0, ROMA_ERR_BAD_ARG, "roma_kde_debug_lut: bad arguments");
    const float two_var = (float)(2.0 * std * std);
    kde_debug_lut_kernel<<<(0x7C00 + 256) / 256, 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(lut, 1.0f / two_var);
    return check_launch("roma_kde_debug_lut");
}
