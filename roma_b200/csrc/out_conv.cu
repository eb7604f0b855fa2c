// (f-2) ConvRefiner head: out_conv (1x1 convolution, hidden_dim -> 3) evaluated in fp32 straight from the
// channels-last activation in its storage dtype.
//
// Replaces (reference): romatch/models/matcher.py:141 `d = self.out_conv(d.float())` — an fp32 copy of the whole
// activation (a full extra write + read) followed by a cudnn 1x1 convolution and its bias add.  Here the
// activation is read exactly once (16-byte vectors), up-converted in registers (exact), and reduced with fp32
// FMAs: HBM-bound, algorithmic bytes = B*h*w*(C*sizeof(T) + 16).
//
// Mapping: a group of G lanes (G = 1..32, power of two) owns one pixel at a time; lane l of the group holds the
// weights of channel vectors l, l + G, ... (NV per lane) in registers for the whole kernel and walks pixels in a
// grid-stride loop, two pixels in flight.  Partial sums are combined with xor-shuffles inside the group.
// Output: fp32 (npix, 4) interleaved (dx, dy, dcert, 0) so that every pixel is one 16-byte store; the consumer
// (`roma_flow_update_resample`) takes explicit strides.
#include <atomic>
#include "common.cuh"

namespace roma {

constexpr int kOutMax = 3;

template <typename T, int NV, int PX>
__global__ void __launch_bounds__(256)
out_conv_kernel(const T* __restrict__ x, const float* __restrict__ wgt, const float* __restrict__ bias, float4* __restrict__ out,
                long long npix, int C, int OUT, int G, int gshift) {
    constexpr int kVec = Elem<T>::kVec;
    const int lane = threadIdx.x & 31;
    const int gl = lane & (G - 1);                       // lane inside the group
    const int nvec = C / kVec;
    float wv[NV][kOutMax][kVec];
#pragma unroll
    for (int i = 0; i < NV; ++i) {
        const int v = gl + i * G;
#pragma unroll
        for (int o = 0; o < kOutMax; ++o)
#pragma unroll
            for (int k = 0; k < kVec; ++k) wv[i][o][k] = (v < nvec && o < OUT) ? __ldg(wgt + (long long)o * C + v * kVec + k) : 0.f;
    }
    const float b0 = OUT > 0 ? __ldg(bias + 0) : 0.f, b1 = OUT > 1 ? __ldg(bias + 1) : 0.f, b2 = OUT > 2 ? __ldg(bias + 2) : 0.f;

    // PX pixels in flight per group (more for narrow rows, where one pixel is a single 16-byte load per lane)
    const long long groups_total = ((long long)gridDim.x * blockDim.x) >> gshift;
    // the trip count is uniform per warp (the shuffles below need all 32 lanes): a group past the end idles
    const long long warp_first = (((long long)blockIdx.x * blockDim.x + threadIdx.x) & ~31LL) >> gshift;
    for (long long base = warp_first; base < npix; base += PX * groups_total) {
        const long long p0 = base + (lane >> gshift);
        uint4 raw[PX][NV];
#pragma unroll
        for (int u = 0; u < PX; ++u) {
            const long long p = p0 + u * groups_total;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                const int v = gl + i * G;
                raw[u][i] = (v < nvec && p < npix) ? ldg16(x + p * C + v * kVec) : make_uint4(0, 0, 0, 0);
            }
        }
        float s[PX][kOutMax];
#pragma unroll
        for (int u = 0; u < PX; ++u) {
#pragma unroll
            for (int o = 0; o < kOutMax; ++o) s[u][o] = 0.f;
#pragma unroll
            for (int i = 0; i < NV; ++i) {
                float f[kVec];
                unpack16<T>(raw[u][i], f);
#pragma unroll
                for (int o = 0; o < kOutMax; ++o)
#pragma unroll
                    for (int k = 0; k < kVec; ++k) s[u][o] = fmaf(f[k], wv[i][o][k], s[u][o]);
            }
        }
        for (int off = G >> 1; off > 0; off >>= 1) {
#pragma unroll
            for (int u = 0; u < PX; ++u)
#pragma unroll
                for (int o = 0; o < kOutMax; ++o) s[u][o] += __shfl_xor_sync(0xffffffffu, s[u][o], off);
        }
        if (gl == 0) {
#pragma unroll
            for (int u = 0; u < PX; ++u) {
                const long long p = p0 + u * groups_total;
                if (p < npix) out[p] = make_float4(s[u][0] + b0, s[u][1] + b1, s[u][2] + b2, 0.f);
            }
        }
    }
}

// Any width: one warp per pixel, weights re-read through L1 (used only when the register-resident variant does
// not fit, i.e. C > 9 * 32 vectors: fp32 activations wider than 1152 channels).
template <typename T>
__global__ void __launch_bounds__(256)
out_conv_wide_kernel(const T* __restrict__ x, const float* __restrict__ wgt, const float* __restrict__ bias, float4* __restrict__ out,
                     long long npix, int C, int OUT) {
    constexpr int kVec = Elem<T>::kVec;
    const int lane = threadIdx.x & 31;
    const int nvec = C / kVec;
    const long long warps_total = ((long long)gridDim.x * blockDim.x) >> 5;
    for (long long p = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5; p < npix; p += warps_total) {
        float s[kOutMax] = {0.f, 0.f, 0.f};
        for (int v = lane; v < nvec; v += 32) {
            float f[kVec];
            unpack16<T>(ldg16(x + p * C + v * kVec), f);
#pragma unroll
            for (int o = 0; o < kOutMax; ++o) {
                if (o < OUT) {
#pragma unroll
                    for (int k = 0; k < kVec; ++k) s[o] = fmaf(f[k], __ldg(wgt + (long long)o * C + v * kVec + k), s[o]);
                }
            }
        }
#pragma unroll
        for (int o = 0; o < kOutMax; ++o) s[o] = warp_sum(s[o]);
        if (lane == 0)
            out[p] = make_float4(s[0] + __ldg(bias), OUT > 1 ? s[1] + __ldg(bias + 1) : 0.f, OUT > 2 ? s[2] + __ldg(bias + 2) : 0.f, 0.f);
    }
}

template <typename T>
static int launch_out_conv(const void* x, const float* w, const float* b, float* out, long long npix, int C, int OUT, cudaStream_t st) {
    constexpr int kVec = Elem<T>::kVec;
    const int nvec = C / kVec;
    int G = 1, gshift = 0;
    while (G < 32 && G < nvec) { G <<= 1; ++gshift; }
    const int nv = cdiv(nvec, G);
    const int px = nv <= 1 ? 8 : nv <= 2 ? 4 : 2;
    const long long groups_wanted = (npix + px - 1) / px;
    const int groups_per_block = 256 / G;
    long long blocks = (groups_wanted + groups_per_block - 1) / groups_per_block;
    const long long cap = (long long)sm_count() * 8;                   // grid-stride beyond 8 resident blocks per SM
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
#define ROMA_OUT_CONV(NVC, PXC) out_conv_kernel<T, NVC, PXC><<<(unsigned)blocks, 256, 0, st>>>((const T*)x, w, b, (float4*)out, npix, C, OUT, G, gshift)
    if (nv <= 1) ROMA_OUT_CONV(1, 8);
    else if (nv <= 2) ROMA_OUT_CONV(2, 4);
    else if (nv <= 3) ROMA_OUT_CONV(3, 2);
    else if (nv <= 5) ROMA_OUT_CONV(5, 2);
    else if (nv <= 9) ROMA_OUT_CONV(9, 2);
    else {
        long long wb = (npix + 7) / 8;
        if (wb > cap) wb = cap;
        out_conv_wide_kernel<T><<<(unsigned)wb, 256, 0, st>>>((const T*)x, w, b, (float4*)out, npix, C, OUT);
    }
#undef ROMA_OUT_CONV
    return 0;
}

}  // namespace roma

extern "C" int roma_refiner_out_conv(const void* x, const float* weight, const float* bias, float* out,
                                     long long npix, int C, int OUT, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && weight && bias && out, ROMA_ERR_BAD_ARG, "roma_refiner_out_conv: null pointer");
    ROMA_REQUIRE(npix > 0 && C > 0 && OUT >= 1 && OUT <= kOutMax, ROMA_ERR_BAD_ARG, "roma_refiner_out_conv: bad sizes (OUT must be 1..3)");
    const int kvec = dtype == ROMA_F32 ? 4 : 8;
    ROMA_REQUIRE(C % kvec == 0, ROMA_ERR_UNSUPPORTED, "roma_refiner_out_conv: C = %d must be a multiple of %d (pad it)", C, kvec);
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0, ROMA_ERR_ALIGNMENT,
                 "roma_refiner_out_conv: x and out must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int rc;
    switch (dtype) {
        case ROMA_F32:  rc = launch_out_conv<float>(x, weight, bias, out, npix, C, OUT, st); break;
        case ROMA_F16:  rc = launch_out_conv<__half>(x, weight, bias, out, npix, C, OUT, st); break;
        case ROMA_BF16: rc = launch_out_conv<__nv_bfloat16>(x, weight, bias, out, npix, C, OUT, st); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_refiner_out_conv: unknown dtype %d", dtype);
    }
    if (rc) return rc;
    return check_launch("roma_refiner_out_conv");
}

// ---------------------------------------------------------------------------------------------------------------
// Narrow 1x1 convolution: out[p, o] = bias[o] + sum_c weight[o, c] * x[p, c] for OUT <= 32, output in the
// storage dtype, packed (npix, OUT).
//
// Replaces (reference): the finest-scale feature projection `proj["1"]` = Conv2d(64, 9, 1x1) + BatchNorm2d
//   (romatch/models/model_zoo/roma_models.py:131-142; the eval-mode BatchNorm is folded into weight / bias by the
//   caller), and the 24 -> 24 pointwise convolutions of the finest ConvRefiner (matcher.py:77-103, conv2).  cuBLAS runs this N = 9 GEMM at ~13 % of HBM speed; here one thread owns two pixels, streams their
//   channels as 16-byte vectors and keeps the OUT accumulators in registers, the weights (C x OUTP fp32) are
//   broadcast from shared memory.  HBM-bound: algorithmic bytes = npix * (C + OUT) * sizeof(T).
namespace roma {

constexpr int kNarrowMaxOut = 32;
static std::atomic<int> g_narrow_simt{0};     // testing aid: 1 forces the FFMA2 kernel for 16-bit inputs too

template <typename T, int OUTP>
__global__ void __launch_bounds__(128)
pointwise_narrow_kernel(const T* __restrict__ x, const float* __restrict__ wgt, const float* __restrict__ bias, T* __restrict__ out,
                        long long npix, int C, int OUT) {
    constexpr int kVec = Elem<T>::kVec;
    constexpr int NP = OUTP / 2;                       // output channel pairs: one FFMA2 each
    extern __shared__ float wsm[];                     // [C][OUTP], zero padded
    for (int i = threadIdx.x; i < C * OUTP; i += blockDim.x) {
        const int c = i / OUTP, o = i - c * OUTP;
        wsm[i] = o < OUT ? __ldg(wgt + (long long)o * C + c) : 0.f;
    }
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    const long long half = (npix + 1) / 2;
    for (long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x; t < half; t += stride) {
        const long long p0 = t, p1 = t + half;
        const bool has1 = p1 < npix;
        float2 a0[NP], a1[NP];
#pragma unroll
        for (int o = 0; o < NP; ++o)
            a0[o] = a1[o] = make_float2(2 * o < OUT ? __ldg(bias + 2 * o) : 0.f, 2 * o + 1 < OUT ? __ldg(bias + 2 * o + 1) : 0.f);
        const T* x0 = x + p0 * C;
        const T* x1 = x + (has1 ? p1 : p0) * C;
        for (int v = 0; v < C / kVec; ++v) {
            float f0[kVec], f1[kVec];
            unpack16<T>(ldg16(x0 + v * kVec), f0);
            unpack16<T>(ldg16(x1 + v * kVec), f1);
#pragma unroll
            for (int k = 0; k < kVec; ++k) {
                const float4* wr = reinterpret_cast<const float4*>(wsm + (v * kVec + k) * OUTP);
                const float2 d0 = make_float2(f0[k], f0[k]), d1 = make_float2(f1[k], f1[k]);
#pragma unroll
                for (int q = 0; q < OUTP / 4; ++q) {
                    const float4 w4 = wr[q];
                    const float2 wa = make_float2(w4.x, w4.y), wb = make_float2(w4.z, w4.w);
                    a0[2 * q] = __ffma2_rn(d0, wa, a0[2 * q]);         a1[2 * q] = __ffma2_rn(d1, wa, a1[2 * q]);
                    a0[2 * q + 1] = __ffma2_rn(d0, wb, a0[2 * q + 1]); a1[2 * q + 1] = __ffma2_rn(d1, wb, a1[2 * q + 1]);
                }
            }
        }
        T* o0 = out + p0 * OUT;
        T* o1 = out + p1 * OUT;
        if (OUT % kVec == 0 && OUTP % kVec == 0) {     // rows are whole 16-byte vectors (out is 16-byte aligned: checked on the host)
#pragma unroll
            for (int q = 0; q < OUTP / kVec; ++q) {
                if (q * kVec < OUT) {
                    float r0[kVec], r1[kVec];
#pragma unroll
                    for (int k = 0; k < kVec / 2; ++k) {
                        r0[2 * k] = a0[q * kVec / 2 + k].x; r0[2 * k + 1] = a0[q * kVec / 2 + k].y;
                        r1[2 * k] = a1[q * kVec / 2 + k].x; r1[2 * k + 1] = a1[q * kVec / 2 + k].y;
                    }
                    *reinterpret_cast<uint4*>(o0 + q * kVec) = pack16<T>(r0);
                    if (has1) *reinterpret_cast<uint4*>(o1 + q * kVec) = pack16<T>(r1);
                }
            }
        } else {
#pragma unroll
            for (int o = 0; o < OUTP; ++o) {
                if (o < OUT) {
                    o0[o] = Elem<T>::from_f((o & 1) ? a0[o / 2].y : a0[o / 2].x);
                    if (has1) o1[o] = Elem<T>::from_f((o & 1) ? a1[o / 2].y : a1[o / 2].x);
                }
            }
        }
    }
}

// ---- tensor-core variant (16-bit dtypes, C <= 64, OUT <= 32) -------------------------------------------------
// The op is a GEMM with N, K <= 64 over millions of rows: memory-bound as soon as the 576 MACs per pixel leave the
// FP32 pipe, so it runs on warp-level mma.sync m16n8k16 (fp32 accumulate) with the A fragments loaded straight
// from global memory (a quad of lanes reads 16 contiguous bytes of one pixel), the weights converted once per
// thread into B fragments and the C fragments stored as packed 16-bit pairs.  No shared memory, no barrier.
// (tcgen05 would need the A tile staged in shared memory and a TMEM round trip for a kernel whose roofline is HBM.)
template <typename T> __device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <> __device__ __forceinline__ void mma16816<__nv_bfloat16>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <> __device__ __forceinline__ void mma16816<__half>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <typename T> __device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <> __device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}
template <> __device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

constexpr int kMmaTiles = 2;           // 16-pixel tiles in flight per warp

template <typename T, int KS, int NT>
__global__ void __launch_bounds__(256)
narrow_mma_kernel(const T* __restrict__ x, const float* __restrict__ wgt, const float* __restrict__ bias, T* __restrict__ out,
                  long long npix, int C, int OUT) {
    const int lane = threadIdx.x & 31, g = lane >> 2, t = lane & 3;
    // B fragments: B[k][n] = W[n][k];  b0 = (k = 16 ks + 2t, +1 ; n = 8 nt + g), b1 = the same 8 further in k
    uint32_t bf[KS][NT][2];
    float bv[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
        const int n = nt * 8 + g;
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int k = ks * 16 + 2 * t + 8 * h;
                const float w0 = (n < OUT && k < C) ? __ldg(wgt + (long long)n * C + k) : 0.f;
                const float w1 = (n < OUT && k + 1 < C) ? __ldg(wgt + (long long)n * C + k + 1) : 0.f;
                bf[ks][nt][h] = pack2<T>(w0, w1);
            }
        }
        bv[nt][0] = nt * 8 + 2 * t < OUT ? __ldg(bias + nt * 8 + 2 * t) : 0.f;
        bv[nt][1] = nt * 8 + 2 * t + 1 < OUT ? __ldg(bias + nt * 8 + 2 * t + 1) : 0.f;
    }
    const bool pair_store = (OUT & 1) == 0;            // 4-byte aligned channel pairs
    const long long ntiles = (npix + 15) / 16;
    const long long wstride = (((long long)gridDim.x * blockDim.x) >> 5) * kMmaTiles;
    for (long long tile0 = ((((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5)) * kMmaTiles; tile0 < ntiles; tile0 += wstride) {
        uint32_t a[kMmaTiles][KS][4];
#pragma unroll
        for (int u = 0; u < kMmaTiles; ++u) {
            const long long plo = (tile0 + u) * 16 + g, phi = plo + 8;
            const uint32_t* rlo = reinterpret_cast<const uint32_t*>(x + plo * C);
            const uint32_t* rhi = reinterpret_cast<const uint32_t*>(x + phi * C);
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
                const int k0 = ks * 16 + 2 * t, k1 = k0 + 8;
                a[u][ks][0] = (plo < npix && k0 < C) ? __ldg(rlo + (k0 >> 1)) : 0u;
                a[u][ks][1] = (phi < npix && k0 < C) ? __ldg(rhi + (k0 >> 1)) : 0u;
                a[u][ks][2] = (plo < npix && k1 < C) ? __ldg(rlo + (k1 >> 1)) : 0u;
                a[u][ks][3] = (phi < npix && k1 < C) ? __ldg(rhi + (k1 >> 1)) : 0u;
            }
        }
#pragma unroll
        for (int u = 0; u < kMmaTiles; ++u) {
            const long long plo = (tile0 + u) * 16 + g, phi = plo + 8;
#pragma unroll
            for (int nt = 0; nt < NT; ++nt) {
                float c[4] = {bv[nt][0], bv[nt][1], bv[nt][0], bv[nt][1]};
#pragma unroll
                for (int ks = 0; ks < KS; ++ks) mma16816<T>(c, a[u][ks], bf[ks][nt][0], bf[ks][nt][1]);
                const int n = nt * 8 + 2 * t;
                if (pair_store) {
                    if (n < OUT) {
                        if (plo < npix) *reinterpret_cast<uint32_t*>(out + plo * OUT + n) = pack2<T>(c[0], c[1]);
                        if (phi < npix) *reinterpret_cast<uint32_t*>(out + phi * OUT + n) = pack2<T>(c[2], c[3]);
                    }
                } else {
                    if (plo < npix) {
                        if (n < OUT) out[plo * OUT + n] = Elem<T>::from_f(c[0]);
                        if (n + 1 < OUT) out[plo * OUT + n + 1] = Elem<T>::from_f(c[1]);
                    }
                    if (phi < npix) {
                        if (n < OUT) out[phi * OUT + n] = Elem<T>::from_f(c[2]);
                        if (n + 1 < OUT) out[phi * OUT + n + 1] = Elem<T>::from_f(c[3]);
                    }
                }
            }
        }
    }
}

template <typename T, int KS>
static void launch_narrow_mma_nt(const void* x, const float* w, const float* b, void* out, long long npix, int C, int OUT,
                                 unsigned blocks, cudaStream_t st) {
    const int nt = (OUT + 7) / 8;
    switch (nt) {
        case 1: narrow_mma_kernel<T, KS, 1><<<blocks, 256, 0, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        case 2: narrow_mma_kernel<T, KS, 2><<<blocks, 256, 0, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        case 3: narrow_mma_kernel<T, KS, 3><<<blocks, 256, 0, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        default: narrow_mma_kernel<T, KS, 4><<<blocks, 256, 0, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
    }
}

template <typename T>
static bool launch_narrow_mma(const void* x, const float* w, const float* b, void* out, long long npix, int C, int OUT, cudaStream_t st) {
    if constexpr (sizeof(T) != 2) return false;
    else {
        if (C > 64 || OUT > 32 || g_narrow_simt.load()) return false;
        const long long warps = ((npix + 15) / 16 + kMmaTiles - 1) / kMmaTiles;
        long long blocks = (warps + 7) / 8;
        const long long cap = (long long)sm_count() * 16;
        if (blocks > cap) blocks = cap;
        if (blocks < 1) blocks = 1;
        const int ks = (C + 15) / 16;
        if (ks <= 1) launch_narrow_mma_nt<T, 1>(x, w, b, out, npix, C, OUT, (unsigned)blocks, st);
        else if (ks <= 2) launch_narrow_mma_nt<T, 2>(x, w, b, out, npix, C, OUT, (unsigned)blocks, st);
        else launch_narrow_mma_nt<T, 4>(x, w, b, out, npix, C, OUT, (unsigned)blocks, st);
        return true;
    }
}

template <typename T>
static int launch_narrow(const void* x, const float* w, const float* b, void* out, long long npix, int C, int OUT, cudaStream_t st) {
    if (launch_narrow_mma<T>(x, w, b, out, npix, C, OUT, st)) return 0;
    int outp = (OUT + 3) / 4 * 4;
    if (outp == 20) outp = 24;
    if (outp == 28) outp = 32;
    const size_t smem = (size_t)C * outp * sizeof(float);
    ROMA_REQUIRE(smem <= 48 * 1024, ROMA_ERR_UNSUPPORTED, "roma_pointwise_narrow: C * OUT too large (%d x %d)", C, OUT);
    long long blocks = ((npix + 1) / 2 + 127) / 128;
    const long long cap = (long long)sm_count() * 16;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    switch (outp) {
        case 4:  pointwise_narrow_kernel<T, 4><<<(unsigned)blocks, 128, smem, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        case 8:  pointwise_narrow_kernel<T, 8><<<(unsigned)blocks, 128, smem, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        case 12: pointwise_narrow_kernel<T, 12><<<(unsigned)blocks, 128, smem, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        case 16: pointwise_narrow_kernel<T, 16><<<(unsigned)blocks, 128, smem, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        case 24: pointwise_narrow_kernel<T, 24><<<(unsigned)blocks, 128, smem, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
        default: pointwise_narrow_kernel<T, 32><<<(unsigned)blocks, 128, smem, st>>>((const T*)x, w, b, (T*)out, npix, C, OUT); break;
    }
    return 0;
}

}  // namespace roma

extern "C" int roma_pointwise_narrow_set_simt(int on) {
    roma::g_narrow_simt.store(on ? 1 : 0);
    return 0;
}

extern "C" int roma_pointwise_narrow(const void* x, const float* weight, const float* bias, void* out,
                                     long long npix, int C, int OUT, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && weight && bias && out, ROMA_ERR_BAD_ARG, "roma_pointwise_narrow: null pointer");
    ROMA_REQUIRE(npix > 0 && C > 0 && OUT >= 1 && OUT <= kNarrowMaxOut, ROMA_ERR_BAD_ARG, "roma_pointwise_narrow: bad sizes (OUT must be 1..32)");
    const int kvec = dtype == ROMA_F32 ? 4 : 8;
    ROMA_REQUIRE(C % kvec == 0, ROMA_ERR_UNSUPPORTED, "roma_pointwise_narrow: C = %d must be a multiple of %d", C, kvec);
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0, ROMA_ERR_ALIGNMENT,
                 "roma_pointwise_narrow: x and out must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int rc;
    switch (dtype) {
        case ROMA_F32:  rc = launch_narrow<float>(x, weight, bias, out, npix, C, OUT, st); break;
        case ROMA_F16:  rc = launch_narrow<__half>(x, weight, bias, out, npix, C, OUT, st); break;
        case ROMA_BF16: rc = launch_narrow<__nv_bfloat16>(x, weight, bias, out, npix, C, OUT, st); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_pointwise_narrow: unknown dtype %d", dtype);
    }
    if (rc) return rc;
    return check_launch("roma_pointwise_narrow");
}
