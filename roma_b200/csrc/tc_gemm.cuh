// tcgen05 / TMEM batched "NT" GEMM with split-bf16 operands:   C = epi(A * B^T)
//
//   A : M x K, B : N x K, both K-major (row-major with K contiguous) in global memory,
//       element type fp32 (split on the fly into SPLIT bf16 planes), bf16 or fp16 (SPLIT = 1, exact).
//   C : M x N fp32, optional transposed mirror Ct : N x M.
//
// One CTA computes one 128 x 128 tile: operand k-blocks (128 rows x 64 K) are staged in shared memory in
// the canonical K-major SWIZZLE_128B layout (row = 128 bytes, 16-byte chunk j of row r at chunk j ^ (r & 7),
// 8-row groups 1024 bytes apart), one elected thread issues tcgen05.mma (M=128, N=128, K=16, kind::f16) with
// the fp32 accumulator in TMEM (128 lanes x 128 columns), and all 8 warps drain TMEM with tcgen05.ld for
// the fused epilogue.  Staging is double buffered: the loads / bf16 splitting of k-block i+1 overlap the
// asynchronous MMAs of k-block i (completion tracked per stage by tcgen05.commit -> mbarrier).
//
// Split precision: an fp32 operand a is written as a = a0 + a1 (+ a2) with bf16 planes; the product keeps
// the terms a_i*b_j with i + j < SPLIT (3 MMAs for SPLIT=2, ~2^-16 relative; 6 MMAs for SPLIT=3, ~fp32),
// all accumulated in the same fp32 TMEM tile.  This is what lets the reference's true-fp32 GP (TF32 off,
// matcher.py:254-263) run on bf16 tensor cores without the 0.047 error of a plain bf16 Gram.
#pragma once
#include <atomic>
#include <type_traits>
#include "common.cuh"
#include "tma.cuh"

namespace roma {
namespace tc {

constexpr int BM = 128, BN = 128, BK = 64;
constexpr int kThreads = 256;
constexpr int kPlaneBytes = BM * BK * 2;          // 16 KB: one bf16 plane of one operand k-block
template <int SPLIT> constexpr int tmem_cols() { return SPLIT == 1 ? 128 : 256; }

enum Epi : int { EPI_STORE = 0, EPI_SUB = 1, EPI_COS = 2, EPI_TINY = 3 };

struct Problem {
    const void* A; const void* B;
    float* C; float* Ct;
    long long sA, sB, sC, sCt;          // batch strides in elements
    int M, N, K;                         // logical extents (loads beyond them read as zero)
    int lda, ldb, ldc, ldct;
    int tiles_m, tiles_n;
    int lower_only;                      // skip tiles strictly above the diagonal (tn > tm)
    int diag_only;                       // tile t is (t, t): only the diagonal tiles exist (tiles = tiles_m)
    int no_diag;                         // EPI_COS: no diag_add / identity padding (rectangular cross kernel)
    int epi;
    int store_m, store_n;                // store bounds (>= M, N when padded rows/cols must be written)
    // EPI_COS: K = exp((acc / (rn[m]*cn[n] + eps) - 1) * inv_T) (+ diag_add on m == n); padding -> identity
    const float* rn; const float* cn; long long srn, scn;
    float eps, inv_T, diag_add;
    // EPI_TINY (Tiny-RoMa corr_volume + pos_embed, tiny.py:178-199, 241-254): nothing is stored; every 64-column half
    // of a tile reduces its scores s = acc * scale per row m to (max, argmax, online-softmax triple over the columns of
    // the 4x4-strided subset) and writes the record to part[((b * slots + slot) * 6 + field) * M + m],
    // slot = 2 * tn + half, slots = 2 * tiles_n.  Column n is pixel (n / w1, n % w1) of image 1.
    float* part; float scale; int w1; float lx0, lxs, ly0, lys;
    // bf16 planes of the stored result (v = p0 + p1 (+ p2): 2^-17 / 2^-25 relative), same indexing as C (b * sC + m * ldc + n):
    // written by the epilogue when non-null, consumed as ready-made split operands by gemm_planes_kernel (below)
    uint16_t* Pout[3]; int nplanes;     // nplanes = 2 (hi, lo) or 3 (hi, mid, lo); 0 = no plane output
    // split operands of gemm_planes_kernel: the planes of A and B (pre-offset like A and B; element = bf16)
    const uint16_t* Apl[3]; const uint16_t* Bpl[3];
};

struct Launch {
    Problem p[2];
    int nprob;
    int tiles0;                          // tiles of problem 0 (problem 1 follows)
};

// ---------------------------------------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t ncols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T, bf16/fp16 inputs, fp32 accumulate, M=128, N=128, K=16
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
                 ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                   "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                   "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                   "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                 : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major SWIZZLE_128B shared-memory matrix descriptor (sm_100 "version 1"): SBO = 1024 B between 8-row groups.
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
    return (uint64_t)((saddr & 0x3FFFFu) >> 4) | ((uint64_t)(1024u >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor: fp32 accumulate, A/B format fmt (0 = f16, 1 = bf16), K-major both, N=128, M=128
__host__ __device__ constexpr uint32_t instr_desc(uint32_t fmt) {
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}

// ---------------------------------------------------------------------------------------------- operand staging
// A k-block of one operand (128 rows x 64 k) travels global -> registers -> shared; the two halves are split
// so the loads of k-block i+1 are in flight while k-block i is converted, stored and multiplied.
template <typename T, int SPLIT> struct Stager;

// fp32 source: split into SPLIT bf16 planes.  Thread handles 8 float4 (row = idx/16, 4 consecutive k).
template <int SPLIT> struct Stager<float, SPLIT> {
    struct Regs { float4 v[8]; };
    static __device__ __forceinline__ void load(Regs& r, const void* srcv, long long ld, int rows_valid, int k0, int K, int tid) {
        const float* src = reinterpret_cast<const float*>(srcv);
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx = it * kThreads + tid;
            const int row = idx >> 4, c4 = idx & 15;
            const int k = k0 + c4 * 4;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (row < rows_valid && k < K) {
                const float* p = src + (long long)row * ld + k;
                if (k + 4 <= K) v = *reinterpret_cast<const float4*>(p);   // plain load: panels are updated in place
                else { v.x = p[0]; if (k + 1 < K) v.y = p[1]; if (k + 2 < K) v.z = p[2]; }
            }
            r.v[it] = v;
        }
    }
    static __device__ __forceinline__ void store(const Regs& rg, unsigned char* planes, int tid) {
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int idx = it * kThreads + tid;
            const int row = idx >> 4, c4 = idx & 15;
            const uint32_t off = row * 128 + ((((c4 >> 1) ^ (row & 7)) << 4) | ((c4 & 1) << 3));
            float r[4] = {rg.v[it].x, rg.v[it].y, rg.v[it].z, rg.v[it].w};
#pragma unroll
            for (int s = 0; s < SPLIT; ++s) {
                __nv_bfloat16 h[4];
#pragma unroll
                for (int e = 0; e < 4; ++e) { h[e] = __float2bfloat16_rn(r[e]); r[e] -= __bfloat162float(h[e]); }
                uint2 pk;
                pk.x = (uint32_t)__bfloat16_as_ushort(h[0]) | ((uint32_t)__bfloat16_as_ushort(h[1]) << 16);
                pk.y = (uint32_t)__bfloat16_as_ushort(h[2]) | ((uint32_t)__bfloat16_as_ushort(h[3]) << 16);
                *reinterpret_cast<uint2*>(planes + s * kPlaneBytes + off) = pk;
            }
        }
    }
};
// 16-bit source (bf16 / fp16): straight 16-byte copies into the swizzled plane.  Needs ld % 8 == 0.
struct StagerB16 {
    struct Regs { uint4 v[4]; };
    static __device__ __forceinline__ void load(Regs& r, const void* srcv, long long ld, int rows_valid, int k0, int K, int tid) {
        const uint16_t* src = reinterpret_cast<const uint16_t*>(srcv);
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int idx = it * kThreads + tid;
            const int row = idx >> 3, j = idx & 7;
            const int k = k0 + j * 8;
            uint4 v = make_uint4(0u, 0u, 0u, 0u);
            if (row < rows_valid && k + 8 <= K) v = __ldg(reinterpret_cast<const uint4*>(src + (long long)row * ld + k));
            else if (row < rows_valid && k < K) {
                uint16_t* e = reinterpret_cast<uint16_t*>(&v);
                for (int t = 0; t < 8 && k + t < K; ++t) e[t] = src[(long long)row * ld + k + t];
            }
            r.v[it] = v;
        }
    }
    static __device__ __forceinline__ void store(const Regs& r, unsigned char* plane, int tid) {
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int idx = it * kThreads + tid;
            const int row = idx >> 3, j = idx & 7;
            *reinterpret_cast<uint4*>(plane + row * 128 + ((j ^ (row & 7)) << 4)) = r.v[it];
        }
    }
};
template <> struct Stager<__nv_bfloat16, 1> : StagerB16 {};
template <> struct Stager<__half, 1> : StagerB16 {};

// 16-bit operands: 2 stages x 32 KB (3 CTAs/SM); split fp32 operands: 1 stage of 64 / 96 KB (3 / 2 CTAs/SM) —
// co-resident CTAs overlap one tile's epilogue and operand staging with another tile's MMAs.
template <int SPLIT> constexpr int num_stages() { return SPLIT == 1 ? 2 : 1; }
template <int SPLIT> constexpr size_t smem_bytes() { return (size_t)num_stages<SPLIT>() * 2 * SPLIT * kPlaneBytes + 1024; }

// ---------------------------------------------------------------------------------------------- epilogue
// Warp w drains lanes 32*(w&3).., columns 64*(w>>2).. of the 128 x 128 accumulator tile and applies the fused
// epilogue (store / subtract-from-C / cosine kernel).  Shared by the register-staged and the TMA-fed main loops.
template <int SPLIT>
__device__ __forceinline__ void tile_epilogue(const Problem& P, int b, int m0, int n0, uint32_t tmem_base, int warp, int lane) {
    const int q = warp & 3, hf = warp >> 2;
    const int m = m0 + q * 32 + lane;
    float* Cb = P.C ? P.C + (long long)b * P.sC : nullptr;
    float* Ctb = P.Ct ? P.Ct + (long long)b * P.sCt : nullptr;
    if (P.epi == EPI_TINY) {
        float best = -INFINITY; int bi = 0x7fffffff;
        float mx = -INFINITY, ssum = 0.f, sx = 0.f, sy = 0.f;
#pragma unroll 1
        for (int cc = 0; cc < 2; ++cc) {
            const int c0 = hf * 64 + cc * 32;
            float v[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
            if (SPLIT > 1) {
                float lo[32];
                tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + 128u + (uint32_t)c0, lo);
#pragma unroll
                for (int i = 0; i < 32; ++i) v[i] += lo[i];
            }
            const int nb0 = n0 + c0;
            int y1 = nb0 / P.w1, x1 = nb0 - y1 * P.w1;            // the same for every row: warp-uniform branches below
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                if (nb0 + i < P.N) {
                    const float sj = v[i] * P.scale;
                    if (sj > best) { best = sj; bi = nb0 + i; }        // columns ascend: the first maximum is kept
                    if (((x1 | y1) & 3) == 0) {
                        const float gx = P.lx0 + P.lxs * (float)(x1 >> 2), gy = P.ly0 + P.lys * (float)(y1 >> 2);
                        if (sj > mx) { const float sc = expf(mx - sj); ssum *= sc; sx *= sc; sy *= sc; mx = sj; }
                        const float e = expf(sj - mx);
                        ssum += e; sx = fmaf(e, gx, sx); sy = fmaf(e, gy, sy);
                    }
                }
                if (++x1 == P.w1) { x1 = 0; ++y1; }
            }
        }
        if (m < P.M) {
            const int slots = 2 * P.tiles_n, slot = 2 * (n0 / BN) + hf;
            float* rec = P.part + ((long long)b * slots + slot) * 6 * P.M + m;
            rec[0] = best; rec[(long long)P.M] = __int_as_float(bi); rec[2ll * P.M] = mx;
            rec[3ll * P.M] = ssum; rec[4ll * P.M] = sx; rec[5ll * P.M] = sy;
        }
        return;
    }
    float rn = 0.f;
    if (P.epi == EPI_COS && m < P.M) rn = P.rn[(long long)b * P.srn + m];
#pragma unroll 1
    for (int cc = 0; cc < 2; ++cc) {
        const int c0 = hf * 64 + cc * 32;
        float v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c0, v);
        if (SPLIT > 1) {
            float lo[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + 128u + (uint32_t)c0, lo);
#pragma unroll
            for (int i = 0; i < 32; ++i) v[i] += lo[i];
        }
        const int nb0 = n0 + c0;
        if (P.epi == EPI_COS) {
            // column norms of this 32-column group: one coalesced load, broadcast by shuffle
            const float cn_lane = (nb0 + lane < P.N) ? __ldg(P.cn + (long long)b * P.scn + nb0 + lane) : 1.f;
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                const int n = nb0 + i;
                const float cn_i = __shfl_sync(0xffffffffu, cn_lane, i);
                float k;
                if (m < P.M && n < P.N) {
                    const float c = v[i] / (rn * cn_i + P.eps);
                    k = expf((c - 1.f) * P.inv_T);
                    if (m == n && !P.no_diag) k += P.diag_add;
                } else {
                    k = (m == n && !P.no_diag) ? 1.f : 0.f;     // padding: identity block keeps the factorisation trivial there
                }
                v[i] = k;
            }
        }
        if (Cb && m < P.store_m) {
            float* crow = Cb + (long long)m * P.ldc + nb0;
            if (nb0 + 32 <= P.store_n && (P.ldc & 3) == 0) {
#pragma unroll
                for (int i = 0; i < 32; i += 4) {
                    float4 o = make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]);
                    if (P.epi == EPI_SUB) {
                        const float4 old = *reinterpret_cast<const float4*>(crow + i);
                        o = make_float4(old.x - o.x, old.y - o.y, old.z - o.z, old.w - o.w);
                        v[i] = o.x; v[i + 1] = o.y; v[i + 2] = o.z; v[i + 3] = o.w;
                    }
                    *reinterpret_cast<float4*>(crow + i) = o;
                }
            } else {
#pragma unroll
                for (int i = 0; i < 32; ++i) {
                    if (nb0 + i < P.store_n) {
                        if (P.epi == EPI_SUB) v[i] = crow[i] - v[i];
                        crow[i] = v[i];
                    }
                }
            }
        }
        if (Ctb && m < P.store_m) {
#pragma unroll
            for (int i = 0; i < 32; ++i)
                if (nb0 + i < P.store_n) Ctb[(long long)(nb0 + i) * P.ldct + m] = v[i];      // lanes -> consecutive m: coalesced
        }
        if (P.nplanes && m < P.store_m && nb0 + 32 <= P.store_n) {                           // finished panel -> bf16 planes
            const long long off = (long long)b * P.sC + (long long)m * P.ldc + nb0;
#pragma unroll
            for (int pl = 0; pl < 3; ++pl) {
                if (pl >= P.nplanes) break;
                uint16_t* dst = P.Pout[pl] + off;
#pragma unroll
                for (int i = 0; i < 32; i += 8) {
                    uint4 hv;
                    uint32_t* h32 = reinterpret_cast<uint32_t*>(&hv);
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        const __nv_bfloat16 h0 = __float2bfloat16_rn(v[i + 2 * e]), h1 = __float2bfloat16_rn(v[i + 2 * e + 1]);
                        v[i + 2 * e] -= __bfloat162float(h0);                                  // the remainder feeds the next plane
                        v[i + 2 * e + 1] -= __bfloat162float(h1);
                        h32[e] = (uint32_t)__bfloat16_as_ushort(h0) | ((uint32_t)__bfloat16_as_ushort(h1) << 16);
                    }
                    *reinterpret_cast<uint4*>(dst + i) = hv;
                }
            }
        }
    }
}

// ---------------------------------------------------------------------------------------------- kernel
template <typename T, int SPLIT>
__global__ void __launch_bounds__(kThreads, 2) gemm_nt_kernel(const Launch L) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t mma_bar[2];
    __shared__ uint32_t tmem_slot;
    constexpr int STAGES = num_stages<SPLIT>();

    // ---- tile decode ----
    int t = blockIdx.x;
    const int pi = (L.nprob > 1 && t >= L.tiles0) ? 1 : 0;
    if (pi) t -= L.tiles0;
    const Problem& P = L.p[pi];
    const int tm = P.diag_only ? t : t / P.tiles_n, tn = P.diag_only ? t : t - tm * P.tiles_n;
    if (P.lower_only && tn > tm) return;
    const int b = blockIdx.y;
    const int m0 = tm * BM, n0 = tn * BN;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
    auto stageA = [&](int s) { return smem + (size_t)s * (2 * SPLIT * kPlaneBytes); };
    auto stageB = [&](int s) { return smem + (size_t)s * (2 * SPLIT * kPlaneBytes) + SPLIT * kPlaneBytes; };

    const char* Ab = reinterpret_cast<const char*>(P.A) + ((long long)b * P.sA + (long long)m0 * P.lda) * (long long)sizeof(T);
    const char* Bb = reinterpret_cast<const char*>(P.B) + ((long long)b * P.sB + (long long)n0 * P.ldb) * (long long)sizeof(T);
    const int rowsA = P.M - m0 < BM ? P.M - m0 : BM;
    const int rowsB = P.N - n0 < BN ? P.N - n0 : BN;
    const int nkb = (P.K + BK - 1) / BK;
    constexpr uint32_t kFmt = std::is_same<T, __half>::value ? 0u : 1u;
    constexpr uint32_t idesc = instr_desc(kFmt);

    using St = Stager<T, SPLIT>;
    typename St::Regs ra, rb;
    St::load(ra, Ab, P.lda, rowsA, 0, P.K, tid);           // first k-block in flight during the TMEM / barrier setup
    St::load(rb, Bb, P.ldb, rowsB, 0, P.K, tid);

    if (warp == 0) tmem_alloc(&tmem_slot, tmem_cols<SPLIT>());
    if (tid == 0) { mbar_init(&mma_bar[0], 1); mbar_init(&mma_bar[1], 1); fence_barrier_init(); }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    for (int kb = 0; kb < nkb; ++kb) {
        const int s = kb % STAGES;
        if (kb >= STAGES) mbar_wait(&mma_bar[s], (uint32_t)((kb / STAGES - 1) & 1));    // MMAs that read this stage are done
        St::store(ra, stageA(s), tid);
        St::store(rb, stageB(s), tid);
        if (kb + 1 < nkb) {
            St::load(ra, Ab, P.lda, rowsA, (kb + 1) * BK, P.K, tid);
            St::load(rb, Bb, P.ldb, rowsB, (kb + 1) * BK, P.K, tid);
        }
        fence_proxy_async();
        __syncthreads();
        if (tid == 0) {
            tc_fence_after();
            const uint32_t a0 = smem_u32(stageA(s)), b0 = smem_u32(stageB(s));
#pragma unroll
            for (int kk = 0; kk < BK / 16; ++kk) {
#pragma unroll
                for (int i = 0; i < SPLIT; ++i) {
#pragma unroll
                    for (int j = 0; j < SPLIT; ++j) {
                        if (i + j >= SPLIT) continue;
                        const uint64_t ad = smem_desc_sw128(a0 + i * kPlaneBytes + kk * 32);
                        const uint64_t bd = smem_desc_sw128(b0 + j * kPlaneBytes + kk * 32);
                        // The tensor core's fp32 accumulate truncates, so every extra MMA into the same tile costs
                        // ~2^-24 of the ACCUMULATOR's magnitude.  The small cross terms (i + j > 0) therefore get
                        // their own TMEM tile (columns 128..255): their rounding is relative to a 2^-8 smaller sum,
                        // and the leading a0*b0 tile sees exactly the K/16 accumulations of a plain bf16 GEMM.
                        if (i + j == 0) umma_f16(tmem_base, ad, bd, idesc, (kb | kk) ? 1u : 0u);
                        else umma_f16(tmem_base + 128u, ad, bd, idesc, (kb | kk | (i + j - 1) | i) ? 1u : 0u);
                    }
                }
            }
            umma_commit(&mma_bar[s]);
        }
    }
    {   // all MMAs complete when the last commit fires
        const int last = nkb - 1;
        mbar_wait(&mma_bar[last % STAGES], (uint32_t)((last / STAGES) & 1));
        tc_fence_after();
    }

    tile_epilogue<SPLIT>(P, b, m0, n0, tmem_base, warp, lane);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, tmem_cols<SPLIT>());
}

template <typename T, int SPLIT>
static int launch_gemm_nt(const Launch& L, int batch, cudaStream_t st, const char* what) {
    static DeviceOnce::Flags configured;     // per instantiation and device
    constexpr size_t smem = smem_bytes<SPLIT>();
    if (DeviceOnce::pending(configured)) {
        cudaError_t e = cudaFuncSetAttribute(gemm_nt_kernel<T, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return fail((int)e, "%s: smem opt-in (%zu B): %s", what, smem, cudaGetErrorString(e));
        DeviceOnce::done(configured);
    }
    int tiles = L.tiles0 + (L.nprob > 1 ? (L.p[1].diag_only ? L.p[1].tiles_m : L.p[1].tiles_m * L.p[1].tiles_n) : 0);
    if (tiles <= 0 || batch <= 0) return ROMA_OK;
    gemm_nt_kernel<T, SPLIT><<<dim3(tiles, batch), kThreads, smem, st>>>(L);
    return check_launch(what);
}

// ---------------------------------------------------------------------------------------------- TMA-fed main loop
// 16-bit operands (bf16 / fp16 features as stored: the Gram matrices of the GP) need no conversion, so their
// k-blocks go global -> shared by cp.async.bulk.tensor (3-D maps: K, rows, batch; SWIZZLE_128B lands them directly
// in the UMMA K-major layout; rows / K beyond the tensor are zero-filled by the copy engine).  Warp-specialised:
// lane 0 of warp 0 produces (waits `empty`, arms `full` with the byte count, issues the two box loads), lane 0 of
// warp 1 issues tcgen05.mma and releases the stage with tcgen05.commit -> `empty`; all 8 warps run the epilogue.
struct TmaMaps { CUtensorMap a[2], b[2]; };
constexpr int kTmaStages = 2;                        // 2 x 32 KB: three CTAs per SM (the epilogue, not the main loop, bounds the Gram)
constexpr size_t kTmaSmem = (size_t)kTmaStages * 2 * kPlaneBytes + 1024;

template <typename T>
__global__ void __launch_bounds__(kThreads, 3) gemm_nt_tma_kernel(const Launch L, const __grid_constant__ TmaMaps maps) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[kTmaStages], empty_bar[kTmaStages], done_bar;
    __shared__ uint32_t tmem_slot;

    int t = blockIdx.x;
    const int pi = (L.nprob > 1 && t >= L.tiles0) ? 1 : 0;
    if (pi) t -= L.tiles0;
    const Problem& P = L.p[pi];
    const int tm = P.diag_only ? t : t / P.tiles_n, tn = P.diag_only ? t : t - tm * P.tiles_n;
    if (P.lower_only && tn > tm) return;
    const int b = blockIdx.y;
    const int m0 = tm * BM, n0 = tn * BN;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int nkb = (P.K + BK - 1) / BK;
    constexpr uint32_t kFmt = std::is_same<T, __half>::value ? 0u : 1u;
    constexpr uint32_t idesc = instr_desc(kFmt);

    if (warp == 0) tmem_alloc(&tmem_slot, 128);
    if (tid == 0) {
        for (int s = 0; s < kTmaStages; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
        mbar_init(&done_bar, 1);
        fence_barrier_init();
        tma::prefetch_map(&maps.a[pi]);
        tma::prefetch_map(&maps.b[pi]);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {                                         // ---- producer ----
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % kTmaStages;
                if (kb >= kTmaStages) mbar_wait(&empty_bar[s], (uint32_t)((kb / kTmaStages - 1) & 1));
                unsigned char* sa = smem + (size_t)s * 2 * kPlaneBytes;
                tma::mbar_expect_tx(&full_bar[s], 2u * kPlaneBytes);
                tma::load_3d(sa, &maps.a[pi], &full_bar[s], kb * BK, m0, b);
                tma::load_3d(sa + kPlaneBytes, &maps.b[pi], &full_bar[s], kb * BK, n0, b);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {                                         // ---- MMA issuer ----
            for (int kb = 0; kb < nkb; ++kb) {
                const int s = kb % kTmaStages;
                mbar_wait(&full_bar[s], (uint32_t)((kb / kTmaStages) & 1));
                tc_fence_after();
                const uint32_t a0 = smem_u32(smem + (size_t)s * 2 * kPlaneBytes), b0 = a0 + kPlaneBytes;
#pragma unroll
                for (int kk = 0; kk < BK / 16; ++kk)
                    umma_f16(tmem_base, smem_desc_sw128(a0 + kk * 32), smem_desc_sw128(b0 + kk * 32), idesc, (kb | kk) ? 1u : 0u);
                umma_commit(&empty_bar[s]);                      // stage reusable once these MMAs have read it
            }
            umma_commit(&done_bar);                              // accumulator complete
        }
        __syncwarp();
    }
    mbar_wait(&done_bar, 0);
    tc_fence_after();
    tile_epilogue<1>(P, b, m0, n0, tmem_base, warp, lane);
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 128);
}

// Operand map: (K, rows, batch) with box (64, 128, 1).  Returns false when the operand cannot be described
// (strides not multiples of 16 bytes, misaligned base): the caller then uses the register-staged kernel.
template <typename T>
static bool operand_map(CUtensorMap* map, const void* base, int rows, int K, int ld, long long batch_stride, int batch, int dtype) {
    const uint64_t esz = sizeof(T);
    if ((reinterpret_cast<uintptr_t>(base) & 15u) || ((uint64_t)ld * esz) % 16 || (batch > 1 && ((uint64_t)batch_stride * esz) % 16)) return false;
    const uint64_t dims[3] = {(uint64_t)K, (uint64_t)rows, (uint64_t)batch};
    const uint64_t strides[2] = {(uint64_t)ld * esz, (uint64_t)(batch > 1 ? batch_stride : (long long)rows * ld) * esz};
    const uint32_t box[3] = {(uint32_t)BK, (uint32_t)BM, 1u};
    return tma::encode_fn() && tma::encode_tiled(map, dtype, 3, base, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B) == 0;
}

extern std::atomic<int> g_tc_tma;      // 1 (default): TMA-fed main loop for 16-bit operands; 0: register-staged (testing aid)

template <typename T>
static int launch_gemm_nt_16(const Launch& L, int batch, cudaStream_t st, const char* what, int dtype) {
    static_assert(sizeof(T) == 2, "16-bit operands");
    int tiles = L.tiles0 + (L.nprob > 1 ? (L.p[1].diag_only ? L.p[1].tiles_m : L.p[1].tiles_m * L.p[1].tiles_n) : 0);
    if (tiles <= 0 || batch <= 0) return ROMA_OK;
    TmaMaps maps;
    bool ok = g_tc_tma.load() != 0;
    for (int i = 0; ok && i < L.nprob; ++i) {
        const Problem& P = L.p[i];
        ok = operand_map<T>(&maps.a[i], P.A, P.M, P.K, P.lda, P.sA, batch, dtype) &&
             operand_map<T>(&maps.b[i], P.B, P.N, P.K, P.ldb, P.sB, batch, dtype);
    }
    if (!ok) return launch_gemm_nt<T, 1>(L, batch, st, what);
    if (L.nprob < 2) { maps.a[1] = maps.a[0]; maps.b[1] = maps.b[0]; }
    static DeviceOnce::Flags configured;
    if (DeviceOnce::pending(configured)) {
        cudaError_t e = cudaFuncSetAttribute(gemm_nt_tma_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kTmaSmem);
        if (e != cudaSuccess) return fail((int)e, "%s: smem opt-in (%zu B): %s", what, kTmaSmem, cudaGetErrorString(e));
        DeviceOnce::done(configured);
    }
    gemm_nt_tma_kernel<T><<<dim3(tiles, batch), kThreads, kTmaSmem, st>>>(L, maps);
    return check_launch(what);
}

// ---------------------------------------------------------------------------------------------- split operands from planes
// C = epi(A * B^T) for fp32-grade operands that already exist as bf16 planes (hi, [mid,] lo — written once by the epilogue
// of the kernel that produced them): no conversion in the main loop.  Per 64-wide k-block the planes of both operands land
// in shared memory by TMA (SWIZZLE_128B boxes of 16 KB, K-major) and the plane pairs with i + j < SPLIT are multiplied:
// (hi, hi) into the main accumulator tile, the others into the cross-term tile (TMEM columns 128..255, see gemm_nt_kernel).
// Warp-specialised like gemm_nt_tma_kernel: warp 0 lane 0 produces, warp 1 lane 0 issues tcgen05.mma and releases the
// stage by tcgen05.commit, all 8 warps run the epilogue (EPI_SUB reads and rewrites 64 KB of C; EPI_STORE may emit planes).
struct PlaneMaps { CUtensorMap a[2][3], b[2][3]; };      // [problem][plane]
// One stage = ALL planes of both operands for one 64-wide k-block (2 * SPLIT x 16 KB: 64 / 96 KB), so every plane is fetched
// once per k-block (a first version moved one (A plane, B plane) pair per stage: 6 pairs = 12 plane loads per k-block at 3
// planes, and the update GEMMs became L2-bandwidth-bound at 8 TB/s).  One stage per CTA, two CTAs per SM: while one CTA's
// MMAs run the other's TMA loads land; the epilogue of one overlaps the main loop of the other.
template <int SPLIT> constexpr size_t planes_smem() { return (size_t)2 * SPLIT * kPlaneBytes + 1024; }

// SPLIT = 2: pairs (0,0) | (0,1) (1,0); SPLIT = 3: (0,0) | (0,1) (1,0) (0,2) (1,1) (2,0)  — the first into the main tile
template <int SPLIT>
__global__ void __launch_bounds__(kThreads, 2) gemm_planes_kernel(const Launch L, const __grid_constant__ PlaneMaps maps) {
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar, empty_bar, done_bar;
    __shared__ uint32_t tmem_slot;

    int t = blockIdx.x;
    const int pi = (L.nprob > 1 && t >= L.tiles0) ? 1 : 0;
    if (pi) t -= L.tiles0;
    const Problem& P = L.p[pi];
    const int tm = P.diag_only ? t : t / P.tiles_n, tn = P.diag_only ? t : t - tm * P.tiles_n;
    if (P.lower_only && tn > tm) return;
    const int b = blockIdx.y;
    const int m0 = tm * BM, n0 = tn * BN;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    unsigned char* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
    const int nkb = (P.K + BK - 1) / BK;
    constexpr uint32_t idesc = instr_desc(1u);

    if (warp == 0) tmem_alloc(&tmem_slot, 256);
    if (tid == 0) {
        mbar_init(&full_bar, 1); mbar_init(&empty_bar, 1); mbar_init(&done_bar, 1);
        fence_barrier_init();
#pragma unroll
        for (int q = 0; q < SPLIT; ++q) { tma::prefetch_map(&maps.a[pi][q]); tma::prefetch_map(&maps.b[pi][q]); }
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = tmem_slot;

    if (warp == 0) {
        if (lane == 0) {                                         // ---- producer ----
            for (int kb = 0; kb < nkb; ++kb) {
                if (kb > 0) mbar_wait(&empty_bar, (uint32_t)((kb - 1) & 1));
                tma::mbar_expect_tx(&full_bar, (uint32_t)(2 * SPLIT) * kPlaneBytes);
#pragma unroll
                for (int q = 0; q < SPLIT; ++q) {
                    tma::load_3d(smem + (size_t)q * kPlaneBytes, &maps.a[pi][q], &full_bar, kb * BK, m0, b);
                    tma::load_3d(smem + (size_t)(SPLIT + q) * kPlaneBytes, &maps.b[pi][q], &full_bar, kb * BK, n0, b);
                }
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {                                         // ---- MMA issuer ----
            const uint32_t a0 = smem_u32(smem), b0 = a0 + SPLIT * kPlaneBytes;
            for (int kb = 0; kb < nkb; ++kb) {
                mbar_wait(&full_bar, (uint32_t)(kb & 1));
                tc_fence_after();
#pragma unroll
                for (int kk = 0; kk < BK / 16; ++kk) {
#pragma unroll
                    for (int i = 0; i < SPLIT; ++i) {
#pragma unroll
                        for (int j = 0; j < SPLIT; ++j) {
                            if (i + j >= SPLIT) continue;
                            const uint64_t ad = smem_desc_sw128(a0 + i * kPlaneBytes + kk * 32);
                            const uint64_t bd = smem_desc_sw128(b0 + j * kPlaneBytes + kk * 32);
                            if (i + j == 0) umma_f16(tmem_base, ad, bd, idesc, (kb | kk) ? 1u : 0u);
                            else umma_f16(tmem_base + 128u, ad, bd, idesc, (kb | kk | (i + j - 1) | i) ? 1u : 0u);
                        }
                    }
                }
                umma_commit(&empty_bar);
            }
            umma_commit(&done_bar);
        }
        __syncwarp();
    }
    mbar_wait(&done_bar, 0);
    tc_fence_after();
    tile_epilogue<2>(P, b, m0, n0, tmem_base, warp, lane);      // <2>: main + cross tile
    tc_fence_before();
    __syncthreads();
    if (warp == 0) tmem_dealloc(tmem_base, 256);
}

// *launched = false (nothing launched) when an operand cannot be described to TMA; the caller then takes the register-staged kernel
template <int SPLIT>
static int launch_gemm_planes(const Launch& L, int batch, cudaStream_t st, const char* what, bool* launched) {
    *launched = false;
    int tiles = L.tiles0 + (L.nprob > 1 ? (L.p[1].diag_only ? L.p[1].tiles_m : L.p[1].tiles_m * L.p[1].tiles_n) : 0);
    if (tiles <= 0 || batch <= 0) { *launched = true; return ROMA_OK; }
    PlaneMaps maps;
    bool ok = true;
    for (int i = 0; ok && i < L.nprob; ++i) {
        const Problem& P = L.p[i];
        for (int q = 0; ok && q < SPLIT; ++q)
            ok = P.Apl[q] && P.Bpl[q] &&
                 operand_map<__nv_bfloat16>(&maps.a[i][q], P.Apl[q], P.M, P.K, P.lda, P.sA, batch, ROMA_BF16) &&
                 operand_map<__nv_bfloat16>(&maps.b[i][q], P.Bpl[q], P.N, P.K, P.ldb, P.sB, batch, ROMA_BF16);
        for (int q = SPLIT; ok && q < 3; ++q) { maps.a[i][q] = maps.a[i][0]; maps.b[i][q] = maps.b[i][0]; }
    }
    if (!ok) return ROMA_OK;
    if (L.nprob < 2) for (int q = 0; q < 3; ++q) { maps.a[1][q] = maps.a[0][q]; maps.b[1][q] = maps.b[0][q]; }
    static DeviceOnce::Flags configured;
    if (DeviceOnce::pending(configured)) {
        cudaError_t e = cudaFuncSetAttribute(gemm_planes_kernel<SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)planes_smem<SPLIT>());
        if (e != cudaSuccess) return fail((int)e, "%s: smem opt-in (%zu B): %s", what, planes_smem<SPLIT>(), cudaGetErrorString(e));
        DeviceOnce::done(configured);
    }
    gemm_planes_kernel<SPLIT><<<dim3(tiles, batch), kThreads, planes_smem<SPLIT>(), st>>>(L, maps);
    *launched = true;
    return check_launch(what);
}

static inline Problem make_problem(const void* A, const void* B, float* C, int M, int N, int K, int lda, int ldb, int ldc,
                                   long long sA, long long sB, long long sC, int epi) {
    Problem p{};
    p.A = A; p.B = B; p.C = C; p.Ct = nullptr;
    p.sA = sA; p.sB = sB; p.sC = sC; p.sCt = 0;
    p.M = M; p.N = N; p.K = K; p.lda = lda; p.ldb = ldb; p.ldc = ldc; p.ldct = 0;
    p.tiles_m = (M + BM - 1) / BM; p.tiles_n = (N + BN - 1) / BN;
    p.lower_only = 0; p.diag_only = 0; p.no_diag = 0; p.epi = epi; p.store_m = M; p.store_n = N;
    p.rn = p.cn = nullptr; p.srn = p.scn = 0; p.eps = 0.f; p.inv_T = 0.f; p.diag_add = 0.f;
    p.part = nullptr; p.scale = 1.f; p.w1 = 1; p.lx0 = p.lxs = p.ly0 = p.lys = 0.f;
    p.nplanes = 0;
    for (int i = 0; i < 3; ++i) { p.Pout[i] = nullptr; p.Apl[i] = nullptr; p.Bpl[i] = nullptr; }
    return p;
}

}  // namespace tc
}  // namespace roma
