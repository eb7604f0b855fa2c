// K3b: streaming Gaussian kde over 4-D matches (romatch/utils/kde.py:4-12) + the sampling weights of
// matcher.py:490-491.  No N x N temporary: each thread owns one query point, reference points stream
// through shared memory, partial sums per reference split are combined in a fixed order (deterministic,
// no atomics) by a finalize kernel.
//
// half != 0 reproduces the reference's fp16 rounding chain on CUDA:
//   x.half() -> cdist (ATen _euclidean_dist: |a|^2 + |b|^2 - 2ab via one fp16 matmul over 6-wide augmented
//   vectors, fp32 accumulate, rounded to half; clamp_min(0); sqrt) -> **2 -> neg -> * (1/(2 std^2)) -> exp
//   -> row sum (fp32 accumulate, rounded to half).
// The bound is SFU/FP32 issue (one ex2 per pair), not HBM: 0.4 MB of compulsory traffic at N=40000.
#include "common.cuh"

namespace roma {

constexpr int kKdeTile = 1024;    // reference points per shared-memory tile
constexpr int kKdeThreads = 256;

__device__ __forceinline__ float rh(float v) { return __half2float(__float2half_rn(v)); }   // round to half

template <bool kHalf>
__global__ void __launch_bounds__(kKdeThreads)
kde_partial_kernel(const float* __restrict__ x, int N, int M, int down, float inv_two_var,
                   float* __restrict__ partial, int nsplit) {
    __shared__ float4 ref[kKdeTile];
    __shared__ float refn[kKdeTile];
    const int i = blockIdx.x * kKdeThreads + threadIdx.x;
    const int split = blockIdx.y;
    const int per = (M + nsplit - 1) / nsplit;
    const int j_begin = split * per, j_end = min(M, j_begin + per);

    float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
    float qn = 0.f;
    if (i < N) {
        q = reinterpret_cast<const float4*>(x)[i];
        if (kHalf) {
            q = make_float4(rh(q.x), rh(q.y), rh(q.z), rh(q.w));
            // x.pow(2).sum(-1): squares rounded to half, fp32-accumulated sum rounded to half
            qn = rh(rh(q.x * q.x) + rh(q.y * q.y) + rh(q.z * q.z) + rh(q.w * q.w));
        }
    }
    float acc = 0.f;
    for (int j0 = j_begin; j0 < j_end; j0 += kKdeTile) {
        const int n = min(kKdeTile, j_end - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < n; t += kKdeThreads) {
            float4 r = reinterpret_cast<const float4*>(x)[(long long)(j0 + t) * down];
            if (kHalf) {
                r = make_float4(rh(r.x), rh(r.y), rh(r.z), rh(r.w));
                refn[t] = rh(rh(r.x * r.x) + rh(r.y * r.y) + rh(r.z * r.z) + rh(r.w * r.w));
            }
            ref[t] = r;
        }
        __syncthreads();
        if (i < N) {
#pragma unroll 4
            for (int t = 0; t < n; ++t) {
                const float4 r = ref[t];
                float e;
                if (kHalf) {
                    // [-2a, |a|^2, 1] . [b, 1, |b|^2]: products are exact in fp32, accumulated in fp32
                    float d2 = (-2.f * q.x) * r.x;
                    d2 = fmaf(-2.f * q.y, r.y, d2);
                    d2 = fmaf(-2.f * q.z, r.z, d2);
                    d2 = fmaf(-2.f * q.w, r.w, d2);
                    d2 = d2 + qn;
                    d2 = d2 + refn[t];
                    d2 = rh(d2);
                    const float d = rh(sqrtf(fmaxf(d2, 0.f)));
                    const float s = rh(d * d);                  // cdist(...)**2 in half
                    const float a = rh(-s * inv_two_var);       // / (2*std**2): ATen multiplies by the fp32 reciprocal
                    e = rh(expf(a));
                } else {
                    const float dx = q.x - r.x, dy = q.y - r.y, dz = q.z - r.z, dw = q.w - r.w;
                    const float d2 = fmaf(dx, dx, fmaf(dy, dy, fmaf(dz, dz, dw * dw)));
                    e = expf(-d2 * inv_two_var);
                }
                acc += e;
            }
        }
    }
    if (i < N) partial[(long long)split * N + i] = acc;
}

// ---- half mode, table-driven -----------------------------------------------------------------------
// After d2 = half(|a|^2 + |b|^2 - 2ab) the rest of the reference chain (clamp_min(0), sqrt, **2, neg, /(2 std^2),
// exp, each rounded to half) is a pure function of ONE non-negative half value, i.e. of at most 31745 bit
// patterns, and it is exactly 0 beyond a cutoff (exp underflows in half).  kde_lut_kernel tabulates it with the
// very same op sequence as kde_partial_kernel<true>; the streaming kernel then needs 4 FMA + 2 FADD + one
// F2F + one shared-memory lookup per pair instead of two MUFU and ten conversions.  Results are bit-identical
// to the direct chain (tests/test_kernels_gpu.py::test_kde_lut_equals_direct).
constexpr int kLutMax = 14336;          // table entries that fit next to the reference tile (56 KB)
constexpr int kQPT = 4;                 // queries per thread

__global__ void kde_lut_kernel(float* __restrict__ lut, int* __restrict__ poison, float inv_two_var, int bound) {
    const int bits = blockIdx.x * blockDim.x + threadIdx.x;         // 0 .. 0x7C00 (= +inf)
    if (bits > 0x7C00) return;
    const float d2 = __half2float(__ushort_as_half((unsigned short)bits));
    const float d = rh(sqrtf(fmaxf(d2, 0.f)));
    const float s = rh(d * d);
    const float a = rh(-s * inv_two_var);
    const float e = rh(expf(a));
    lut[bits] = e;
    if (e != 0.f && bits >= bound) atomicExch(poison, 1);          // host-side cutoff estimate violated: fail loudly (NaN)
}

// Packed arithmetic: queries are processed two at a time, their coordinates side by side in 64-bit register
// pairs, so the distance chain (1 mul, 3 fma, 2 add — the order of kde_partial_kernel<true>, hence the same bits)
// is 6 FMUL2/FFMA2/FADD2 for two pairs; the two half indices come out of one F2FP pack, are clamped at 0 / cutoff
// with one 16-bit SIMD max and min each, and address the table twice.  Reference points sit in shared memory
// already duplicated (x,x,y,y | z,z,w,w | n,n) so that a broadcast LDS.128 yields the packed operand.
__global__ void __launch_bounds__(kKdeThreads)
kde_partial_lut_kernel(const float* __restrict__ x, int N, int M, int down, const float* __restrict__ lut_g,
                       int cutoff, float* __restrict__ partial, int nsplit) {
    extern __shared__ float ksm[];
    float* lut = ksm;                                               // [cutoff + 1]; every entry >= cutoff is 0
    float4* ref = reinterpret_cast<float4*>(ksm + ((cutoff + 1 + 3) & ~3));    // [tile][2]: (x,x,y,y), (z,z,w,w)
    float2* refn = reinterpret_cast<float2*>(ref + 2 * kKdeTile);              // [tile]: (n,n)
    for (int t = threadIdx.x; t <= cutoff; t += kKdeThreads) lut[t] = (t < cutoff) ? lut_g[t] : 0.f;

    const int split = blockIdx.y;
    const int per = (M + nsplit - 1) / nsplit;
    const int j_begin = split * per, j_end = min(M, j_begin + per);
    const int i0 = (blockIdx.x * kKdeThreads + threadIdx.x) * kQPT;
    constexpr int kPairs = kQPT / 2;

    float2 qx[kPairs], qy[kPairs], qz[kPairs], qw[kPairs], qn[kPairs];
    float acc[kQPT];
#pragma unroll
    for (int u = 0; u < kQPT; ++u) {
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (i0 + u < N) v = reinterpret_cast<const float4*>(x)[i0 + u];
        v = make_float4(rh(v.x), rh(v.y), rh(v.z), rh(v.w));
        const float n = rh(rh(v.x * v.x) + rh(v.y * v.y) + rh(v.z * v.z) + rh(v.w * v.w));
        if (u & 1) { qx[u / 2].y = -2.f * v.x; qy[u / 2].y = -2.f * v.y; qz[u / 2].y = -2.f * v.z; qw[u / 2].y = -2.f * v.w; qn[u / 2].y = n; }
        else       { qx[u / 2].x = -2.f * v.x; qy[u / 2].x = -2.f * v.y; qz[u / 2].x = -2.f * v.z; qw[u / 2].x = -2.f * v.w; qn[u / 2].x = n; }
        acc[u] = 0.f;
    }
    const unsigned cut2 = (unsigned)cutoff | ((unsigned)cutoff << 16);
    for (int j0 = j_begin; j0 < j_end; j0 += kKdeTile) {
        const int n = min(kKdeTile, j_end - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < n; t += kKdeThreads) {
            float4 r = reinterpret_cast<const float4*>(x)[(long long)(j0 + t) * down];
            r = make_float4(rh(r.x), rh(r.y), rh(r.z), rh(r.w));
            const float rn = rh(rh(r.x * r.x) + rh(r.y * r.y) + rh(r.z * r.z) + rh(r.w * r.w));
            ref[2 * t] = make_float4(r.x, r.x, r.y, r.y);
            ref[2 * t + 1] = make_float4(r.z, r.z, r.w, r.w);
            refn[t] = make_float2(rn, rn);
        }
        __syncthreads();
#pragma unroll 2
        for (int t = 0; t < n; ++t) {
            const float4 rxy = ref[2 * t], rzw = ref[2 * t + 1];
            const float2 rn = refn[t];
            const float2 rx = make_float2(rxy.x, rxy.y), ry = make_float2(rxy.z, rxy.w);
            const float2 rz = make_float2(rzw.x, rzw.y), rw = make_float2(rzw.z, rzw.w);
#pragma unroll
            for (int pr = 0; pr < kPairs; ++pr) {
                float2 d2 = __fmul2_rn(qx[pr], rx);
                d2 = __ffma2_rn(qy[pr], ry, d2);
                d2 = __ffma2_rn(qz[pr], rz, d2);
                d2 = __ffma2_rn(qw[pr], rw, d2);
                d2 = __fadd2_rn(d2, qn[pr]);
                d2 = __fadd2_rn(d2, rn);
                // two half bit patterns; negative (sign bit -> negative int16) clamps to entry 0 = clamp_min(0), large
                // values and +inf clamp to `cutoff` whose entry is 0; NaN never occurs
                const __half2 h2 = __floats2half2_rn(d2.x, d2.y);
                unsigned hb = *reinterpret_cast<const unsigned*>(&h2);
                hb = __vminu2(__vmaxs2(hb, 0u), cut2);
                acc[2 * pr] += lut[hb & 0xffffu];
                acc[2 * pr + 1] += lut[hb >> 16];
            }
        }
    }
#pragma unroll
    for (int u = 0; u < kQPT; ++u)
        if (i0 + u < N) partial[(long long)split * N + i0 + u] = acc[u];
}

// ---- half mode, table-driven, distances on the tensor cores ------------------------------------------------------
// The reference's half cdist IS a matrix product (ATen _euclidean_dist: [-2a, |a|^2, 1] . [b, 1, |b|^2] over 6-wide
// augmented half vectors, fp32 accumulate, rounded to half; on CUDA it runs as a tensor-core HGEMM).  Here the same
// product is one warp-level mma.sync.m16n8k8 (f16 x f16 -> f32, K padded 6 -> 8 with zeros; every product is exact
// in fp32) per 16 queries x 8 reference points: the 6 FMUL2/FFMA2/FADD2 per pair of the SIMT table kernel disappear
// and what is left per pair is the half pack, the two 16-bit SIMD clamps, the table lookup and the fp32 add.
// A warp keeps kKdeQT query tiles (A fragments) in registers and streams the reference points (B fragments, one
// conflict-free LDS.32 per thread per 8 points) against all of them.  Only the fp32 summation order inside the
// product differs from the sequential chain of kde_partial_kernel<true>: a pair can land on the neighbouring half
// value of d2 when the fp32 sum is within an ulp of a rounding boundary (tests/test_kernels_gpu.py bounds the effect
// on the densities: at most one half ulp on a small fraction of the rows).
constexpr int kKdeQT = 8;                // 16-query tiles per warp: 128 queries per warp, 1024 per block

__device__ __forceinline__ void mma_m16n8k8_f16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t b0) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5}, {%6}, {%7, %7, %7, %7};"
                 : "=f"(d[0]), "=f"(d[1]), "=f"(d[2]), "=f"(d[3]) : "r"(a0), "r"(a1), "r"(b0), "f"(0.f));
}
__device__ __forceinline__ uint32_t pack_h2(float lo, float hi) {
    const __half2 h = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&h);
}

__global__ void __launch_bounds__(kKdeThreads)
kde_partial_mma_kernel(const float* __restrict__ x, int N, int M, int down, const float* __restrict__ lut_g,
                       int cutoff, float* __restrict__ partial, int per, int sym, float* __restrict__ colpart) {
    extern __shared__ float ksm[];
    float* lut = ksm;                                               // [cutoff + 1]; every entry >= cutoff is 0
    uint4* ref = reinterpret_cast<uint4*>(ksm + ((cutoff + 1 + 3) & ~3));      // [tile]: halves (x, y, z, w, 1, |r|^2, 0, 0)
    // Symmetric mode (reference points = the queries): k(i, j) = k(j, i) exactly (the tensor core sums the same six exact
    // products), so a warp owning rows [q0, q0 + 128) evaluates only j >= q0: the square on the diagonal as before
    // (row sums only) and, for j >= q0 + 128, every value is added to its row sum AND to a column sum that is written,
    // per warp range and without atomics, to colpart[q0 / 128][j]; kde_finalize adds the column sums of the ranges above.
    const int split = blockIdx.y;
    const int j_begin = split * per, j_end = min(M, j_begin + per);       // per is a multiple of 8
    if (sym && j_end <= (int)(blockIdx.x * (kKdeThreads / 32)) * (16 * kKdeQT)) return;   // whole chunk below the diagonal
    for (int t = threadIdx.x; t <= cutoff; t += kKdeThreads) lut[t] = (t < cutoff) ? lut_g[t] : 0.f;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
    const int q0 = (blockIdx.x * (kKdeThreads / 32) + warp) * (16 * kKdeQT);

    // A fragments: row g (a0) and g + 8 (a1) of each query tile, columns 2 t4, 2 t4 + 1 of (-2x, -2y, -2z, -2w, |q|^2, 1, 0, 0)
    uint32_t a[kKdeQT][2];
    float acc[kKdeQT][2];
#pragma unroll
    for (int qt = 0; qt < kKdeQT; ++qt) {
#pragma unroll
        for (int hr = 0; hr < 2; ++hr) {
            const int q = q0 + qt * 16 + g + 8 * hr;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            float one = 0.f;
            if (q < N) { v = reinterpret_cast<const float4*>(x)[q]; one = 1.f; }
            v = make_float4(rh(v.x), rh(v.y), rh(v.z), rh(v.w));
            float n = rh(rh(v.x * v.x) + rh(v.y * v.y) + rh(v.z * v.z) + rh(v.w * v.w));
            if (q >= N) n = 60000.f;                                  // rows past the end: beyond the cutoff, contribute 0
            const float lo = t4 == 0 ? -2.f * v.x : t4 == 1 ? -2.f * v.z : t4 == 2 ? n : 0.f;
            const float hi = t4 == 0 ? -2.f * v.y : t4 == 1 ? -2.f * v.w : t4 == 2 ? one : 0.f;
            a[qt][hr] = pack_h2(lo, hi);
            acc[qt][hr] = 0.f;
        }
    }
    const unsigned cut2 = (unsigned)cutoff | ((unsigned)cutoff << 16);
    const uint32_t* refw = reinterpret_cast<const uint32_t*>(ref) + g * 4 + t4;     // word t4 of point g of a group of 8
    for (int j0 = j_begin; j0 < j_end; j0 += kKdeTile) {
        const int n = min(kKdeTile, j_end - j0);
        const int groups = (n + 7) >> 3;
        __syncthreads();
        for (int t = threadIdx.x; t < groups * 8; t += kKdeThreads) {
            uint4 pk;
            if (t < n) {
                float4 r = reinterpret_cast<const float4*>(x)[(long long)(j0 + t) * down];
                r = make_float4(rh(r.x), rh(r.y), rh(r.z), rh(r.w));
                const float rn = rh(rh(r.x * r.x) + rh(r.y * r.y) + rh(r.z * r.z) + rh(r.w * r.w));
                pk = make_uint4(pack_h2(r.x, r.y), pack_h2(r.z, r.w), pack_h2(1.f, rn), 0u);
            } else {
                pk = make_uint4(0u, 0u, pack_h2(0.f, 60000.f), 0u);       // tail of the last group: far beyond the cutoff -> contributes 0
            }
            ref[t] = pk;
        }
        __syncthreads();
        // groups of 8 reference points: [0, s_skip) lie below this warp's rows (symmetric mode: left to the column sums of
        // the warp ranges above), [s_skip, s_sym) meet the diagonal square (row sums only), [s_sym, groups) are symmetric
        const int s_skip = sym ? min(max((q0 - j0) >> 3, 0), groups) : 0;
        const int s_sym = sym ? min(max((q0 + 16 * kKdeQT - j0) >> 3, 0), groups) : groups;
#pragma unroll 2
        for (int s = s_skip; s < s_sym; ++s) {
            const uint32_t b0 = refw[s * 32];
#pragma unroll
            for (int qt = 0; qt < kKdeQT; ++qt) {
                float d[4];
                mma_m16n8k8_f16(d, a[qt][0], a[qt][1], b0);
                // half bit patterns of d2; negative -> entry 0 (clamp_min(0)), beyond the table / inf -> `cutoff` (0)
                unsigned u0 = __vminu2(__vmaxs2(pack_h2(d[0], d[1]), 0u), cut2);
                unsigned u1 = __vminu2(__vmaxs2(pack_h2(d[2], d[3]), 0u), cut2);
                acc[qt][0] += lut[u0 & 0xffffu];
                acc[qt][1] += lut[u1 & 0xffffu];
                acc[qt][0] += lut[u0 >> 16];
                acc[qt][1] += lut[u1 >> 16];
            }
        }
        for (int s = s_sym; s < groups; ++s) {
            const uint32_t b0 = refw[s * 32];
            float c0 = 0.f, c1 = 0.f;                                 // columns 2 t4, 2 t4 + 1 of the group, rows g and g + 8 of all tiles
#pragma unroll
            for (int qt = 0; qt < kKdeQT; ++qt) {
                float d[4];
                mma_m16n8k8_f16(d, a[qt][0], a[qt][1], b0);
                unsigned u0 = __vminu2(__vmaxs2(pack_h2(d[0], d[1]), 0u), cut2);
                unsigned u1 = __vminu2(__vmaxs2(pack_h2(d[2], d[3]), 0u), cut2);
                const float v00 = lut[u0 & 0xffffu], v10 = lut[u1 & 0xffffu], v01 = lut[u0 >> 16], v11 = lut[u1 >> 16];
                acc[qt][0] += v00;
                acc[qt][1] += v10;
                acc[qt][0] += v01;
                acc[qt][1] += v11;
                c0 += v00; c0 += v10;
                c1 += v01; c1 += v11;
            }
            // the eight row groups of the warp hold the same two columns: fixed-order butterfly over g
            c0 += __shfl_xor_sync(0xffffffffu, c0, 4);  c1 += __shfl_xor_sync(0xffffffffu, c1, 4);
            c0 += __shfl_xor_sync(0xffffffffu, c0, 8);  c1 += __shfl_xor_sync(0xffffffffu, c1, 8);
            c0 += __shfl_xor_sync(0xffffffffu, c0, 16); c1 += __shfl_xor_sync(0xffffffffu, c1, 16);
            if (g == 0) {
                const int j = j0 + s * 8 + 2 * t4;
                float* cp = colpart + (long long)(q0 / (16 * kKdeQT)) * N + j;
                if (j < N) cp[0] = c0;
                if (j + 1 < N) cp[1] = c1;
            }
        }
    }
    // the four threads of a quad hold the columns of the same query rows: fixed-order butterfly, thread 0 stores
#pragma unroll
    for (int qt = 0; qt < kKdeQT; ++qt) {
#pragma unroll
        for (int hr = 0; hr < 2; ++hr) {
            float v = acc[qt][hr];
            v += __shfl_xor_sync(0xffffffffu, v, 1);
            v += __shfl_xor_sync(0xffffffffu, v, 2);
            const int q = q0 + qt * 16 + g + 8 * hr;
            if (t4 == 0 && q < N) partial[(long long)split * N + q] = v;
        }
    }
}

// sym_per > 0: symmetric tensor-core mode with sym_per reference points per split and qper queries per block: the row
// partial of split k exists only if its chunk reaches the block's rows, and the column sums of the warp ranges above
// row i (colpart[w][i], w < i / 128) complete the density; fixed summation order, no atomics anywhere.
__global__ void kde_finalize_kernel(const float* __restrict__ partial, int N, int nsplit, int half,
                                    float* __restrict__ density, float* __restrict__ p_out, const int* __restrict__ poison,
                                    int sym_per, int qper, const float* __restrict__ colpart) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float s = 0.f;
    if (sym_per > 0) {
        const int qb = (i / qper) * qper;
        for (int k = 0; k < nsplit; ++k)
            if (min(N, (k + 1) * sym_per) > qb) s += partial[(long long)k * N + i];
        const int nw = i / (16 * kKdeQT);
#pragma unroll 8
        for (int w = 0; w < nw; ++w) s += colpart[(long long)w * N + i];      // independent loads, fixed summation order
    } else {
        for (int k = 0; k < nsplit; ++k) s += partial[(long long)k * N + i];
    }
    if (poison && *poison) s = __int_as_float(0x7fc00000);
    if (half) s = rh(s);
    density[i] = s;
    if (p_out) {
        // matcher.py:490-491: p = 1/(density+1); p[density < 10] = 1e-7   (in the dtype of density)
        float p;
        if (half) { p = rh(1.f / rh(s + 1.f)); if (s < 10.f) p = rh(1e-7f); }
        else { p = 1.f / (s + 1.f); if (s < 10.f) p = 1e-7f; }
        p_out[i] = p;
    }
}

}  // namespace roma


namespace roma {
// =====================================================================================================================
// Exact half mode: the density is BIT-IDENTICAL to the reference's CUDA chain
//     (-torch.cdist(x, x)**2 / (2 * std**2)).exp().sum(dim=-1)          romatch/utils/kde.py:4-12, x half
// on this GPU / torch build (pinned on the B200 by tests/test_kde_exact_gpu.py against the unmodified chain):
//   * d2: ATen's _euclidean_dist runs x1_ @ x2_^T (6-wide augmented half vectors) as cuBLAS'
//     cutlass_80_tensorop_f16_s16816gemm (HMMA m16n8k16, fp32 accumulate).  One warp-level mma.sync over the same six
//     products gives the same fp32 value for every pair (67 M pairs compared, 0 differences; the sequential-FMA form of
//     mode 1 differs on 0.05 % of them) — kde_partial_mma_kernel already used it.
//   * clamp_min(0), sqrt, **2, neg, * (1 / (2 std^2)), exp: elementwise half ops = the table (0 differences over all
//     31745 non-negative half values).
//   * sum(dim=-1): ATen's reduce kernel (ATen/native/cuda/Reduce.cuh, sum_functor<Half, float, Half>) accumulates in
//     fp32 in a FIXED order that depends only on the row length M and the row's alignment: block (bw, bh) = (32, 16);
//     for M >= 128 every thread keeps 8 accumulators over aligned 16-byte vectors (input_vec_size 8 for 16-bit types),
//     vector index idx = tl, tl + T, ... with T = 512 cooperating threads per row when ceil(M / 32) >= 256 (else T = 32,
//     one warp per row), unaligned head and tail elements go to accumulator 0 of the threads tx < 8, the 8 accumulators
//     are combined left to right, then a shuffle-down tree over tx (16, 8, 4, 2, 1) and a shared-memory tree over ty
//     (8, 4, 2, 1); the fp32 result is rounded to half once.  fp32 addition is not associative, so the kernels below
//     add the same terms in the same order:
//       - kde_exact_fused_kernel (M % 8 == 0, T = 512: the sizes sample() produces, N = 4 * num): accumulators
//         (row, tl, c) live in registers — a column group of 8 of the MMA tile IS one aligned vector — and the
//         512 per-thread values of a row go to the workspace; kde_exact_tree_kernel applies the two trees.
//       - any other shape: kde_scores_kernel writes the half scores of a chunk of rows to the workspace and
//         kde_aten_sum_kernel is a literal restatement of ATen's thread_reduce / block_x_reduce / block_y_reduce.
// =====================================================================================================================
struct AtenSumCfg { int vectorized, bw, bh, split; };

// points -> the two 8-wide half operand rows of the GEMM (zero padded): q = x1_ = (-2a, |a|^2, 1), r = x2_ = (a, 1, |a|^2)
__global__ void kde_pack_kernel(const float* __restrict__ x, int N, int M, int down, uint4* __restrict__ qpack, uint4* __restrict__ rpack) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < N) {
        float4 v = reinterpret_cast<const float4*>(x)[i];
        v = make_float4(rh(v.x), rh(v.y), rh(v.z), rh(v.w));
        const float n = rh(rh(v.x * v.x) + rh(v.y * v.y) + rh(v.z * v.z) + rh(v.w * v.w));
        qpack[i] = make_uint4(pack_h2(-2.f * v.x, -2.f * v.y), pack_h2(-2.f * v.z, -2.f * v.w), pack_h2(n, 1.f), 0u);
    }
    if (i < M) {
        float4 v = reinterpret_cast<const float4*>(x)[(long long)i * down];
        v = make_float4(rh(v.x), rh(v.y), rh(v.z), rh(v.w));
        const float n = rh(rh(v.x * v.x) + rh(v.y * v.y) + rh(v.z * v.z) + rh(v.w * v.w));
        rpack[i] = make_uint4(pack_h2(v.x, v.y), pack_h2(v.z, v.w), pack_h2(1.f, n), 0u);
    }
}

// ---- fused: T = 512 threads per row, rows 16-byte aligned (M % 8 == 0) ----------------------------------------------
// Work item = (block of 8 x 16 rows, set of kFusedTL consecutive tl values): warp w owns rows [rb*128 + 16 w, +16) and, for
// every k, the column groups G = 512 k + tl0 + u (u < kFusedTL) — i.e. all terms of the accumulators (row, tl0 + u, c).
// Per group one m16n8k8 MMA (A = the warp's query rows, held in registers; B = 8 reference points, one coalesced
// 4-byte load per lane from the packed operands: the 8 warps of the block read the same line, L1 hits), two packs, the
// clamps, four table lookups and four fp32 adds in k order.  No shared-memory tile, no barrier in the main loop.
constexpr int kFusedTL = 16;            // tl values per work item: 64 accumulators per thread

__global__ void __launch_bounds__(kKdeThreads, 2)
kde_exact_fused_kernel(const uint4* __restrict__ qpack, const uint4* __restrict__ rpack, int N, int M,
                       const float* __restrict__ lut_g, int cutoff, float* __restrict__ V, int nrb, int nitems) {
    extern __shared__ float ksm[];
    float* lut = ksm;
    for (int t = threadIdx.x; t <= cutoff; t += kKdeThreads) lut[t] = (t < cutoff) ? lut_g[t] : 0.f;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
    const unsigned cut2 = (unsigned)cutoff | ((unsigned)cutoff << 16);
    const int ngroups = M >> 3;                                   // column groups (aligned vectors) per row
    const uint32_t* rw = reinterpret_cast<const uint32_t*>(rpack) + g * 4 + t4;   // word t4 of point g of a group
    for (int item = blockIdx.x; item < nitems; item += gridDim.x) {
        const int rb = item % nrb, ts = item / nrb;               // consecutive blocks share the reference columns
        const int row0 = rb * 128 + warp * 16;
        if (row0 >= N) continue;
        const int tl0 = ts * kFusedTL;
        uint32_t a0 = 0u, a1 = 0u;
        {
            const int q0 = row0 + g, q1 = row0 + g + 8;
            const uint32_t far = t4 == 2 ? pack_h2(60000.f, 0.f) : 0u;          // rows past the end: beyond the cutoff
            a0 = q0 < N ? reinterpret_cast<const uint32_t*>(qpack + q0)[t4] : far;
            a1 = q1 < N ? reinterpret_cast<const uint32_t*>(qpack + q1)[t4] : far;
        }
        float acc[kFusedTL][4];
#pragma unroll
        for (int u = 0; u < kFusedTL; ++u) { acc[u][0] = acc[u][1] = acc[u][2] = acc[u][3] = 0.f; }
        for (int G0 = tl0; G0 < ngroups; G0 += 512) {             // k = G0 / 512: ascending = ATen's loop order
            if (G0 + kFusedTL <= ngroups) {
                // all kFusedTL groups exist (every iteration but possibly the last): straight-line code, so the 16 MMAs,
                // packs, clamps, table loads and adds of one iteration interleave freely (a branch per group — the tail form
                // below — exposes the HMMA and LDS latency once per group: measured 0.80 -> see DESIGN.md at N = 40000)
                uint32_t b[kFusedTL];
#pragma unroll
                for (int u = 0; u < kFusedTL; ++u) b[u] = __ldg(rw + (long long)(G0 + u) * 32);
#pragma unroll
                for (int u0 = 0; u0 < kFusedTL; u0 += 4) {
                    float d[4][4];
#pragma unroll
                    for (int v = 0; v < 4; ++v) mma_m16n8k8_f16(d[v], a0, a1, b[u0 + v]);
                    unsigned ix[4][2];
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        // clamp to [0, cutoff] (half bit patterns of non-negative values order like integers), then x 4 in
                        // the packed domain: 4 * cutoff < 2^16, so each half becomes the BYTE offset of its table entry
                        ix[v][0] = __vminu2(__vmaxs2(pack_h2(d[v][0], d[v][1]), 0u), cut2) << 2;
                        ix[v][1] = __vminu2(__vmaxs2(pack_h2(d[v][2], d[v][3]), 0u), cut2) << 2;
                    }
                    const char* lutb = reinterpret_cast<const char*>(lut);
                    float e[4][4];
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        e[v][0] = *reinterpret_cast<const float*>(lutb + (ix[v][0] & 0xffffu));
                        e[v][1] = *reinterpret_cast<const float*>(lutb + (ix[v][0] >> 16));
                        e[v][2] = *reinterpret_cast<const float*>(lutb + (ix[v][1] & 0xffffu));
                        e[v][3] = *reinterpret_cast<const float*>(lutb + (ix[v][1] >> 16));
                    }
#pragma unroll
                    for (int v = 0; v < 4; ++v) {
                        acc[u0 + v][0] += e[v][0];               // (row g,     c = 2 t4)
                        acc[u0 + v][1] += e[v][1];               // (row g,     c = 2 t4 + 1)
                        acc[u0 + v][2] += e[v][2];               // (row g + 8, c = 2 t4)
                        acc[u0 + v][3] += e[v][3];               // (row g + 8, c = 2 t4 + 1)
                    }
                }
                continue;
            }
            uint32_t b[kFusedTL];
#pragma unroll
            for (int u = 0; u < kFusedTL; ++u) b[u] = (G0 + u < ngroups) ? __ldg(rw + (long long)(G0 + u) * 32) : 0xffffffffu;
#pragma unroll
            for (int u = 0; u < kFusedTL; ++u) {
                if (G0 + u < ngroups) {                               // warp-uniform
                    float d[4];
                    mma_m16n8k8_f16(d, a0, a1, b[u]);
                    const unsigned u0 = __vminu2(__vmaxs2(pack_h2(d[0], d[1]), 0u), cut2);
                    const unsigned u1 = __vminu2(__vmaxs2(pack_h2(d[2], d[3]), 0u), cut2);
                    acc[u][0] += lut[u0 & 0xffffu];               // (row g,     c = 2 t4)
                    acc[u][1] += lut[u0 >> 16];                   // (row g,     c = 2 t4 + 1)
                    acc[u][2] += lut[u1 & 0xffffu];               // (row g + 8, c = 2 t4)
                    acc[u][3] += lut[u1 >> 16];                   // (row g + 8, c = 2 t4 + 1)
                }
            }
        }
        // combine the 8 accumulators of (row, tl) left to right: ((((((v0 + v1) + v2) + v3) + v4) + v5) + v6) + v7,
        // v(2 t4), v(2 t4 + 1) held by lane t4 of the quad: the running sum walks through the quad
#pragma unroll
        for (int u = 0; u < kFusedTL; ++u) {
#pragma unroll
            for (int hr = 0; hr < 2; ++hr) {
                const float e0 = acc[u][2 * hr], e1 = acc[u][2 * hr + 1];
                float s = e0 + e1;                                 // lane t4 = 0: v0 + v1
#pragma unroll
                for (int step = 1; step < 4; ++step) {
                    const float prev = __shfl_up_sync(0xffffffffu, s, 1);
                    if (t4 == step) s = (prev + e0) + e1;
                }
                const int q = row0 + g + 8 * hr;
                if (t4 == 3 && q < N && tl0 + u < 512) V[(long long)q * 512 + tl0 + u] = s;
            }
        }
    }
}

// the two trees of ATen's reduce over the 512 per-thread values of a row (one warp per row: lane = tx, v[ty])
__global__ void __launch_bounds__(256)
kde_exact_tree_kernel(const float* __restrict__ V, int N, float* __restrict__ density, float* __restrict__ p_out,
                      const int* __restrict__ poison) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (row >= N) return;
    float v[16];
#pragma unroll
    for (int ty = 0; ty < 16; ++ty) v[ty] = V[(long long)row * 512 + ty * 32 + lane];
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) {
#pragma unroll
        for (int ty = 0; ty < 16; ++ty) v[ty] = v[ty] + __shfl_down_sync(0xffffffffu, v[ty], off);
    }
#pragma unroll
    for (int off = 8; off > 0; off >>= 1) {
#pragma unroll
        for (int ty = 0; ty < 8; ++ty) if (ty < off) v[ty] = v[ty] + v[ty + off];
    }
    if (lane != 0) return;
    float s = v[0];
    if (poison && *poison) s = __int_as_float(0x7fc00000);
    s = rh(s);
    density[row] = s;
    if (p_out) { float p = rh(1.f / rh(s + 1.f)); if (s < 10.f) p = rh(1e-7f); p_out[row] = p; }
}

// ---- generic: scores of a chunk of rows to the workspace, then ATen's reduction restated literally ------------------
// E[(r - r0) * pitch + j] = half bits of exp(...) for rows [r0, r0 + rows) x all M columns; 16-bit table in shared memory
__global__ void __launch_bounds__(kKdeThreads)
kde_scores_kernel(const uint4* __restrict__ qpack, const uint4* __restrict__ rpack, int N, int M, int r0, int rows,
                  const float* __restrict__ lut_g, int cutoff, unsigned short* __restrict__ E, int pitch, int gper) {
    extern __shared__ unsigned short lut16[];                                // table as half bits (28 KB at std = 0.1)
    for (int t = threadIdx.x; t <= cutoff; t += kKdeThreads)
        lut16[t] = (t < cutoff) ? __half_as_ushort(__float2half_rn(lut_g[t])) : (unsigned short)0;
    __syncthreads();
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, t4 = lane & 3;
    const int lr0 = (blockIdx.x * (kKdeThreads / 32) + warp) * 16;          // first local row of this warp
    if (lr0 >= rows) return;
    const unsigned cut2 = (unsigned)cutoff | ((unsigned)cutoff << 16);
    const int q0 = r0 + lr0 + g, q1 = q0 + 8;
    const bool ok0 = lr0 + g < rows && q0 < N, ok1 = lr0 + g + 8 < rows && q1 < N;
    const uint32_t a0 = ok0 ? reinterpret_cast<const uint32_t*>(qpack + q0)[t4] : 0u;
    const uint32_t a1 = ok1 ? reinterpret_cast<const uint32_t*>(qpack + q1)[t4] : 0u;
    const int ngroups = (M + 7) >> 3;
    const int G_begin = blockIdx.y * gper, G_end = min(ngroups, G_begin + gper);
    for (int G = G_begin; G < G_end; ++G) {
        const int jr = G * 8 + g;                                            // reference point of this lane's B row
        const uint32_t b0 = jr < M ? __ldg(reinterpret_cast<const uint32_t*>(rpack + jr) + t4) : 0u;
        float d[4];
        mma_m16n8k8_f16(d, a0, a1, b0);
        const unsigned u0 = __vminu2(__vmaxs2(pack_h2(d[0], d[1]), 0u), cut2);
        const unsigned u1 = __vminu2(__vmaxs2(pack_h2(d[2], d[3]), 0u), cut2);
        const int j = G * 8 + 2 * t4;
        if (ok0) {
            unsigned short* e = E + (long long)(lr0 + g) * pitch + j;
            if (j < M) e[0] = lut16[u0 & 0xffffu];
            if (j + 1 < M) e[1] = lut16[u0 >> 16];
        }
        if (ok1) {
            unsigned short* e = E + (long long)(lr0 + g + 8) * pitch + j;
            if (j < M) e[0] = lut16[u1 & 0xffffu];
            if (j + 1 < M) e[1] = lut16[u1 >> 16];
        }
    }
}

// ATen's ReduceOp::run for a contiguous inner-dimension sum of halfs with fp32 accumulation (Reduce.cuh: thread_reduce
// with input_vectorized_thread_reduce_impl / thread_reduce_impl, block_x_reduce, block_y_reduce), same block shape,
// same per-thread element order; `rowM` = the row length in the REFERENCE's matrix (decides the alignment shift of row
// r: (r * M) % 8 elements), E rows here are pitch-strided and always aligned (the shift is applied arithmetically).
__global__ void kde_aten_sum_kernel(const unsigned short* __restrict__ E, int pitch, int r0, int rows, int N, int M,
                                    AtenSumCfg cfg, float* __restrict__ density, float* __restrict__ p_out,
                                    const int* __restrict__ poison) {
    extern __shared__ float red[];
    const int tx = threadIdx.x, ty = threadIdx.y, bw = blockDim.x, bh = blockDim.y;
    const int lrow = cfg.split ? (int)blockIdx.x : (int)blockIdx.x * bh + ty;
    const bool live = lrow < rows && r0 + lrow < N;
    const int T = cfg.split ? bw * bh : bw;                       // threads cooperating on one row
    const int tl = cfg.split ? tx + ty * bw : tx;                 // config.input_idx()
    const bool tail_thread = !cfg.split || ty == 0;               // should_reduce_tail()
    const unsigned short* data = E + (long long)lrow * pitch;
    auto at = [&](int j) { return __half2float(__ushort_as_half(data[j])); };
    float value = 0.f;
    if (live) {
        if (cfg.vectorized) {
            int end = M, base = 0;                                // data + base is the (virtually) aligned start
            float head = 0.f;
            const int shift0 = (int)(((long long)(r0 + lrow) * M) & 7);
            if (shift0 > 0) {
                // data -= shift; end += shift; threads shift <= tx < 8 take element tx - shift; then skip one vector
                if (tx >= shift0 && tx < 8 && tail_thread) head = head + at(tx - shift0);
                end = M + shift0 - 8;
                base = 8 - shift0;
            }
            float vl[8];
            vl[0] = head;
#pragma unroll
            for (int c = 1; c < 8; ++c) vl[c] = 0.f;
            int idx = tl;
            while (idx * 8 + 7 < end) {
#pragma unroll
                for (int c = 0; c < 8; ++c) vl[c] = vl[c] + at(base + idx * 8 + c);
                idx += T;
            }
            const int tail_start = end - (end & 7);               // end may be negative only when M < 8: not vectorized then
            if (tail_thread) {
                const int j = tail_start + tx;
                if (j < end) vl[0] = vl[0] + at(base + j);
            }
#pragma unroll
            for (int c = 1; c < 8; ++c) vl[0] = vl[0] + vl[c];
            value = vl[0];
        } else {
            // thread_reduce_impl, vt0 = 4: four accumulators over idx, idx + T, idx + 2T, idx + 3T
            float vl[4] = {0.f, 0.f, 0.f, 0.f};
            int idx = tl;
            while (idx + 3 * T < M) {
#pragma unroll
                for (int i = 0; i < 4; ++i) vl[i] = vl[i] + at(idx + i * T);
                idx += 4 * T;
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                if (idx >= M) break;
                vl[i] = vl[i] + at(idx);
                idx += T;
            }
            value = ((vl[0] + vl[1]) + vl[2]) + vl[3];
        }
    }
    // block_x_reduce (always: the reduction runs along x)
    int dim_x = bw;
    if (dim_x > 32) {
        const int address_base = tx + ty * bw;
        red[address_base] = value;
        for (int offset = dim_x / 2; offset >= 32; offset >>= 1) {
            __syncthreads();
            if (tx < offset && tx + offset < bw) {
                value = value + red[address_base + offset];
                red[address_base] = value;
            }
        }
        dim_x = 32;
    }
    __syncthreads();
    for (int offset = dim_x >> 1; offset > 0; offset >>= 1) value = value + __shfl_down_sync(0xffffffffu, value, offset);
    if (cfg.split) {                                              // block_y_reduce
        red[tx + ty * bw] = value;
        for (int offset = bh / 2; offset > 0; offset >>= 1) {
            __syncthreads();
            if (ty < offset && ty + offset < bh) {
                value = value + red[tx + (ty + offset) * bw];
                red[tx + ty * bw] = value;
            }
        }
    }
    if (live && tx == 0 && (!cfg.split || ty == 0)) {
        float s = value;
        if (poison && *poison) s = __int_as_float(0x7fc00000);
        s = rh(s);
        density[r0 + lrow] = s;
        if (p_out) { float p = rh(1.f / rh(s + 1.f)); if (s < 10.f) p = rh(1e-7f); p_out[r0 + lrow] = p; }
    }
}

}  // namespace roma

namespace roma {
static int last_pow2(int n) {
    n |= n >> 1; n |= n >> 2; n |= n >> 4; n |= n >> 8; n |= n >> 16;
    const int r = n - (n >> 1);
    return r < 1 ? 1 : r;
}
// setReduceConfig (Reduce.cuh) for sum(dim=-1) of a contiguous (rows, M) half matrix, input_vec_size = 8, 512 threads max
static AtenSumCfg aten_sum_config(int rows, int M) {
    AtenSumCfg c;
    c.vectorized = M >= 128 ? 1 : 0;
    const int dim0 = c.vectorized ? M / 8 : M, dim1 = rows;
    const int d0 = dim0 < 512 ? last_pow2(dim0) : 512, d1 = dim1 < 512 ? last_pow2(dim1) : 512;
    c.bw = d0 < 32 ? d0 : 32;
    c.bh = d1 < 512 / c.bw ? d1 : 512 / c.bw;
    c.bw = d0 < 512 / c.bh ? d0 : 512 / c.bh;
    const int vpt = (M + c.bw - 1) / c.bw;
    const int thr = c.bh * 16 < 256 ? c.bh * 16 : 256;
    c.split = vpt >= thr ? 1 : 0;
    return c;
}
}  // namespace roma

namespace roma {
static int kde_nsplit(int N) {
    // enough CTAs for >= 2 waves of 148 SMs x 4 resident CTAs
    const int qblocks = (N + kKdeThreads - 1) / kKdeThreads;
    int ns = (2 * 4 * sm_count() + qblocks - 1) / qblocks;
    if (ns < 1) ns = 1;
    if (ns > 64) ns = 64;
    return ns;
}
}  // namespace roma

namespace roma {
constexpr size_t kKdeLutBytes = 32768 * sizeof(float) + 256;      // table + cutoff word, at the END of the workspace
}

namespace roma {
constexpr int kKdeSymMaxN = 65536;       // symmetric mode keeps ceil(N / 128) x N column partials: 134 MB at the limit
static size_t kde_colpart_bytes(int N) {
    return N <= kKdeSymMaxN ? (size_t)((N + 16 * kKdeQT - 1) / (16 * kKdeQT)) * (size_t)N * sizeof(float) : 0;
}
}  // namespace roma

namespace roma {
static size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
// workspace of the exact mode: [qpack N x 16][rpack M x 16][table 32768 f + flag][main: V (N x 512 f) or a chunk of scores]
static size_t kde_exact_fixed_bytes(int N) { return align256((size_t)N * 32) + align256(kKdeLutBytes); }
}  // namespace roma

extern "C" size_t roma_kde_workspace_bytes(int N) {
    if (N <= 0) return 0;
    const size_t legacy = (size_t)64 * (size_t)N * sizeof(float) + roma::kde_colpart_bytes(N) + roma::kKdeLutBytes;
    const size_t exact = roma::kde_exact_fixed_bytes(N) + (size_t)N * 512 * sizeof(float) + 4096;   // V, or >= 1024 rows of scores
    return legacy > exact ? legacy : exact;
}

// 0 (default) = exact: bit-identical to the reference's CUDA op chain (tensor-core distances, table, ATen's summation
// order).  Testing / opt-in: 1 = the op chain per pair with sequential fp32 FMAs, 2 = table + sequential FMAs, 3 = tensor-core
// distances + table, every (i, j) and (j, i), free summation order, 4 = as 3 with every pair evaluated once (fastest; the
// density may differ from the reference's in the last half bit on ~0.03 % of the rows)
namespace roma { static int g_kde_direct = 0; }
extern "C" int roma_kde_set_direct(int direct) { roma::g_kde_direct = (direct >= 1 && direct <= 4) ? direct : 0; return ROMA_OK; }

namespace roma {
static int kde_exact(const float* x, int N, int M, int down, float inv_two_var, float* density, float* p_out,
                     void* workspace, size_t workspace_bytes, cudaStream_t st) {
    const size_t fixed = kde_exact_fixed_bytes(N);
    const int pitch = (M + 7) & ~7;
    ROMA_REQUIRE(workspace_bytes >= fixed + (size_t)16 * pitch * 2, ROMA_ERR_WORKSPACE,
                 "roma_kde: workspace too small for the exact half mode (%zu bytes, see roma_kde_workspace_bytes)", workspace_bytes);
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 15u) == 0, ROMA_ERR_ALIGNMENT, "roma_kde: workspace must be 16-byte aligned");
    char* base = reinterpret_cast<char*>(workspace);
    uint4* qpack = reinterpret_cast<uint4*>(base);
    uint4* rpack = qpack + N;
    float* lut = reinterpret_cast<float*>(base + align256((size_t)N * 32));
    int* poison = reinterpret_cast<int*>(lut + 32768);
    char* mainbuf = base + fixed;
    const size_t main_bytes = workspace_bytes - fixed;
    // table: entries >= cutoff are exactly 0 (exp underflows in half); host-side estimate, verified on the device (poison)
    int cutoff = 0x7C00;
    {
        const float s_max = 1.125f * 17.33f / inv_two_var;
        if (s_max < 60000.f) {
            const int cb = (int)__half_as_ushort(__float2half_rn(s_max)) + 2;
            if (cb < cutoff) cutoff = cb;
        }
    }
    int rc;
    cudaMemsetAsync(poison, 0, sizeof(int), st);
    kde_lut_kernel<<<(0x7C00 + 256) / 256, 256, 0, st>>>(lut, poison, inv_two_var, cutoff < 0x7C00 ? cutoff : 0x7C01);
    kde_pack_kernel<<<cdiv(N, 256), 256, 0, st>>>(x, N, M, down, qpack, rpack);
    if ((rc = check_launch("roma_kde(exact: table, pack)", 2))) return rc;
    const AtenSumCfg cfg = aten_sum_config(N, M);
    int dev = 0;
    cudaGetDevice(&dev);
    static bool cfg_done[64] = {false};
    if (dev >= 0 && dev < 64 && !cfg_done[dev]) {
        cudaError_t e = cudaFuncSetAttribute(kde_exact_fused_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)((kLutMax + 4) * sizeof(float)));
        if (e == cudaSuccess) e = cudaFuncSetAttribute(kde_scores_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(0x7C04 * sizeof(unsigned short)));
        if (e != cudaSuccess) return fail((int)e, "roma_kde: smem opt-in: %s", cudaGetErrorString(e));
        cfg_done[dev] = true;
    }
    const bool fused = cfg.vectorized && cfg.split && cfg.bw == 32 && cfg.bh == 16 && (M & 7) == 0 && cutoff <= kLutMax &&
                       main_bytes >= (size_t)N * 512 * sizeof(float);
    if (fused) {
        float* V = reinterpret_cast<float*>(mainbuf);
        const int nrb = cdiv(N, 128), nitems = nrb * (512 / kFusedTL);
        int grid = 2 * sm_count();
        if (grid > nitems) grid = nitems;
        const size_t smem = ((size_t)(cutoff + 1 + 3) & ~(size_t)3) * sizeof(float);
        kde_exact_fused_kernel<<<grid, kKdeThreads, smem, st>>>(qpack, rpack, N, M, lut, cutoff, V, nrb, nitems);
        kde_exact_tree_kernel<<<cdiv(N, 8), 256, 0, st>>>(V, N, density, p_out, poison);
        return check_launch("roma_kde(exact, fused)", 2);
    }
    // generic: chunks of rows through the workspace
    long long chunk = (long long)(main_bytes / ((size_t)pitch * 2));
    chunk = chunk > N ? N : chunk;
    if (chunk < N) chunk &= ~15ll;
    ROMA_REQUIRE(chunk >= 1, ROMA_ERR_WORKSPACE, "roma_kde: workspace too small for one chunk of scores");
    unsigned short* E = reinterpret_cast<unsigned short*>(mainbuf);
    const int ngroups = (M + 7) >> 3;
    const size_t smem_scores = (size_t)(cutoff + 2) * sizeof(unsigned short);
    for (long long r0 = 0; r0 < N; r0 += chunk) {
        const int rows = (int)((N - r0) < chunk ? (N - r0) : chunk);
        const int rblocks = cdiv(cdiv(rows, 16), kKdeThreads / 32);
        int csplit = cdiv(4 * sm_count(), rblocks);
        if (csplit > ngroups) csplit = ngroups;
        if (csplit < 1) csplit = 1;
        const int gper = cdiv(ngroups, csplit);
        csplit = cdiv(ngroups, gper);
        kde_scores_kernel<<<dim3(rblocks, csplit), kKdeThreads, smem_scores, st>>>(qpack, rpack, N, M, (int)r0, rows, lut, cutoff, E, pitch, gper);
        const int sblocks = cfg.split ? rows : cdiv(rows, cfg.bh);
        kde_aten_sum_kernel<<<sblocks, dim3(cfg.bw, cfg.bh), (size_t)cfg.bw * cfg.bh * sizeof(float), st>>>(
            E, pitch, (int)r0, rows, N, M, cfg, density, p_out, poison);
        if ((rc = check_launch("roma_kde(exact, generic)", 2))) return rc;
    }
    return ROMA_OK;
}
}  // namespace roma

extern "C" int roma_kde(const float* x, int N, int down, double std, int half,
                           float* density, float* p_out, void* workspace, size_t workspace_bytes, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && density && workspace, ROMA_ERR_BAD_ARG, "roma_kde: null pointer");
    ROMA_REQUIRE(N > 0 && down >= 1 && std > 0.0, ROMA_ERR_BAD_ARG, "roma_kde: bad arguments N=%d down=%d std=%g", N, down, std);
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15u) == 0, ROMA_ERR_ALIGNMENT, "roma_kde: x must be 16-byte aligned");
    const int M = (N + down - 1) / down;
    int nsplit = kde_nsplit(N);
    const int max_split = (int)(workspace_bytes / ((size_t)N * sizeof(float)));
    ROMA_REQUIRE(max_split >= 1, ROMA_ERR_WORKSPACE, "roma_kde: workspace too small");
    if (nsplit > max_split) nsplit = max_split;
    if (nsplit > (M + kKdeTile - 1) / kKdeTile) nsplit = (M + kKdeTile - 1) / kKdeTile;
    // 2*std**2 evaluated in double like the Python scalar, reciprocal taken in fp32 like ATen's div-by-scalar
    const float two_var = (float)(2.0 * std * std);
    const float inv_two_var = 1.0f / two_var;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (half && g_kde_direct == 0) return kde_exact(x, N, M, down, inv_two_var, density, p_out, workspace, workspace_bytes, st);
    const int legacy = g_kde_direct;                    // 1, 2, 3, 4 (half) — or anything for fp32
    float* partial = reinterpret_cast<float*>(workspace);
    dim3 grid((N + kKdeThreads - 1) / kKdeThreads, nsplit);
    int rc;
    int* poison = nullptr;
    int sym_per = 0, sym_qper = 0;
    float* colpart = nullptr;
    // table-driven half path: the cutoff is a property of std alone; evaluate it on the host with the same
    // half arithmetic (exp(-s/(2 std^2)) < 2^-25 rounds to 0) so no device round trip is needed to size smem.
    bool use_lut = half && legacy != 1 && workspace_bytes >= (size_t)N * sizeof(float) + kKdeLutBytes;
    int cutoff_bound = 0;
    if (use_lut) {
        // s beyond which exp underflows to 0 in half: exp(-s*inv) < 2^-25  <=>  s > 25 ln2 / inv ; +12.5% margin
        const float s_max = 1.125f * 17.33f / inv_two_var;
        __half hs = __float2half_rn(s_max);
        cutoff_bound = (int)__half_as_ushort(hs) + 2;
        use_lut = s_max < 60000.f && cutoff_bound <= kLutMax;
    }
    if (use_lut) {
        float* lut = reinterpret_cast<float*>(reinterpret_cast<char*>(workspace) + workspace_bytes - kKdeLutBytes);
        poison = reinterpret_cast<int*>(lut + 32768);
        if (nsplit > (int)((workspace_bytes - kKdeLutBytes) / ((size_t)N * sizeof(float)))) nsplit = (int)((workspace_bytes - kKdeLutBytes) / ((size_t)N * sizeof(float)));
        cudaMemsetAsync(poison, 0, sizeof(int), st);
        kde_lut_kernel<<<(0x7C00 + 256) / 256, 256, 0, st>>>(lut, poison, inv_two_var, cutoff_bound);
        const size_t smem = ((size_t)(cutoff_bound + 1 + 3) & ~(size_t)3) * sizeof(float) + kKdeTile * (2 * sizeof(float4) + sizeof(float2));
        static DeviceOnce::Flags cfg;
        if (DeviceOnce::pending(cfg)) {
            cudaError_t e = cudaFuncSetAttribute(kde_partial_lut_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(((kLutMax + 4) * sizeof(float)) + kKdeTile * 40));
            if (e != cudaSuccess) return fail((int)e, "roma_kde: smem opt-in: %s", cudaGetErrorString(e));
            DeviceOnce::done(cfg);
        }
        if (legacy == 2) {
            const int qblocks = (N + kKdeThreads * kQPT - 1) / (kKdeThreads * kQPT);
            int ns = (2 * 3 * sm_count() + qblocks - 1) / qblocks;
            ns = ns < 1 ? 1 : (ns > nsplit ? nsplit : ns);
            if (ns > (M + 255) / 256) ns = (M + 255) / 256;
            nsplit = ns;
            kde_partial_lut_kernel<<<dim3(qblocks, nsplit), kKdeThreads, smem, st>>>(x, N, M, down, lut, cutoff_bound, partial, nsplit);
        } else {
            static DeviceOnce::Flags cfg_mma;
            if (DeviceOnce::pending(cfg_mma)) {
                cudaError_t e = cudaFuncSetAttribute(kde_partial_mma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(((kLutMax + 4) * sizeof(float)) + kKdeTile * 16));
                if (e != cudaSuccess) return fail((int)e, "roma_kde: smem opt-in: %s", cudaGetErrorString(e));
                DeviceOnce::done(cfg_mma);
            }
            const size_t smem_mma = ((size_t)(cutoff_bound + 1 + 3) & ~(size_t)3) * sizeof(float) + kKdeTile * sizeof(uint4);
            const int qper = (kKdeThreads / 32) * 16 * kKdeQT;                  // queries per block
            const int qblocks = (N + qper - 1) / qper;
            // symmetric mode: half of the (block, split) pairs exit at once, so twice the splits keep the same number of waves
            const bool sym = down == 1 && legacy == 4 && N <= kKdeSymMaxN &&
                             workspace_bytes >= (size_t)64 * N * sizeof(float) + kde_colpart_bytes(N) + kKdeLutBytes;
            // >= 4 waves of 3 resident blocks per SM (uniform blocks: the tail of the last wave is what is lost)
            const size_t avail = workspace_bytes - kKdeLutBytes - (sym ? kde_colpart_bytes(N) : 0);
            int cap = (int)(avail / ((size_t)N * sizeof(float)));
            if (cap > 64) cap = 64;
            int ns = ((sym ? 8 : 4) * 3 * sm_count() + qblocks - 1) / qblocks;
            ns = ns < 1 ? 1 : (ns > cap ? cap : ns);
            if (ns > (M + 255) / 256) ns = (M + 255) / 256;
            const int per = ((M + ns - 1) / ns + 7) & ~7;                       // reference points per split, a multiple of 8
            nsplit = (M + per - 1) / per;
            if (sym) {
                sym_per = per; sym_qper = qper;
                colpart = partial + (size_t)64 * N;
            }
            kde_partial_mma_kernel<<<dim3(qblocks, nsplit), kKdeThreads, smem_mma, st>>>(x, N, M, down, lut, cutoff_bound, partial, per,
                                                                                     sym ? 1 : 0, colpart);
        }
        if ((rc = check_launch("roma_kde(lut)", 2))) return rc;
    } else {
        if (half) kde_partial_kernel<true><<<grid, kKdeThreads, 0, st>>>(x, N, M, down, inv_two_var, partial, nsplit);
        else      kde_partial_kernel<false><<<grid, kKdeThreads, 0, st>>>(x, N, M, down, inv_two_var, partial, nsplit);
        if ((rc = check_launch("roma_kde(partial)"))) return rc;
    }
    kde_finalize_kernel<<<(N + 127) / 128, 128, 0, st>>>(partial, N, nsplit, half, density, p_out, poison, sym_per, sym_qper, colpart);
    return check_launch("roma_kde(finalize)");
}
