// fp32 SIMT building blocks of the GP posterior (K1): token norms, generic tiled GEMM with a fused
// epilogue, positional-basis generation, blocked Cholesky panels and block substitutions.
// These are the precision-reference path (true fp32 like the reference, TF32 off, matcher.py:254-263)
// and the fallback for shapes the tcgen05 path does not cover.
#pragma once
#include "common.cuh"

namespace roma {
namespace gp {

constexpr int kNB = 32;   // Cholesky / substitution block size

// ---- token L2 norms: one warp per token ---------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) norms_kernel(const T* __restrict__ x, float* __restrict__ out, long long ntok, int C) {
    const long long tok = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (tok >= ntok) return;
    const T* p = x + tok * C;
    float s = 0.f;
    for (int c = lane; c < C; c += 32) { const float v = Elem<T>::to_f(p[c]); s = fmaf(v, v, s); }
    s = warp_sum(s);
    if (lane == 0) out[tok] = sqrtf(s);
}

// ---- generic 64x64x16 SIMT GEMM ------------------------------------------------------------------
// C[m][n] = epi( sum_k A[m][k] * B'[k][n] ),  B' = B^T when kBT (B stored N x K row-major) else B (K x N).
struct GemmShape { int M, N, K; long long lda, ldb, ldc; long long sA, sB, sC; };   // s* = batch strides

struct EpiStore {   // plain store
    __device__ __forceinline__ float operator()(float acc, int, int, int) const { return acc; }
};
struct EpiCosKernel {   // K = exp((acc/(xn*yn+eps) - 1)/T) (+ sigma on the diagonal)   matcher.py:154-163, 259-261
    const float* xn; const float* yn; long long sxn, syn; float eps, inv_T, diag_add;
    __device__ __forceinline__ float operator()(float acc, int b, int m, int n) const {
        const float c = acc / (xn[b * sxn + m] * yn[b * syn + n] + eps);
        float k = expf((c - 1.f) * inv_T);
        if (m == n) k += diag_add;
        return k;
    }
};

template <typename TA, typename TB, bool kBT, typename Epi>
__global__ void __launch_bounds__(256)
gemm64_kernel(const TA* __restrict__ A, const TB* __restrict__ Bm, float* __restrict__ Cm, GemmShape s, Epi epi) {
    __shared__ float As[16][64 + 4];
    __shared__ float Bs[16][64 + 4];
    const int b = blockIdx.z;
    const int m0 = blockIdx.y * 64, n0 = blockIdx.x * 64;
    const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
    A += (long long)b * s.sA; Bm += (long long)b * s.sB; Cm += (long long)b * s.sC;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < s.K; k0 += 16) {
        {   // A tile: 64 rows x 16 k, row-major source; thread -> (row = tid/4, k = (tid%4)*4 .. +4)
            const int r = tid >> 2, kk = (tid & 3) * 4;
            const int m = m0 + r;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int k = k0 + kk + e;
                As[kk + e][r] = (m < s.M && k < s.K) ? Elem<TA>::to_f(A[(long long)m * s.lda + k]) : 0.f;
            }
        }
        if (kBT) {   // B stored N x K
            const int r = tid >> 2, kk = (tid & 3) * 4;
            const int n = n0 + r;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int k = k0 + kk + e;
                Bs[kk + e][r] = (n < s.N && k < s.K) ? Elem<TB>::to_f(Bm[(long long)n * s.ldb + k]) : 0.f;
            }
        } else {     // B stored K x N: thread -> (k = tid/16, n = (tid%16)*4 .. +4)
            const int kk = tid >> 4, nn = (tid & 15) * 4;
            const int k = k0 + kk;
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                const int n = n0 + nn + e;
                Bs[kk][nn + e] = (k < s.K && n < s.N) ? Elem<TB>::to_f(Bm[(long long)k * s.ldb + n]) : 0.f;
            }
        }
        __syncthreads();
#pragma unroll
        for (int k = 0; k < 16; ++k) {
            const float4 a4 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
            const float4 b4 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
            const float a[4] = {a4.x, a4.y, a4.z, a4.w}, bb[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int m = m0 + ty * 4 + i;
        if (m >= s.M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int n = n0 + tx * 4 + j;
            if (n < s.N) Cm[(long long)m * s.ldc + n] = epi(acc[i][j], b, m, n);
        }
    }
}

// ---- positional basis f = cos(8*pi*(W g + bias)) : (B, N, D) -------------------------------------
// matcher.py:232-247 (pixel-centre grid, x then y) and :224 (fourier basis).
__global__ void __launch_bounds__(256)
posenc_kernel(float* __restrict__ Z, const float* __restrict__ pw, const float* __restrict__ pb, int B, int hy, int wy, int D) {
    const long long n = (long long)B * hy * wy * D;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride) {
        const int d = (int)(idx % D);
        const long long t = idx / D;
        const int tok = (int)(t % ((long long)hy * wy));
        const int y = tok / wy, x = tok - y * wy;
        const float gx = pixel_center(x, wy), gy = pixel_center(y, hy);
        // conv 1x1: w[d][0]*gx + w[d][1]*gy + bias, then * fp32(8*pi), then cos
        const float lin = fmaf(pw[2 * d + 1], gy, pw[2 * d] * gx) + pb[d];
        Z[idx] = cosf(25.132741228718345f * lin);
    }
}

// ---- left-looking blocked Cholesky panel ---------------------------------------------------------
// One launch per 32-column panel k.  CTA (rb, b) owns off-diagonal rows R = k+32+rb*64 .. +64 of the
// panel; every CTA redundantly forms and factors the 32x32 diagonal block (cheap) so that no CTA has
// to wait for another.  The factored diagonal block goes to Ld (never into A, which other CTAs are
// still reading); off-diagonal blocks of L overwrite A in place (strictly lower part).
__global__ void __launch_bounds__(256)
chol_panel_kernel(float* __restrict__ A, float* __restrict__ Ld, int N, int k, long long sA, long long sLd) {
    __shared__ float Lr[96][kNB + 1];     // rows [diag 32 | own 64] x 32 k-chunk
    __shared__ float P[96][kNB + 1];      // accumulated product / result
    const int b = blockIdx.y;
    A += (long long)b * sA;
    Ld += (long long)b * sLd + (long long)(k / kNB) * kNB * kNB;
    const int tid = threadIdx.x;
    const int own0 = k + kNB + blockIdx.x * 64;
    auto grow = [&](int r) { return r < kNB ? k + r : own0 + (r - kNB); };   // global row of local row r

    // each thread: 3 rows x 4 cols of the 96x32 product
    const int pr = (tid >> 3) * 3, pc = (tid & 7) * 4;
    float acc[3][4];
#pragma unroll
    for (int i = 0; i < 3; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int j0 = 0; j0 < k; j0 += kNB) {
        for (int e = tid; e < 96 * kNB; e += 256) {
            const int r = e >> 5, c = e & 31;
            const int gr = grow(r);
            Lr[r][c] = (gr < N) ? A[(long long)gr * N + j0 + c] : 0.f;
        }
        __syncthreads();
#pragma unroll 8
        for (int c = 0; c < kNB; ++c) {
            float a[3], bb[4];
#pragma unroll
            for (int i = 0; i < 3; ++i) a[i] = Lr[pr + i][c];
#pragma unroll
            for (int j = 0; j < 4; ++j) bb[j] = Lr[pc + j][c];
#pragma unroll
            for (int i = 0; i < 3; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
        }
        __syncthreads();
    }
    // P = A[rows, k:k+32] - acc ; virtual identity padding past N
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int r = pr + i, gr = grow(r);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gc = k + pc + j;
            float v;
            if (gr < N && gc < N) v = A[(long long)gr * N + gc] - acc[i][j];
            else v = (gr == gc) ? 1.f : 0.f;
            P[r][pc + j] = v;
        }
    }
    __syncthreads();
    // potrf of the 32x32 diagonal block P[0:32][0:32] by warp 0 (lower triangle, column by column)
    if (tid < 32) {
        for (int j = 0; j < kNB; ++j) {
            const float djj = sqrtf(P[j][j]);
            __syncwarp();
            if (tid == j) P[j][j] = djj;
            else if (tid > j) P[tid][j] = P[tid][j] / djj;
            __syncwarp();
            // trailing update of columns j+1.. : P[i][c] -= P[i][j]*P[c][j]  for i>=c>j ; lane = row i
            if (tid > j) {
                const float lij = P[tid][j];
                for (int c = j + 1; c <= tid; ++c) P[tid][c] = fmaf(-lij, P[c][j], P[tid][c]);
            }
            __syncwarp();
        }
    }
    __syncthreads();
    // trsm: own rows X * L_kk^T = S, one thread per row
    if (tid < 64) {
        const int r = kNB + tid;
        if (grow(r) < N) {
            float xr[kNB];
#pragma unroll
            for (int j = 0; j < kNB; ++j) {
                float s = P[r][j];
#pragma unroll
                for (int t = 0; t < kNB; ++t)
                    if (t < j) s = fmaf(-xr[t], P[j][t], s);
                xr[j] = s / P[j][j];
            }
#pragma unroll
            for (int j = 0; j < kNB; ++j) P[r][j] = xr[j];
        }
    }
    __syncthreads();
    // write back: off-diagonal rows to A (in place), diagonal block to Ld by CTA 0
    for (int e = tid; e < 64 * kNB; e += 256) {
        const int r = kNB + (e >> 5), c = e & 31;
        const int gr = grow(r), gc = k + c;
        if (gr < N && gc < N) A[(long long)gr * N + gc] = P[r][c];
    }
    if (blockIdx.x == 0) {
        for (int e = tid; e < kNB * kNB; e += 256) {
            const int r = e >> 5, c = e & 31;
            Ld[e] = (c <= r) ? P[r][c] : 0.f;
        }
    }
}

// ---- block substitution: Z <- L^-1 Z (forward) or Z <- L^-T Z (backward), 32 RHS columns per CTA --
// The RHS columns are independent, so each CTA walks ALL block rows for its own 32 columns with no
// inter-CTA dependency: the whole triangular solve is one launch.
template <bool kBackward>
__global__ void __launch_bounds__(256)
subst_kernel(const float* __restrict__ A, const float* __restrict__ Ld, float* __restrict__ Z,
             int N, int D, long long sA, long long sLd, long long sZ) {
    __shared__ float Lb[kNB][kNB + 1];
    __shared__ float Zb[kNB][kNB + 1];
    __shared__ float Tt[kNB][kNB + 1];
    const int b = blockIdx.y, c0 = blockIdx.x * kNB;
    A += (long long)b * sA; Ld += (long long)b * sLd; Z += (long long)b * sZ;
    const int tid = threadIdx.x;
    const int tr = tid >> 3, tc = (tid & 7) * 4;   // thread -> row tr, cols tc..tc+4 of the 32x32 tile
    const int nb = (N + kNB - 1) / kNB;
    for (int step = 0; step < nb; ++step) {
        const int kb = kBackward ? nb - 1 - step : step;
        float acc[4] = {0.f, 0.f, 0.f, 0.f};
        const int jb_begin = kBackward ? kb + 1 : 0, jb_end = kBackward ? nb : kb;
        for (int jb = jb_begin; jb < jb_end; ++jb) {
            for (int e = tid; e < kNB * kNB; e += 256) {
                const int r = e >> 5, c = e & 31;
                float lv;
                if (!kBackward) {   // L[kb rows][jb cols]
                    const int gr = kb * kNB + r, gc = jb * kNB + c;
                    lv = (gr < N && gc < N) ? A[(long long)gr * N + gc] : 0.f;
                    Lb[r][c] = lv;
                } else {            // (L[jb rows][kb cols])^T -> Lb[kb-local][jb-local]; consecutive threads read
                    const int rr = c, cc = r;   // consecutive columns of one L row (coalesced)
                    const int gr = jb * kNB + cc, gc = kb * kNB + rr;
                    lv = (gr < N && gc < N) ? A[(long long)gr * N + gc] : 0.f;
                    Lb[rr][cc] = lv;
                }
                const int zr = jb * kNB + r, zc = c0 + c;
                Zb[r][c] = (zr < N && zc < D) ? Z[(long long)zr * D + zc] : 0.f;
            }
            __syncthreads();
#pragma unroll 8
            for (int t = 0; t < kNB; ++t) {
                const float l = Lb[tr][t];
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[j] = fmaf(l, Zb[t][tc + j], acc[j]);
            }
            __syncthreads();
        }
        // T = Z[kb] - acc ; diagonal solve with Ld[kb]
        for (int e = tid; e < kNB * kNB; e += 256) {
            const int r = e >> 5, c = e & 31;
            Lb[r][c] = Ld[(long long)kb * kNB * kNB + e];   // lower triangular (identity-padded past N)
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int zr = kb * kNB + tr, zc = c0 + tc + j;
            Tt[tr][tc + j] = ((zr < N && zc < D) ? Z[(long long)zr * D + zc] : 0.f) - acc[j];
        }
        __syncthreads();
        if (tid < kNB) {   // one thread per RHS column
            float xcol[kNB];
            if (!kBackward) {
#pragma unroll
                for (int r = 0; r < kNB; ++r) {
                    float s = Tt[r][tid];
#pragma unroll
                    for (int t = 0; t < kNB; ++t)
                        if (t < r) s = fmaf(-Lb[r][t], xcol[t], s);
                    xcol[r] = s / Lb[r][r];
                }
            } else {
#pragma unroll
                for (int r = kNB - 1; r >= 0; --r) {
                    float s = Tt[r][tid];
#pragma unroll
                    for (int t = 0; t < kNB; ++t)
                        if (t > r) s = fmaf(-Lb[t][r], xcol[t], s);
                    xcol[r] = s / Lb[r][r];
                }
            }
#pragma unroll
            for (int r = 0; r < kNB; ++r) Tt[r][tid] = xcol[r];
        }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int zr = kb * kNB + tr, zc = c0 + tc + j;
            if (zr < N && zc < D) Z[(long long)zr * D + zc] = Tt[tr][tc + j];
        }
        __syncthreads();   // Z rows of this block are re-read (as Zb) by later steps of this same CTA
    }
}

}  // namespace gp
}  // namespace roma
