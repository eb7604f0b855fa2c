// K1 on tensor cores: the GP posterior mean as a chain of tcgen05 GEMMs (tc_gemm.cuh) around a small
// fp32 SIMT kernel that factors and inverts one 128x128 diagonal block per step.
//
//   mu = K_xy (K_yy + sigma I)^-1 f                                  (matcher.py:249-264)
//
// With K_yy + sigma I = L L^T:   mu = K_xy L^-T L^-1 f = W^T Z,   W = L^-1 K_xy^T,   Z = L^-1 f
// so ONE forward substitution applied to the stacked right-hand side [f | K_xy^T] is enough: no backward sweep,
// no explicit inverse of K_yy.
//
// Workspace per batch element (Mp = Ny rounded up to 128; fp32, every matrix K-major for the UMMA operands):
//   A  (Mp x Mp)        K_yy + sigma I (lower triangle, identity on the padding) -> L in place, block column
//                        by block column
//   R  ((D + Nx) x Mp)  the right-hand sides TRANSPOSED and stacked: rows 0..D-1 = f^T, rows D.. = K_xy;
//                        overwritten by Z^T and W^T
//   Ld (nb x 128 x 128) L_kk^-1 of every diagonal block
//   xn, yn              inverse row norms of x and y (CosKernel)
// Keeping the right-hand sides transposed makes every product an "A * B^T, both operands K-major" GEMM, which is
// the only shape `tc::gemm_nt_kernel` implements.  Left-looking blocked Cholesky, block k of nb = Mp / 128:
//   update  (k > 0)  A[k:, k]  -= L[k:, 0:k] * L[k, 0:k]^T        R[:, k] -= R[:, 0:k] * L[k, 0:k]^T     (one launch)
//   diag             L_kk = chol(A_kk),  Ld_k = L_kk^-1            (fp32 SIMT, one CTA per batch element)
//   panel            R[:, k] = R[:, k] * Ld_k^T                    L[k+1:, k] = A[k+1:, k] * Ld_k^T       (one launch)
//   posterior        mu = R[D:, :] * R[0:D, :]^T = W^T Z           (Nx x D, token-major = channels-last)
// Launches per call: 2 norms + 1 basis + 1 Gram (K_yy and K_xy as two problems of one launch) + nb diag +
// (2 nb - 1) update/panel + 1 posterior; the Gram runs on bf16/fp16 operands as stored, the solve on split-bf16
// planes of the fp32 workspace with the cross terms in a second TMEM accumulator (tc_gemm.cuh).
#pragma once
#include "tc_gemm.cuh"
#include "gp_simt.cuh"
#include <atomic>
#include <mutex>

namespace roma {
namespace gp {

constexpr int kDB = 128;                 // diagonal block = UMMA tile

// R[0:D] = f^T = cos(8*pi*(W g + bias))^T : (D x Mp) per batch element (row stride Mp, batch stride sR), zero
// for m >= M.  f does not depend on the batch element: each thread evaluates 4 consecutive m once and stores
// them to every batch element (matcher.py:232-247, :224).
__global__ void __launch_bounds__(256)
posenc_t_kernel(float* __restrict__ R, const float* __restrict__ pw, const float* __restrict__ pb,
                int B, int hy, int wy, int D, int Mp, long long sR) {
    const int M = hy * wy, q = Mp >> 2;
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= D * q) return;
    const int d = idx / q, m4 = (idx - d * q) * 4;
    const float wx = pw[2 * d], wyy = pw[2 * d + 1], bias = pb[d];
    float v[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
        const int m = m4 + e;
        v[e] = 0.f;
        if (m < M) {
            const int y = m / wy, x = m - y * wy;
            const float lin = fmaf(wyy, pixel_center(y, hy), wx * pixel_center(x, wy)) + bias;
            v[e] = cosf(25.132741228718345f * lin);
        }
    }
    const float4 o = make_float4(v[0], v[1], v[2], v[3]);
    float* dst = R + (long long)d * Mp + m4;
    for (int b = 0; b < B; ++b) *reinterpret_cast<float4*>(dst + (long long)b * sR) = o;
}

// ---- Cholesky + inverse of one 128x128 diagonal block, one CTA per batch element -------------------
// S (lower) <- chol(A_kk); V <- S^-1 (lower).  L_kk is written back into A (lower part of the block; nothing
// reads it again, it is kept for inspection) and V into Ld.
constexpr int kDiagLd = kDB + 1;
constexpr size_t kDiagSmem = (size_t)2 * kDB * kDiagLd * sizeof(float) + 32 * 33 * sizeof(float) + kDB * sizeof(float);

// pre != 0: the block first receives the rank-128 update of the block column just finished, S -= L_{k,k-1} L_{k,k-1}^T (fp32
// FMAs, lower triangle), which takes the syrk launch for THIS tile — and one cross-stream hand-off — off the dependent chain.
__global__ void __launch_bounds__(256)
diag_block_kernel(float* __restrict__ A, float* __restrict__ Ld, int ld, int k0, long long sA, long long sLd, int kblk, int pre) {
    extern __shared__ float dsm[];
    float (*S)[kDiagLd] = reinterpret_cast<float (*)[kDiagLd]>(dsm);
    float (*V)[kDiagLd] = reinterpret_cast<float (*)[kDiagLd]>(dsm + kDB * kDiagLd);
    float (*Tm)[33] = reinterpret_cast<float (*)[33]>(dsm + 2 * kDB * kDiagLd);
    float* invd = dsm + 2 * kDB * kDiagLd + 32 * 33;                 // 1 / L_jj
    const int b = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float* Ab = A + (long long)b * sA + (long long)k0 * ld + k0;
    float* Lo = Ld + (long long)b * sLd + (long long)kblk * kDB * kDB;

    for (int e = tid; e < kDB * (kDB / 4); e += 256) {
        const int r = e >> 5, c = (e & 31) * 4;
        const float4 v = *reinterpret_cast<const float4*>(Ab + (long long)r * ld + c);
        S[r][c] = (c <= r) ? v.x : 0.f; S[r][c + 1] = (c + 1 <= r) ? v.y : 0.f;
        S[r][c + 2] = (c + 2 <= r) ? v.z : 0.f; S[r][c + 3] = (c + 3 <= r) ? v.w : 0.f;
        V[r][c] = 0.f; V[r][c + 1] = 0.f; V[r][c + 2] = 0.f; V[r][c + 3] = 0.f;
    }
    __syncthreads();
    if (pre) {
        const float* Lp = Ab - kDB;                                   // L_{k,k-1}: rows of this block, the 128 columns to its left
        for (int e = tid; e < kDB * (kDB / 4); e += 256) {
            const int r = e >> 5, c = (e & 31) * 4;
            const float4 v = *reinterpret_cast<const float4*>(Lp + (long long)r * ld + c);
            V[r][c] = v.x; V[r][c + 1] = v.y; V[r][c + 2] = v.z; V[r][c + 3] = v.w;
        }
        __syncthreads();
        // 4 x 4 blocks of the lower triangle: 32 x 33 / 2 = 528 blocks over 256 threads
        for (int e = tid; e < 528; e += 256) {
            int rr = (int)((sqrtf(8.f * (float)e + 1.f) - 1.f) * 0.5f);
            while (rr * (rr + 1) / 2 > e) --rr;
            while ((rr + 1) * (rr + 2) / 2 <= e) ++rr;
            const int cc = e - rr * (rr + 1) / 2;
            const int r = rr * 4, c = cc * 4;
            float acc[4][4];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
            for (int t = 0; t < kDB; ++t) {
                float ar[4], ac[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) { ar[i] = V[r + i][t]; ac[i] = V[c + i][t]; }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], ac[j], acc[i][j]);
            }
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (c + j <= r + i) S[r + i][c + j] -= acc[i][j];
        }
        __syncthreads();
        for (int e = tid; e < kDB * kDB; e += 256) V[e >> 7][e & 127] = 0.f;
        __syncthreads();
    }

    for (int jb = 0; jb < 4; ++jb) {
        const int j0 = jb * 32;
        if (warp == 0) {
            // potrf of the 32x32 block in registers: lane = row, a[c] = S[j0+lane][j0+c]; the column of L needed
            // by the rank-1 update comes from the owning lanes by shuffle (no shared-memory round trips)
            float a[32];
#pragma unroll
            for (int c = 0; c < 32; ++c) a[c] = S[j0 + lane][j0 + c];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                // 1/sqrt once per column (MUFU.RSQ, 2 ulp) instead of an IEEE sqrt and 32 divisions on the critical path;
                // the same reciprocal scales the column here, the panel rows below and the inverse further down
                const float piv = __shfl_sync(0xffffffffu, a[j], j);
                const float rinv = rsqrtf(piv);
                a[j] = (lane == j) ? piv * rinv : a[j] * rinv;
                if (lane == j) invd[j0 + j] = rinv;
#pragma unroll
                for (int k = j + 1; k < 32; ++k) {
                    const float lk = __shfl_sync(0xffffffffu, a[j], k);
                    a[k] = fmaf(-a[j], lk, a[k]);               // meaningful for lane >= k
                }
            }
#pragma unroll
            for (int c = 0; c < 32; ++c)
                if (c <= lane) S[j0 + lane][j0 + c] = a[c];
        }
        __syncthreads();
        // panel below: X * Ljj^T = S_panel by forward substitution, one thread per row (row in registers)
        {
            const int r = j0 + 32 + tid;
            if (r < kDB) {
                float x[32];
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    float s = S[r][j0 + c];
#pragma unroll
                    for (int t = 0; t < 32; ++t)
                        if (t < c) s = fmaf(-x[t], S[j0 + c][j0 + t], s);
                    x[c] = s * invd[j0 + c];
                }
#pragma unroll
                for (int c = 0; c < 32; ++c) S[r][j0 + c] = x[c];
            }
        }
        __syncthreads();
        // trailing update inside the block: S[r][c] -= sum_t S[r][j0+t] * S[c][j0+t], r >= c >= j0+32 (4x4 per thread)
        {
            const int n4 = (kDB - (j0 + 32)) >> 2;
            for (int e = tid; e < n4 * n4; e += 256) {
                const int rr = e / n4, cc = e - rr * n4;
                if (cc > rr) continue;
                const int r = j0 + 32 + rr * 4, c = j0 + 32 + cc * 4;
                float acc[4][4];
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
#pragma unroll 8
                for (int t = 0; t < 32; ++t) {
                    float ar[4], ac[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) { ar[i] = S[r + i][j0 + t]; ac[i] = S[c + i][j0 + t]; }
#pragma unroll
                    for (int i = 0; i < 4; ++i)
#pragma unroll
                        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(ar[i], ac[j], acc[i][j]);
                }
#pragma unroll
                for (int i = 0; i < 4; ++i)
#pragma unroll
                    for (int j = 0; j < 4; ++j) S[r + i][c + j] -= acc[i][j];
            }
        }
        __syncthreads();
    }
    // inverses of the four 32x32 diagonal blocks, one warp each: lane = column c of the inverse
    if (warp < 4) {
        const int j0 = warp * 32;
        float x[32];
#pragma unroll
        for (int i = 0; i < 32; ++i) {
            float s = (i == lane) ? 1.f : 0.f;
#pragma unroll
            for (int t = 0; t < 32; ++t)
                if (t < i) s = fmaf(-S[j0 + i][j0 + t], x[t], s);            // x[t] == 0 for t < lane
            x[i] = (i >= lane) ? s * invd[j0 + i] : 0.f;
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) V[j0 + i][j0 + lane] = x[i];
    }
    __syncthreads();
    // off-diagonal blocks of the inverse: V_ij = -V_ii * sum_{t=j}^{i-1} S_it V_tj   (i > j)
    for (int j = 0; j < 3; ++j) {
        for (int i = j + 1; i < 4; ++i) {
            const int r = tid >> 3, c4 = (tid & 7) * 4;        // thread -> row r, 4 columns
            float acc[4] = {0.f, 0.f, 0.f, 0.f};
            for (int t = j; t < i; ++t) {
#pragma unroll 8
                for (int kk = 0; kk < 32; ++kk) {
                    const float a = S[i * 32 + r][t * 32 + kk];
#pragma unroll
                    for (int e = 0; e < 4; ++e) acc[e] = fmaf(a, V[t * 32 + kk][j * 32 + c4 + e], acc[e]);
                }
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) Tm[r][c4 + e] = acc[e];
            __syncthreads();
            float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll 8
            for (int kk = 0; kk < 32; ++kk) {
                const float a = V[i * 32 + r][i * 32 + kk];
#pragma unroll
                for (int e = 0; e < 4; ++e) o[e] = fmaf(a, Tm[kk][c4 + e], o[e]);
            }
#pragma unroll
            for (int e = 0; e < 4; ++e) V[i * 32 + r][j * 32 + c4 + e] = -o[e];
            __syncthreads();
        }
    }
    for (int e = tid; e < kDB * kDB; e += 256) {
        const int r = e >> 7, c = e & 127;
        if (c <= r) Ab[(long long)r * ld + c] = S[r][c];
        Lo[e] = V[r][c];                                  // L_kk^-1
    }
}

// ---- two-lane schedule of the factorisation ---------------------------------------------------------
// The dependent chain of the blocked Cholesky is  diag(k) -> panel_A(k) -> syrk(k) -> diag(k+1)  (small launches,
// <= 12 x B tiles); everything else — the left-looking updates of the block column below the diagonal and ALL
// right-hand-side work — only has to be finished one step later.  The chain therefore runs on a high-priority
// helper stream owned by the library (one per device, created on first use, never destroyed) while the bulk stays
// on the caller's stream; cudaEvents order the two lanes.  Fork/join are both on the caller's stream, so the call
// is still "asynchronous on the passed stream" (and capturable into a CUDA graph).
constexpr int kMaxLaneBlocks = 64;
struct Lanes {
    cudaStream_t chain = nullptr, abulk = nullptr;      // dependent chain (diag, panel A) / its feeders (update A, syrk)
    cudaEvent_t fork = nullptr, diag[kMaxLaneBlocks], upd[kMaxLaneBlocks], panel[kMaxLaneBlocks], syrk[kMaxLaneBlocks];
    bool ok = false;
    std::mutex enqueue;   // held while one call enqueues its two-lane schedule: cudaStreamWaitEvent snapshots an event at call
                          // time, so whole-call mutual exclusion of the ENQUEUE makes concurrent host threads on one device safe
};
static Lanes* lanes_for_current_device() {
    static std::mutex mu;
    static Lanes* table[64] = {nullptr};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    std::lock_guard<std::mutex> lock(mu);
    if (table[dev]) return table[dev]->ok ? table[dev] : nullptr;
    Lanes* l = new Lanes();
    table[dev] = l;
    int least = 0, greatest = 0;
    if (cudaDeviceGetStreamPriorityRange(&least, &greatest) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (cudaStreamCreateWithPriority(&l->chain, cudaStreamNonBlocking, greatest) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    if (cudaStreamCreateWithPriority(&l->abulk, cudaStreamNonBlocking, greatest) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    bool ok = cudaEventCreateWithFlags(&l->fork, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; ok && i < kMaxLaneBlocks; ++i)
        ok = cudaEventCreateWithFlags(&l->diag[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&l->upd[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&l->panel[i], cudaEventDisableTiming) == cudaSuccess &&
             cudaEventCreateWithFlags(&l->syrk[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) { cudaGetLastError(); return nullptr; }
    l->ok = true;
    return l;
}
extern std::atomic<int> g_lanes;         // 1 (default): two-lane schedule; 0: everything on the caller's stream, left-looking

struct WorkspaceTC {
    float *xn, *yn, *A, *R, *Ld;
    uint16_t *Apl[3], *Rpl[3];           // bf16 planes (hi, mid, lo) of the finished panels of A (= L) and R (= [Z^T ; W^T])
    int Mp, nb;
    size_t bytes;
};

static inline WorkspaceTC carve_tc(void* base, int B, int Nx, int Ny, int D) {
    WorkspaceTC w;
    size_t off = 0;
    auto take = [&](size_t nfloats) {
        float* p = base ? reinterpret_cast<float*>(reinterpret_cast<char*>(base) + off) : nullptr;
        off = (off + nfloats * sizeof(float) + 255) / 256 * 256;
        return p;
    };
    w.Mp = (Ny + kDB - 1) / kDB * kDB;
    w.nb = w.Mp / kDB;
    w.xn = take((size_t)B * Nx);
    w.yn = take((size_t)B * Ny);
    w.A = take((size_t)B * w.Mp * w.Mp);
    w.R = take((size_t)B * (D + Nx) * w.Mp);
    w.Ld = take((size_t)B * w.nb * kDB * kDB);
    for (int q = 0; q < 3; ++q) w.Apl[q] = reinterpret_cast<uint16_t*>(take(((size_t)B * w.Mp * w.Mp + 1) / 2));
    for (int q = 0; q < 3; ++q) w.Rpl[q] = reinterpret_cast<uint16_t*>(take(((size_t)B * (D + Nx) * w.Mp + 1) / 2));
    w.bytes = off;
    return w;
}

extern std::atomic<int> g_planes;        // 1: finished panels are kept as bf16 hi / lo planes and the update / syrk / posterior
                                         // GEMMs read them through TMA (gemm_planes_kernel); 0: fp32 operands re-split on the fly

// The solve, all as "X * Y^T with both operands K-major" tcgen05 GEMMs.  R = [f^T ; K_xy] ((D + Nx) x Mp) is the
// TRANSPOSED right-hand side block: forward substitution turns it into [Z^T ; W^T] = [f^T ; K_xy] L^-T, and
//   mu = K_xy (L L^T)^-1 f = (K_xy L^-T)(L^-1 f) = W^T Z,
// so no backward sweep is needed.  Left-looking blocked Cholesky, per 128-column block k:
//   update   T_ik = A_ik - sum_{j<k} L_ij L_kj^T (i >= k)      R_k -= R_{0:k} L_{k,0:k}^T        (K = 128 k)
//   diag     L_kk = chol(T_kk), L_kk^-1                                           (fp32 SIMT, one CTA per element)
//   panel    L_ik = T_ik L_kk^-T (i > k)                        R_k  = R_k L_kk^-T               (K = 128)
template <typename T, int SOLVE_SPLIT>
static int run_tc(const T* x, const T* y, const float* pos_w, const float* pos_b, float* mu,
                  int B, int Nx, int Ny, int C, int D, int hy, int wy, float Tk, float sigma, float eps,
                  const WorkspaceTC& w, cudaStream_t st) {
    using namespace tc;
    int rc;
    const int Mp = w.Mp, nb = w.nb, RR = D + Nx;
    const long long sA = (long long)Mp * Mp, sR = (long long)RR * Mp, sLd = (long long)nb * kDB * kDB;
    constexpr int kGramSplit = std::is_same<T, float>::value ? 3 : 1;
    float* Kxy = w.R + (long long)D * Mp;

    // 1. norms, RHS
    norms_kernel<T><<<cdiv((long long)B * Nx * 32, 256), 256, 0, st>>>(x, w.xn, (long long)B * Nx, C);
    norms_kernel<T><<<cdiv((long long)B * Ny * 32, 256), 256, 0, st>>>(y, w.yn, (long long)B * Ny, C);
    posenc_t_kernel<<<cdiv((long long)D * (Mp / 4), 256), 256, 0, st>>>(w.R, pos_w, pos_b, B, hy, wy, D, Mp, sR);
    if ((rc = check_launch("gp prologue", 3))) return rc;

    // 2. Gram tiles with the cosine/exp epilogue: K_yy + sigma I (lower, identity padded) and K_xy, one launch
    {
        Launch L{};
        Problem& p0 = L.p[0];
        p0 = make_problem(y, y, w.A, Ny, Ny, C, C, C, Mp, (long long)Ny * C, (long long)Ny * C, sA, EPI_COS);
        p0.tiles_m = p0.tiles_n = nb; p0.lower_only = 1; p0.store_m = p0.store_n = Mp;
        p0.rn = w.yn; p0.cn = w.yn; p0.srn = p0.scn = Ny; p0.eps = eps; p0.inv_T = 1.f / Tk; p0.diag_add = sigma;
        Problem& p1 = L.p[1];
        p1 = make_problem(x, y, Kxy, Nx, Ny, C, C, C, Mp, (long long)Nx * C, (long long)Ny * C, sR, EPI_COS);
        p1.rn = w.xn; p1.cn = w.yn; p1.srn = Nx; p1.scn = Ny; p1.eps = eps; p1.inv_T = 1.f / Tk; p1.diag_add = 0.f;
        p1.no_diag = 1; p1.tiles_n = nb; p1.store_n = Mp;          // zero-fill the padded columns of K_xy
        L.nprob = 2; L.tiles0 = p0.tiles_m * p0.tiles_n;
        if constexpr (sizeof(T) == 2) {
            if ((rc = launch_gemm_nt_16<T>(L, B, st, "gp gram", std::is_same<T, __half>::value ? ROMA_F16 : ROMA_BF16))) return rc;
        } else {
            if ((rc = launch_gemm_nt<T, kGramSplit>(L, B, st, "gp gram"))) return rc;
        }
    }

    static DeviceOnce::Flags diag_cfg;
    if (DeviceOnce::pending(diag_cfg)) {
        cudaError_t e = cudaFuncSetAttribute(diag_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDiagSmem);
        if (e != cudaSuccess) return fail((int)e, "gp diag: smem opt-in: %s", cudaGetErrorString(e));
        DeviceOnce::done(diag_cfg);
    }

    Lanes* lanes = (g_lanes.load() != 0 && nb > 1 && nb <= kMaxLaneBlocks) ? lanes_for_current_device() : nullptr;
    const bool planes = lanes != nullptr && g_planes.load() != 0 && w.Apl[0] != nullptr;
    std::unique_lock<std::mutex> lanes_lock;
    if (lanes) {
        lanes_lock = std::unique_lock<std::mutex>(lanes->enqueue);
        // 3'. two-lane schedule.  Diagonal tiles are kept up to date right-looking, the rest of the block column and the
        //     right-hand sides left-looking, one step behind the chain.  The chain holds two launches per step:
        //       chain  (sc):  [wait syrk(k-2)] diag(k) (+ its own rank-128 update from column k-1) . [wait upd(k)] panel_A(k)
        //       feeder (sb):  [wait panel(k-1)] update_A(k): A[k+1:, k] . syrk(k-1): diagonal tiles i >= k+1
        //       caller (st):  [wait panel(k-1)] update_R(k): R[:, k] . [wait diag(k)] panel_R(k)   — the right-hand sides only trail
        //     (round 1 / early round 2: syrk(k) sat on the chain between panel_A(k) and diag(k+1), and update_A waited behind
        //     the 17 right-hand-side tiles of the same launch.)
        cudaStream_t sc = lanes->chain, sb = lanes->abulk;
#define ROMA_CU(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return fail((int)e_, "gp lanes: %s", cudaGetErrorString(e_)); } while (0)
        ROMA_CU(cudaEventRecord(lanes->fork, st));
        ROMA_CU(cudaStreamWaitEvent(sc, lanes->fork, 0));
        ROMA_CU(cudaStreamWaitEvent(sb, lanes->fork, 0));
        for (int k = 0; k < nb; ++k) {
            const int k0 = k * kDB, rem = Mp - k0 - kDB;
            const float* Lk = w.A + (long long)k0 * Mp;                 // L_{k,0:k}: 128 rows x k0
            const float* Linv = w.Ld + (long long)k * kDB * kDB;
            float* Ap = w.A + (long long)(k0 + kDB) * Mp + k0;          // block column k below the diagonal
            const long long offL = (long long)k0 * Mp;
            // ---- chain: diag(k) with the rank-128 update of column k-1 folded in; it needs the syrk of columns <= k-2 (bulk) ----
            if (k >= 2) ROMA_CU(cudaStreamWaitEvent(sc, lanes->syrk[k - 2], 0));
            diag_block_kernel<<<B, 256, kDiagSmem, sc>>>(w.A, w.Ld, Mp, k0, sA, sLd, k, k > 0 ? 1 : 0);
            if ((rc = check_launch("gp diag"))) return rc;
            ROMA_CU(cudaEventRecord(lanes->diag[k], sc));
            // ---- bulk: the part of the left-looking update that gates the chain first: A[k+1:, k] -= L[k+1:, 0:k] L[k, 0:k]^T ----
            if (k > 0) {
                ROMA_CU(cudaStreamWaitEvent(sb, lanes->panel[k - 1], 0));       // L[:, 0:k] complete
                ROMA_CU(cudaStreamWaitEvent(st, lanes->panel[k - 1], 0));
                if (rem > 0) {
                    Launch L{};
                    L.p[0] = make_problem(Lk + (long long)kDB * Mp, Lk, Ap, rem, kDB, k0, Mp, Mp, Mp, sA, sA, sA, EPI_SUB);
                    L.nprob = 1; L.tiles0 = L.p[0].tiles_m;
                    bool done = false;
                    if (planes) {
                        for (int q = 0; q < SOLVE_SPLIT; ++q) { L.p[0].Apl[q] = w.Apl[q] + offL + (long long)kDB * Mp; L.p[0].Bpl[q] = w.Apl[q] + offL; }
                        if ((rc = launch_gemm_planes<SOLVE_SPLIT>(L, B, sb, "gp update A (planes)", &done))) return rc;
                    }
                    if (!done && (rc = launch_gemm_nt<float, SOLVE_SPLIT>(L, B, sb, "gp update A"))) return rc;
                }
                ROMA_CU(cudaEventRecord(lanes->upd[k], sb));
            }
            // ---- chain: panel A(k) ----
            if (rem > 0) {
                if (k > 0) ROMA_CU(cudaStreamWaitEvent(sc, lanes->upd[k], 0));
                Launch L{};
                L.p[0] = make_problem(Ap, Linv, Ap, rem, kDB, kDB, Mp, kDB, Mp, sA, sLd, sA, EPI_STORE);
                const long long offP = (long long)(k0 + kDB) * Mp + k0;
                if (planes) { L.p[0].nplanes = SOLVE_SPLIT; for (int q = 0; q < SOLVE_SPLIT; ++q) L.p[0].Pout[q] = w.Apl[q] + offP; }   // the finished panel, split once
                L.nprob = 1; L.tiles0 = L.p[0].tiles_m;
                if ((rc = launch_gemm_nt<float, SOLVE_SPLIT>(L, B, sc, "gp panel A"))) return rc;
                ROMA_CU(cudaEventRecord(lanes->panel[k], sc));
            }
            // ---- bulk: diagonal tiles i >= k+1 -= L_{i,k-1} L_{i,k-1}^T (tile k took its own update inside diag(k)) ----
            if (k > 0) {
                if (rem > 0) {
                    float* Aq = w.A + (long long)(k0 + kDB) * Mp + (k0 - kDB);      // block column k-1, rows below block k
                    const long long offQ = (long long)(k0 + kDB) * Mp + (k0 - kDB);
                    Launch S{};
                    S.p[0] = make_problem(Aq, Aq, w.A + (long long)(k0 + kDB) * Mp + (k0 + kDB), rem, rem, kDB, Mp, Mp, Mp, sA, sA, sA, EPI_SUB);
                    S.p[0].diag_only = 1;
                    S.nprob = 1; S.tiles0 = S.p[0].tiles_m;
                    bool done = false;
                    if (planes) {
                        for (int q = 0; q < SOLVE_SPLIT; ++q) S.p[0].Apl[q] = S.p[0].Bpl[q] = w.Apl[q] + offQ;
                        if ((rc = launch_gemm_planes<SOLVE_SPLIT>(S, B, sb, "gp syrk (planes)", &done))) return rc;
                    }
                    if (!done && (rc = launch_gemm_nt<float, SOLVE_SPLIT>(S, B, sb, "gp syrk"))) return rc;
                }
                ROMA_CU(cudaEventRecord(lanes->syrk[k - 1], sb));
                // ---- bulk: right-hand sides R[:, k] -= R[:, 0:k] L[k, 0:k]^T ----
                Launch L{};
                L.p[0] = make_problem(w.R, Lk, w.R + k0, RR, kDB, k0, Mp, Mp, Mp, sR, sA, sR, EPI_SUB);
                L.nprob = 1; L.tiles0 = L.p[0].tiles_m;
                bool done = false;
                if (planes) {
                    for (int q = 0; q < SOLVE_SPLIT; ++q) { L.p[0].Apl[q] = w.Rpl[q]; L.p[0].Bpl[q] = w.Apl[q] + offL; }
                    if ((rc = launch_gemm_planes<SOLVE_SPLIT>(L, B, st, "gp update R (planes)", &done))) return rc;
                }
                if (!done && (rc = launch_gemm_nt<float, SOLVE_SPLIT>(L, B, st, "gp update R"))) return rc;
            }
            ROMA_CU(cudaStreamWaitEvent(st, lanes->diag[k], 0));
            {
                Launch L{};
                L.p[0] = make_problem(w.R + k0, Linv, w.R + k0, RR, kDB, kDB, Mp, kDB, Mp, sR, sLd, sR, EPI_STORE);
                if (planes) { L.p[0].nplanes = SOLVE_SPLIT; for (int q = 0; q < SOLVE_SPLIT; ++q) L.p[0].Pout[q] = w.Rpl[q] + k0; }
                L.nprob = 1; L.tiles0 = L.p[0].tiles_m;
                if ((rc = launch_gemm_nt<float, SOLVE_SPLIT>(L, B, st, "gp panel R"))) return rc;
            }
        }
        if (nb >= 2) ROMA_CU(cudaStreamWaitEvent(st, lanes->syrk[nb - 2], 0));          // join the feeder lane (its last record)
#undef ROMA_CU
        // the chain's last operation is diag(nb-1), which panel_R(nb-1) waited for: all lanes are joined
    } else
    // 3. left-looking blocked Cholesky with the transposed right-hand sides riding along
    for (int k = 0; k < nb; ++k) {
        const int k0 = k * kDB, rows = Mp - k0, rem = rows - kDB;
        if (k > 0) {
            const float* Lk = w.A + (long long)k0 * Mp;                 // L_{k,0:k}: 128 rows x k0
            Launch L{};
            L.p[0] = make_problem(Lk, Lk, w.A + (long long)k0 * Mp + k0, rows, kDB, k0, Mp, Mp, Mp, sA, sA, sA, EPI_SUB);
            L.p[1] = make_problem(w.R, Lk, w.R + k0, RR, kDB, k0, Mp, Mp, Mp, sR, sA, sR, EPI_SUB);
            L.nprob = 2; L.tiles0 = L.p[0].tiles_m;
            if ((rc = launch_gemm_nt<float, SOLVE_SPLIT>(L, B, st, "gp update"))) return rc;
        }
        diag_block_kernel<<<B, 256, kDiagSmem, st>>>(w.A, w.Ld, Mp, k0, sA, sLd, k, 0);
        if ((rc = check_launch("gp diag"))) return rc;
        {
            const float* Linv = w.Ld + (long long)k * kDB * kDB;
            Launch L{};
            L.p[0] = make_problem(w.R + k0, Linv, w.R + k0, RR, kDB, kDB, Mp, kDB, Mp, sR, sLd, sR, EPI_STORE);
            L.nprob = 1; L.tiles0 = L.p[0].tiles_m;
            if (rem > 0) {
                float* Ap = w.A + (long long)(k0 + kDB) * Mp + k0;
                L.p[1] = make_problem(Ap, Linv, Ap, rem, kDB, kDB, Mp, kDB, Mp, sA, sLd, sA, EPI_STORE);
                L.nprob = 2;
            }
            if ((rc = launch_gemm_nt<float, SOLVE_SPLIT>(L, B, st, "gp panel"))) return rc;
        }
    }
    // 4. mu = W^T Z  (token-major == channels-last)
    {
        Launch L{};
        L.p[0] = make_problem(Kxy, w.R, mu, Nx, D, Ny, Mp, Mp, D, sR, sR, (long long)Nx * D, EPI_STORE);
        L.nprob = 1; L.tiles0 = L.p[0].tiles_m * L.p[0].tiles_n;
        bool done = false;
        if (planes) {
            for (int s = 0; s < SOLVE_SPLIT; ++s) { L.p[0].Apl[s] = w.Rpl[s] + (long long)D * Mp; L.p[0].Bpl[s] = w.Rpl[s]; }
            if ((rc = launch_gemm_planes<SOLVE_SPLIT>(L, B, st, "gp posterior (planes)", &done))) return rc;
        }
        if (!done && (rc = launch_gemm_nt<float, SOLVE_SPLIT>(L, B, st, "gp posterior"))) return rc;
    }
    return ROMA_OK;
}

}  // namespace gp
}  // namespace roma
