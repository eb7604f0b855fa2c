// K1 on tensor cores: the GP posterior mean as a chain of tcgen05 GEMMs (tc_gemm.cuh) around a small
// fp32 SIMT kernel that factors and inverts one 128x128 diagonal block per step.
//
//   mu = K_xy (K_yy + sigma I)^-1 f                                  (matcher.py:249-264)
//
// Workspace layout per batch element (Mp = M rounded up to 128, all fp32, K-major for the UMMA operands):
//   A   (Mp x Mp)   lower triangle: K_yy + sigma I -> L (in place);  upper triangle: L^T mirror
//   Kxy (Nx x Mp)   cross kernel
//   Ft  (D  x Mp)   RHS TRANSPOSED: f^T -> Z^T = (L^-1 f)^T -> alpha^T = (L^-T L^-1 f)^T
//   Ld  (nb x 2 x 128 x 128)   per diagonal block: L_kk^-1 and its transpose
// Keeping the RHS transposed makes every product of the solve an "A * B^T with both operands K-major" GEMM:
//   panel      L_ik   = A_ik  * (L_kk^-1)^T          Z^T_k  = F^T_k * (L_kk^-1)^T
//   update     A_ij  -= L_ik  * L_jk^T  (i>=j>k)     F^T_j -= Z^T_k * L_jk^T  (j>k)
//   backward   a^T_k  = Z^T_k * ((L_kk^-1)^T)^T      Z^T_j -= a^T_k * (L^T mirror rows j)^T  (j<k)
//   posterior  mu     = Kxy   * (a^T)^T
// Padding rows/cols (M..Mp) of A are the identity, of Ft zero, so they factor trivially and contribute 0.
#pragma once
#include "tc_gemm.cuh"
#include "gp_simt.cuh"

namespace roma {
namespace gp {

constexpr int kDB = 128;                 // diagonal block = UMMA tile
constexpr int kSolveSplit = 3;           // bf16 planes per fp32 operand in the solve GEMMs

// f^T = cos(8*pi*(W g + bias))^T : (B, D, Mp), zero for m >= M   (matcher.py:232-247, :224)
__global__ void __launch_bounds__(256)
posenc_t_kernel(float* __restrict__ Ft, const float* __restrict__ pw, const float* __restrict__ pb,
                int B, int hy, int wy, int D, int Mp) {
    const long long n = (long long)B * D * Mp;
    const long long stride = (long long)gridDim.x * blockDim.x;
    const int M = hy * wy;
    for (long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride) {
        const int m = (int)(idx % Mp);
        const int d = (int)((idx / Mp) % D);
        float v = 0.f;
        if (m < M) {
            const int y = m / wy, x = m - y * wy;
            const float gx = pixel_center(x, wy), gy = pixel_center(y, hy);
            const float lin = fmaf(pw[2 * d + 1], gy, pw[2 * d] * gx) + pb[d];
            v = cosf(25.132741228718345f * lin);
        }
        Ft[idx] = v;
    }
}

// ---- Cholesky + inverse of one 128x128 diagonal block, one CTA per batch element -------------------
// S (lower) <- chol(A_kk); V <- S^-1 (lower).  Writes L_kk back into A (lower part of the block, upper
// zeroed is NOT needed: nothing reads the diagonal block of L again) and V, V^T into Ld.
constexpr int kDiagLd = kDB + 1;
constexpr size_t kDiagSmem = (size_t)2 * kDB * kDiagLd * sizeof(float) + 32 * 33 * sizeof(float);

__global__ void __launch_bounds__(256)
diag_block_kernel(float* __restrict__ A, float* __restrict__ Ld, int ld, int k0, long long sA, long long sLd, int kblk) {
    extern __shared__ float dsm[];
    float (*S)[kDiagLd] = reinterpret_cast<float (*)[kDiagLd]>(dsm);
    float (*V)[kDiagLd] = reinterpret_cast<float (*)[kDiagLd]>(dsm + kDB * kDiagLd);
    float (*Tm)[33] = reinterpret_cast<float (*)[33]>(dsm + 2 * kDB * kDiagLd);
    const int b = blockIdx.x, tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    float* Ab = A + (long long)b * sA + (long long)k0 * ld + k0;
    float* Lo = Ld + (long long)b * sLd + (long long)kblk * 2 * kDB * kDB;

    for (int e = tid; e < kDB * kDB; e += 256) {
        const int r = e >> 7, c = e & 127;
        S[r][c] = (c <= r) ? Ab[(long long)r * ld + c] : 0.f;
        V[r][c] = 0.f;
    }
    __syncthreads();

    for (int jb = 0; jb < 4; ++jb) {
        const int j0 = jb * 32;
        if (warp == 0) {
            // potrf of the 32x32 block (left-looking inside the block), lane = row
            const int r = j0 + lane;
            for (int j = 0; j < 32; ++j) {
                float s = S[r][j0 + j];
                for (int t = 0; t < j; ++t) s = fmaf(-S[r][j0 + t], S[j0 + j][j0 + t], s);
                const float d = sqrtf(__shfl_sync(0xffffffffu, s, j));
                __syncwarp();
                if (lane == j) S[r][j0 + j] = d;
                else if (lane > j) S[r][j0 + j] = s / d;
                __syncwarp();
            }
            // inverse of the 32x32 lower-triangular block: lane = column c of the inverse
            float x[32];
#pragma unroll
            for (int i = 0; i < 32; ++i) {
                float s = (i == lane) ? 1.f : 0.f;
#pragma unroll
                for (int t = 0; t < 32; ++t)
                    if (t < i) s = fmaf(-S[j0 + i][j0 + t], x[t], s);        // x[t] == 0 for t < lane
                x[i] = (i >= lane) ? s / S[j0 + i][j0 + i] : 0.f;
            }
#pragma unroll
            for (int i = 0; i < 32; ++i) V[j0 + i][j0 + lane] = x[i];
        }
        __syncthreads();
        // panel below: X = S_panel * Vjj^T, one thread per row (row cached in registers; in-place)
        {
            const int r = j0 + 32 + tid;
            if (r < kDB) {
                float row[32], out[32];
#pragma unroll
                for (int t = 0; t < 32; ++t) row[t] = S[r][j0 + t];
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    float s = 0.f;
#pragma unroll
                    for (int t = 0; t < 32; ++t)
                        if (t <= c) s = fmaf(row[t], V[j0 + c][j0 + t], s);
                    out[c] = s;
                }
#pragma unroll
                for (int c = 0; c < 32; ++c) S[r][j0 + c] = out[c];
            }
        }
        __syncthreads();
        // trailing update of the block: S[r][c] -= sum_t S[r][j0+t] * S[c][j0+t], r >= c >= j0+32
        {
            const int n = kDB - (j0 + 32);
            for (int e = tid; e < n * n; e += 256) {
                const int rr = e / n, cc = e - rr * n;
                if (cc > rr) continue;
                const int r = j0 + 32 + rr, c = j0 + 32 + cc;
                float s = S[r][c];
#pragma unroll 8
                for (int t = 0; t < 32; ++t) s = fmaf(-S[r][j0 + t], S[c][j0 + t], s);
                S[r][c] = s;
            }
        }
        __syncthreads();
    }
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
            __syncthreads();
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
    for (int e = tid; e < kDB * kDB; e += 256) {
        const int r = e >> 7, c = e & 127;
        Lo[kDB * kDB + e] = V[c][r];                      // (L_kk^-1)^T
    }
}

struct WorkspaceTC {
    float *xn, *yn, *A, *Kxy, *Ft, *Ld;
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
    w.Kxy = take((size_t)B * Nx * w.Mp);
    w.Ft = take((size_t)B * D * w.Mp);
    w.Ld = take((size_t)B * w.nb * 2 * kDB * kDB);
    w.bytes = off;
    return w;
}

template <typename T>
static int run_tc(const T* x, const T* y, const float* pos_w, const float* pos_b, float* mu,
                  int B, int Nx, int Ny, int C, int D, int hy, int wy, float Tk, float sigma, float eps,
                  const WorkspaceTC& w, cudaStream_t st) {
    using namespace tc;
    int rc;
    const int Mp = w.Mp, nb = w.nb;
    const long long sA = (long long)Mp * Mp, sK = (long long)Nx * Mp, sF = (long long)D * Mp, sLd = (long long)nb * 2 * kDB * kDB;
    constexpr int kGramSplit = std::is_same<T, float>::value ? 3 : 1;

    // 1. norms, RHS
    norms_kernel<T><<<cdiv((long long)B * Nx * 32, 256), 256, 0, st>>>(x, w.xn, (long long)B * Nx, C);
    norms_kernel<T><<<cdiv((long long)B * Ny * 32, 256), 256, 0, st>>>(y, w.yn, (long long)B * Ny, C);
    {
        const long long n = (long long)B * D * Mp;
        long long blocks = (n + 255) / 256;
        const long long cap = (long long)sm_count() * 8;
        posenc_t_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, st>>>(w.Ft, pos_w, pos_b, B, hy, wy, D, Mp);
    }
    if ((rc = check_launch("gp prologue", 3))) return rc;

    // 2. Gram tiles with the cosine/exp epilogue: K_yy + sigma I (lower, identity padded) and K_xy, one launch
    {
        Launch L{};
        Problem& p0 = L.p[0];
        p0 = make_problem(y, y, w.A, Ny, Ny, C, C, C, Mp, (long long)Ny * C, (long long)Ny * C, sA, EPI_COS);
        p0.tiles_m = p0.tiles_n = nb; p0.lower_only = 1; p0.store_m = p0.store_n = Mp;
        p0.rn = w.yn; p0.cn = w.yn; p0.srn = p0.scn = Ny; p0.eps = eps; p0.inv_T = 1.f / Tk; p0.diag_add = sigma;
        Problem& p1 = L.p[1];
        p1 = make_problem(x, y, w.Kxy, Nx, Ny, C, C, C, Mp, (long long)Nx * C, (long long)Ny * C, sK, EPI_COS);
        p1.rn = w.xn; p1.cn = w.yn; p1.srn = Nx; p1.scn = Ny; p1.eps = eps; p1.inv_T = 1.f / Tk; p1.diag_add = 0.f;
        // the cos epilogue adds diag_add on m == n and writes identity on padding: K_xy has neither
        p1.store_m = Nx; p1.store_n = Ny;
        L.nprob = 2; L.tiles0 = p0.tiles_m * p0.tiles_n;
        if ((rc = launch_gemm_nt<T, kGramSplit>(L, B, st, "gp gram"))) return rc;
    }

    static bool diag_cfg = false;
    if (!diag_cfg) {
        cudaError_t e = cudaFuncSetAttribute(diag_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kDiagSmem);
        if (e != cudaSuccess) return fail((int)e, "gp diag: smem opt-in: %s", cudaGetErrorString(e));
        diag_cfg = true;
    }

    // 3. forward: blocked right-looking Cholesky with the forward substitution riding along
    for (int k = 0; k < nb; ++k) {
        const int k0 = k * kDB, rem = Mp - k0 - kDB;            // rows below the block
        diag_block_kernel<<<B, 256, kDiagSmem, st>>>(w.A, w.Ld, Mp, k0, sA, sLd, k);
        if ((rc = check_launch("gp diag"))) return rc;
        const float* Linv = w.Ld + (long long)k * 2 * kDB * kDB;
        {   // panel: Z^T_k = F^T_k Linv^T  [and L_ik = A_ik Linv^T with its transposed mirror]
            Launch L{};
            L.p[0] = make_problem(w.Ft + k0, Linv, w.Ft + k0, D, kDB, kDB, Mp, kDB, Mp, sF, sLd, sF, EPI_STORE);
            L.nprob = 1; L.tiles0 = L.p[0].tiles_m;
            if (rem > 0) {
                float* Ap = w.A + (long long)(k0 + kDB) * Mp + k0;
                L.p[1] = make_problem(Ap, Linv, Ap, rem, kDB, kDB, Mp, kDB, Mp, sA, sLd, sA, EPI_STORE);
                L.p[1].Ct = w.A + (long long)k0 * Mp + (k0 + kDB); L.p[1].ldct = Mp; L.p[1].sCt = sA;
                L.nprob = 2;
            }
            if ((rc = launch_gemm_nt<float, kSolveSplit>(L, B, st, "gp panel"))) return rc;
        }
        if (rem > 0) {   // trailing update
            const float* Lp = w.A + (long long)(k0 + kDB) * Mp + k0;
            Launch L{};
            L.p[0] = make_problem(Lp, Lp, w.A + (long long)(k0 + kDB) * Mp + (k0 + kDB), rem, rem, kDB, Mp, Mp, Mp, sA, sA, sA, EPI_SUB);
            L.p[0].lower_only = 1;
            L.p[1] = make_problem(w.Ft + k0, Lp, w.Ft + k0 + kDB, D, rem, kDB, Mp, Mp, Mp, sF, sA, sF, EPI_SUB);
            L.nprob = 2; L.tiles0 = L.p[0].tiles_m * L.p[0].tiles_n;
            if ((rc = launch_gemm_nt<float, kSolveSplit>(L, B, st, "gp update"))) return rc;
        }
    }
    // 4. backward substitution on the transposed RHS
    for (int k = nb - 1; k >= 0; --k) {
        const int k0 = k * kDB;
        const float* LinvT = w.Ld + (long long)k * 2 * kDB * kDB + kDB * kDB;
        {
            Launch L{};
            L.p[0] = make_problem(w.Ft + k0, LinvT, w.Ft + k0, D, kDB, kDB, Mp, kDB, Mp, sF, sLd, sF, EPI_STORE);
            L.nprob = 1; L.tiles0 = L.p[0].tiles_m;
            if ((rc = launch_gemm_nt<float, kSolveSplit>(L, B, st, "gp back diag"))) return rc;
        }
        if (k > 0) {   // Z^T_j -= a^T_k * L_kj for all j < k, using the L^T mirror rows [0, k0) x cols [k0, k0+128)
            Launch L{};
            L.p[0] = make_problem(w.Ft + k0, w.A + k0, w.Ft, D, k0, kDB, Mp, Mp, Mp, sF, sA, sF, EPI_SUB);
            L.nprob = 1; L.tiles0 = L.p[0].tiles_m * L.p[0].tiles_n;
            if ((rc = launch_gemm_nt<float, kSolveSplit>(L, B, st, "gp back update"))) return rc;
        }
    }
    // 5. mu = K_xy alpha  (token-major == channels-last)
    {
        Launch L{};
        L.p[0] = make_problem(w.Kxy, w.Ft, mu, Nx, D, Ny, Mp, Mp, D, sK, sF, (long long)Nx * D, EPI_STORE);
        L.nprob = 1; L.tiles0 = L.p[0].tiles_m * L.p[0].tiles_n;
        if ((rc = launch_gemm_nt<float, kSolveSplit>(L, B, st, "gp posterior"))) return rc;
    }
    return ROMA_OK;
}

}  // namespace gp
}  // namespace roma
