// K1: GP coarse-matcher posterior mean  mu = K_xy (K_yy + sigma I)^-1 f   (matcher.py:249-264).
//
// Pipeline (per call, all batch elements at once, every stage asynchronous on the caller's stream):
//   1. token norms |x|, |y|
//   2. K_yy + sigma*I and K_xy   — Gram tiles with the cosine / exp epilogue fused (K_xx, which the
//      reference computes and throws away under no_cov=True, matcher.py:255, is never formed)
//   3. f = cos(8*pi*pos_conv(grid))  generated on the fly into the RHS buffer
//   4. blocked Cholesky of K_yy + sigma*I (SPD, well conditioned: lambda_min >= sigma), instead of the
//      reference's explicit LU inverse (torch.linalg.inv, matcher.py:261) — same result to ~1e-5
//   5. alpha = L^-T L^-1 f  by two one-launch block substitutions (RHS columns are independent)
//   6. mu = K_xy alpha, written token-major == channels-last (the rearrange of matcher.py:264 is free)
#include "gp_simt.cuh"
#include "gp_tc.cuh"
#include <atomic>

namespace roma {
namespace gp {

struct Workspace {
    float *xn, *yn, *A, *Ld, *Kxy, *Z;
    size_t bytes;
};

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

static Workspace carve(void* base, int B, int Nx, int Ny, int C, int D) {
    (void)C;
    Workspace w;
    size_t off = 0;
    auto take = [&](size_t nfloats) {
        float* p = base ? reinterpret_cast<float*>(reinterpret_cast<char*>(base) + off) : nullptr;
        off = align_up(off + nfloats * sizeof(float), 256);
        return p;
    };
    const int nb = (Ny + kNB - 1) / kNB;
    w.xn = take((size_t)B * Nx);
    w.yn = take((size_t)B * Ny);
    w.A = take((size_t)B * Ny * Ny);
    w.Ld = take((size_t)B * nb * kNB * kNB);
    w.Kxy = take((size_t)B * Nx * Ny);
    w.Z = take((size_t)B * Ny * D);
    w.bytes = off;
    return w;
}

template <typename T>
static int run(const T* x, const T* y, const float* pos_w, const float* pos_b, float* mu,
               int B, int Nx, int Ny, int C, int D, int hy, int wy, float Tk, float sigma, float eps,
               const Workspace& w, cudaStream_t st) {
    int rc;
    // 1. norms
    norms_kernel<T><<<cdiv((long long)B * Nx * 32, 256), 256, 0, st>>>(x, w.xn, (long long)B * Nx, C);
    norms_kernel<T><<<cdiv((long long)B * Ny * 32, 256), 256, 0, st>>>(y, w.yn, (long long)B * Ny, C);
    if ((rc = check_launch("gp norms", 2))) return rc;
    // 2. Gram + cosine/exp epilogue
    {
        GemmShape s{Ny, Ny, C, C, C, Ny, (long long)Ny * C, (long long)Ny * C, (long long)Ny * Ny};
        EpiCosKernel e{w.yn, w.yn, Ny, Ny, eps, 1.f / Tk, sigma};
        gemm64_kernel<T, T, true, EpiCosKernel><<<dim3(cdiv(Ny, 64), cdiv(Ny, 64), B), 256, 0, st>>>(y, y, w.A, s, e);
    }
    {
        GemmShape s{Nx, Ny, C, C, C, Ny, (long long)Nx * C, (long long)Ny * C, (long long)Nx * Ny};
        EpiCosKernel e{w.xn, w.yn, Nx, Ny, eps, 1.f / Tk, 0.f};
        gemm64_kernel<T, T, true, EpiCosKernel><<<dim3(cdiv(Ny, 64), cdiv(Nx, 64), B), 256, 0, st>>>(x, y, w.Kxy, s, e);
    }
    if ((rc = check_launch("gp gram", 2))) return rc;
    // 3. RHS
    {
        const long long n = (long long)B * Ny * D;
        long long blocks = (n + 255) / 256;
        const long long cap = (long long)sm_count() * 8;
        posenc_kernel<<<(int)(blocks < cap ? blocks : cap), 256, 0, st>>>(w.Z, pos_w, pos_b, B, hy, wy, D);
        if ((rc = check_launch("gp posenc"))) return rc;
    }
    // 4. Cholesky, one launch per 32-column panel
    const int nb = (Ny + kNB - 1) / kNB;
    for (int kb = 0; kb < nb; ++kb) {
        const int k = kb * kNB;
        const int rem = Ny - k - kNB;
        const int gx = rem > 0 ? cdiv(rem, 64) : 1;
        chol_panel_kernel<<<dim3(gx, B), 256, 0, st>>>(w.A, w.Ld, Ny, k, (long long)Ny * Ny, (long long)nb * kNB * kNB);
    }
    if ((rc = check_launch("gp cholesky", nb))) return rc;
    // 5. substitutions
    subst_kernel<false><<<dim3(cdiv(D, kNB), B), 256, 0, st>>>(w.A, w.Ld, w.Z, Ny, D, (long long)Ny * Ny, (long long)nb * kNB * kNB, (long long)Ny * D);
    subst_kernel<true><<<dim3(cdiv(D, kNB), B), 256, 0, st>>>(w.A, w.Ld, w.Z, Ny, D, (long long)Ny * Ny, (long long)nb * kNB * kNB, (long long)Ny * D);
    if ((rc = check_launch("gp substitution", 2))) return rc;
    // 6. mu = K_xy alpha
    {
        GemmShape s{Nx, D, Ny, Ny, D, D, (long long)Nx * Ny, (long long)Ny * D, (long long)Nx * D};
        gemm64_kernel<float, float, false, EpiStore><<<dim3(cdiv(D, 64), cdiv(Nx, 64), B), 256, 0, st>>>(w.Kxy, w.Z, mu, s, EpiStore{});
        if ((rc = check_launch("gp posterior gemm"))) return rc;
    }
    return ROMA_OK;
}

}  // namespace gp
}  // namespace roma

namespace roma { namespace tc { std::atomic<int> g_tc_tma{1}; } }
extern "C" int roma_tc_set_tma(int on) {
    roma::tc::g_tc_tma.store(on ? 1 : 0);
    return 0;
}
namespace roma { namespace gp { static std::atomic<int> g_path{0}; } }   // 0 auto, 1 SIMT fp32, 2 tcgen05 (3-plane solve), 3 tcgen05 (2-plane solve)

namespace roma { namespace gp { std::atomic<int> g_lanes{1}; std::atomic<int> g_planes{1}; } }
extern "C" int roma_gp_set_planes(int on) {
    roma::gp::g_planes.store(on ? 1 : 0);
    return ROMA_OK;
}
extern "C" int roma_gp_set_lanes(int on) {
    roma::gp::g_lanes.store(on ? 1 : 0);
    return ROMA_OK;
}

extern "C" int roma_gp_set_path(int path) {
    if (path < 0 || path > 3) return roma::fail(ROMA_ERR_BAD_ARG, "roma_gp_set_path: path must be 0 (auto), 1 (simt) or 2 (tcgen05)");
    roma::gp::g_path.store(path);
    return ROMA_OK;
}

extern "C" size_t roma_gp_workspace_bytes(int B, int Nx, int Ny, int C, int D) {
    if (B <= 0 || Nx <= 0 || Ny <= 0 || C <= 0 || D <= 0) return 0;
    const size_t a = roma::gp::carve(nullptr, B, Nx, Ny, C, D).bytes;
    const size_t b = roma::gp::carve_tc(nullptr, B, Nx, Ny, D).bytes;
    return a > b ? a : b;
}

extern "C" int roma_gp_posterior(const void* x, const void* y, const float* pos_w, const float* pos_b,
                                 float* mu, int B, int Nx, int Ny, int C, int D, int hy, int wy,
                                 float T, float sigma_noise, float eps, int dtype,
                                 void* workspace, size_t workspace_bytes, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && y && pos_w && pos_b && mu && workspace, ROMA_ERR_BAD_ARG, "roma_gp_posterior: null pointer");
    ROMA_REQUIRE(B > 0 && Nx > 0 && Ny > 0 && C > 0 && D > 0 && hy > 0 && wy > 0, ROMA_ERR_BAD_ARG, "roma_gp_posterior: bad sizes");
    ROMA_REQUIRE(hy * wy == Ny, ROMA_ERR_BAD_ARG, "roma_gp_posterior: hy*wy (%d) != Ny (%d)", hy * wy, Ny);
    ROMA_REQUIRE(T > 0.f, ROMA_ERR_BAD_ARG, "roma_gp_posterior: temperature must be positive");
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, ROMA_ERR_ALIGNMENT, "roma_gp_posterior: workspace must be 256-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    // tcgen05 path: 16-byte aligned rows for the operand staging (16-bit: C % 8, fp32: C % 4)
    const int esz = dtype == ROMA_F32 ? 4 : 2;
    const bool tc_ok = (C * esz) % 16 == 0 && ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15u) == 0 && D % 4 == 0;
    const int path = gp::g_path.load();
    ROMA_REQUIRE(path < 2 || tc_ok, ROMA_ERR_UNSUPPORTED, "roma_gp_posterior: tcgen05 path needs 16-byte aligned token rows (C=%d) and D %% 4 == 0", C);
    if (tc_ok && path != 1) {
        gp::WorkspaceTC w = gp::carve_tc(workspace, B, Nx, Ny, D);
        ROMA_REQUIRE(workspace_bytes >= w.bytes, ROMA_ERR_WORKSPACE, "roma_gp_posterior: workspace %zu < required %zu", workspace_bytes, w.bytes);
#define ROMA_GP_TC(TT, SP) gp::run_tc<TT, SP>((const TT*)x, (const TT*)y, pos_w, pos_b, mu, B, Nx, Ny, C, D, hy, wy, T, sigma_noise, eps, w, st)
        // auto / path 2: 3-plane solve (6 MMAs per k-step, fp32-grade products).  The 2-plane solve (path 3: 3 MMAs, ~2^-16
        // relative) is 20 % faster but NOT accurate enough when K_yy is ill-conditioned: on photographs with flat regions
        // (config 1, sacre_coeur) its 1e-4-level error in mu flips near-tied coarse arg-maxes on 0.14 % of the pixels, while
        // the 3-plane solve agrees with the reference on the same GPU to 3e-6 everywhere (tools/config1_gp_paths.py)
        const bool two = (path == 3);
        switch (dtype) {
            case ROMA_F32:  return two ? ROMA_GP_TC(float, 2) : ROMA_GP_TC(float, 3);
            case ROMA_F16:  return two ? ROMA_GP_TC(__half, 2) : ROMA_GP_TC(__half, 3);
            case ROMA_BF16: return two ? ROMA_GP_TC(__nv_bfloat16, 2) : ROMA_GP_TC(__nv_bfloat16, 3);
            default: return fail(ROMA_ERR_BAD_ARG, "roma_gp_posterior: unknown dtype %d", dtype);
        }
#undef ROMA_GP_TC
    }
    gp::Workspace w = gp::carve(workspace, B, Nx, Ny, C, D);
    ROMA_REQUIRE(workspace_bytes >= w.bytes, ROMA_ERR_WORKSPACE, "roma_gp_posterior: workspace %zu < required %zu", workspace_bytes, w.bytes);
    switch (dtype) {
        case ROMA_F32:  return gp::run<float>((const float*)x, (const float*)y, pos_w, pos_b, mu, B, Nx, Ny, C, D, hy, wy, T, sigma_noise, eps, w, st);
        case ROMA_F16:  return gp::run<__half>((const __half*)x, (const __half*)y, pos_w, pos_b, mu, B, Nx, Ny, C, D, hy, wy, T, sigma_noise, eps, w, st);
        case ROMA_BF16: return gp::run<__nv_bfloat16>((const __nv_bfloat16*)x, (const __nv_bfloat16*)y, pos_w, pos_b, mu, B, Nx, Ny, C, D, hy, wy, T, sigma_noise, eps, w, st);
        default: return fail(ROMA_ERR_BAD_ARG, "roma_gp_posterior: unknown dtype %d", dtype);
    }
}


// Batched C = A * B^T on the tcgen05 path (tc_gemm.cuh), exposed so the tensor-core building block of K1 can be
// unit-tested and timed on its own.  A: batch x M x K, B: batch x N x K (row-major), C: batch x M x N fp32.
extern "C" int roma_tc_gemm_nt(const void* A, const void* Bm, float* Cm, int batch, int M, int N, int K,
                               int dtype, int split, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(A && Bm && Cm && batch > 0 && M > 0 && N > 0 && K > 0, ROMA_ERR_BAD_ARG, "roma_tc_gemm_nt: bad arguments");
    const int esz = dtype == ROMA_F32 ? 4 : 2;
    ROMA_REQUIRE((K * esz) % 16 == 0 && ((reinterpret_cast<uintptr_t>(A) | reinterpret_cast<uintptr_t>(Bm)) & 15u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_tc_gemm_nt: rows must be 16-byte aligned (K=%d)", K);
    tc::Launch L{};
    L.p[0] = tc::make_problem(A, Bm, Cm, M, N, K, K, K, N, (long long)M * K, (long long)N * K, (long long)M * N, tc::EPI_STORE);
    L.nprob = 1; L.tiles0 = L.p[0].tiles_m * L.p[0].tiles_n;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (dtype == ROMA_F32) {
        if (split == 2) return tc::launch_gemm_nt<float, 2>(L, batch, st, "roma_tc_gemm_nt");
        if (split == 3) return tc::launch_gemm_nt<float, 3>(L, batch, st, "roma_tc_gemm_nt");
        return fail(ROMA_ERR_BAD_ARG, "roma_tc_gemm_nt: fp32 operands need split 2 or 3");
    }
    ROMA_REQUIRE(split == 1, ROMA_ERR_BAD_ARG, "roma_tc_gemm_nt: 16-bit operands are exact with split 1");
    if (dtype == ROMA_F16) return tc::launch_gemm_nt_16<__half>(L, batch, st, "roma_tc_gemm_nt", dtype);
    if (dtype == ROMA_BF16) return tc::launch_gemm_nt_16<__nv_bfloat16>(L, batch, st, "roma_tc_gemm_nt", dtype);
    return fail(ROMA_ERR_BAD_ARG, "roma_tc_gemm_nt: unknown dtype %d", dtype);
}
