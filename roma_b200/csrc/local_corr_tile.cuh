// K2, shared-patch form for COHERENT flows (trained weights): local correlation of an 8 x 8 pixel tile as ONE tensor-core
// GEMM against the union of the tile's windows.
//
// Reference: romatch/utils/local_correlation.py:4-48 inside ConvRefiner.forward (matcher.py:105-132).
//
// The per-pixel kernels (local_corr_mma_kernel) drag every pixel's own (2r+2)^2 x C window through L2 -> L1 -> registers:
// 64 KB per pixel at C = 512, r = 3 — 12 GB per launch for 0.8 GB of compulsory HBM traffic, L2-bandwidth-bound.  When the
// flow is coherent the windows of neighbouring pixels overlap almost entirely: the 64 pixels of an 8 x 8 tile need a
// patch of at most PW x PW positions of f1 (PW = 7 + (2r+2) + 11..12 of slack: displacement gradients up to ~1 px / px), so the tile's f0 rows (A, 64 x C) and the patch
// (B, PW^2 x C) are loaded ONCE by TMA (4-D boxes (64 ch, x, y, 1) straight into the UMMA K-major SWIZZLE_128B layout;
// out-of-image positions are zero-filled by the copy engine = padding_mode "zeros"), D = A B^T runs on tcgen05 with the
// fp32 accumulator in TMEM, and each pixel picks its (2r+2)^2 neighbour products out of its TMEM lane and blends them with
// the shared bilinear weights exactly like the per-pixel kernels.  ~13 KB per pixel through L2 instead of 64 + 4.
//
// Tiles whose windows do not fit the patch (incoherent flow, flow discontinuities, non-finite flow) are left to the
// per-pixel kernel: tile_verdict_kernel (64 threads per tile) writes one byte per tile (taken / declined) into a tile map in
// the caller's workspace; this kernel's CTAs (one per tile, raster order) exit at once on a declined tile; the per-pixel
// kernel, launched next in its usual flat form, skips the pixels of the taken tiles.  (Tried and
// measured: the per-pixel kernel re-deriving the verdict per tile — its ~35 k CTAs that only exit cost 0.17 / 0.8 ms at
// s8' / s4'; the per-pixel form folded into this kernel — its 100 KB of shared memory per CTA shrink the L1 the
// per-pixel form lives on: 0.85 -> 1.38 ms at s8' on incoherent flows.)
//
// CTA = 256 threads, 2 CTAs / SM (<= 256 TMEM columns, ~100 KB of shared memory each):
//   warp 0 lane 0  TMA producer (2-stage ring over the C / 64 k-blocks, two patch halves = two passes over K)
//   warp 1 lane 0  tcgen05.mma issuer (M = 128 with rows 64..127 unused, N = patch positions of the pass, K = 16)
//   warps 4, 5     TMEM lanes 0..63 = the tile's pixels: after each pass, window rows -> shared memory
//   warps 2,3,6,7  the streaming segments meanwhile: x copy, x_hat (4-tap blend), displacement embedding, pitch padding
//   all            blend (2r+1)^2 correlations per pixel from the staged products, store
#pragma once
#include "tc_gemm.cuh"

namespace roma {


template <int R> struct TileGeom {
    static constexpr int W2 = 2 * R + 2, NN = W2 * W2, K1 = 2 * R + 1, K = K1 * K1;
    static constexpr int NPASS = 3;                             // the patch is multiplied in NPASS row bands (<= 256 TMEM columns each)
    static constexpr int PW = R == 2 ? 24 : 27;                 // patch side = 7 + W2 + slack: slack 11 (r = 2), 12 (r = 3) positions
    static constexpr int ROWS = PW / NPASS;                     // patch rows per pass: 8 / 9
    static constexpr int NCOL = ROWS * PW;                      // D columns in use per pass: 192 / 243
    static constexpr int NPAD = (NCOL + 15) / 16 * 16;          // MMA N: 192 / 256
    static constexpr int BBYTES = (NPAD * 128 + 1023) / 1024 * 1024;
    // A: 64 rows loaded (8 KB); the MMA reads 128 rows, rows 64..127 = the first 8 KB of B (their products are never read)
    static constexpr int AOFF = 8192;
    static constexpr int STAGE = AOFF + BBYTES;
    static constexpr int DSTRIDE = NN + 1;                      // staged products per pixel (+1: bank spread)
    static constexpr size_t SMEM = 1024 + 2 * (size_t)STAGE + (size_t)kTilePx * DSTRIDE * sizeof(float);
    static_assert(PW % NPASS == 0 && NPAD <= 256 && BBYTES >= 8192, "patch pass must fit one MMA and 256 TMEM columns");
};

struct TileMaps { CUtensorMap f0, f1; };

// Window bases of the tile's pixels and whether their union fits a pw x pw patch.  Called by all 256 threads of the CTA;
// threads 0..63 own one pixel each (m = ty * 8 + tx).  Returns the verdict to every thread; X0 / Y0 = patch origin.
// red: 8 ints of shared scratch.
__device__ __forceinline__ bool tile_window_fits(const GatherParams& p, int b, int ty0, int tx0, int pw, int* red,
                                                 bool& valid, int& x0, int& y0, float& wx0, float& wx1, float& wy0, float& wy1,
                                                 float& gx, float& gy, int& X0, int& Y0, int* rows_needed = nullptr) {
    const int m = threadIdx.x;
    valid = false;
    x0 = y0 = 0; wx0 = wx1 = wy0 = wy1 = 0.f; gx = gy = 0.f;
    int lo_x = INT_MAX, hi_x = INT_MIN, lo_y = INT_MAX, hi_y = INT_MIN;
    if (m < kTilePx) {
        const int y = ty0 + (m >> 3), x = tx0 + (m & 7);
        if (y < p.h && x < p.w) {
            valid = true;
            const float* f = p.flow + (long long)b * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
            gx = __ldg(f);
            gy = __ldg(f + p.fs.c);
            sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
            lo_x = hi_x = x0; lo_y = hi_y = y0;
        }
        lo_x = __reduce_min_sync(0xffffffffu, lo_x); hi_x = __reduce_max_sync(0xffffffffu, hi_x);
        lo_y = __reduce_min_sync(0xffffffffu, lo_y); hi_y = __reduce_max_sync(0xffffffffu, hi_y);
        if ((m & 31) == 0) { red[(m >> 5) * 4 + 0] = lo_x; red[(m >> 5) * 4 + 1] = hi_x; red[(m >> 5) * 4 + 2] = lo_y; red[(m >> 5) * 4 + 3] = hi_y; }
    }
    __syncthreads();
    lo_x = min(red[0], red[4]); hi_x = max(red[1], red[5]);
    lo_y = min(red[2], red[6]); hi_y = max(red[3], red[7]);
    const int W2 = 2 * p.r + 2;
    X0 = lo_x - p.r; Y0 = lo_y - p.r;
    if (rows_needed) *rows_needed = lo_x == INT_MAX ? 0 : (int)min((long long)hi_y - lo_y + W2, (long long)pw);
    // a non-finite flow clamps to +-2^20 in sample_coords: the spans below then exceed any patch
    return lo_x != INT_MAX && (long long)hi_x - lo_x + W2 <= pw && (long long)hi_y - lo_y + W2 <= pw;
}

// The per-pixel kernel's side of the same decision, for FOUR horizontally adjacent tiles at once (one prologue — flow loads,
// window bases, reduction, barrier — per 256 pixels instead of per 64): thread t looks at pixel t & 63 of tile
// 4 * blockIdx.x + (t >> 6).  bit q of the result is set when tile q belongs to local_corr_tile_kernel (or lies outside the image).
// red: 32 ints of shared scratch.  Same arithmetic as tile_window_fits: both kernels must agree on every tile.
constexpr int kPxTilesPerCta = 4;
__device__ __forceinline__ unsigned tiles_fit4(const GatherParams& p, int b, int ty0, int txq0, int pw, int* red) {
    const int t = threadIdx.x, q = t >> 6, m = t & 63;
    int lo_x = INT_MAX, hi_x = INT_MIN, lo_y = INT_MAX, hi_y = INT_MIN;
    const int y = ty0 + (m >> 3), x = (txq0 + q) * kTileSide + (m & 7);
    if (y < p.h && x < p.w) {
        const float* f = p.flow + (long long)b * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
        int x0, y0; float wx0, wx1, wy0, wy1;
        sample_coords(__ldg(f), __ldg(f + p.fs.c), p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
        lo_x = hi_x = x0; lo_y = hi_y = y0;
    }
    lo_x = __reduce_min_sync(0xffffffffu, lo_x); hi_x = __reduce_max_sync(0xffffffffu, hi_x);
    lo_y = __reduce_min_sync(0xffffffffu, lo_y); hi_y = __reduce_max_sync(0xffffffffu, hi_y);
    if ((t & 31) == 0) { int* r = red + (t >> 5) * 4; r[0] = lo_x; r[1] = hi_x; r[2] = lo_y; r[3] = hi_y; }
    __syncthreads();
    const int W2 = 2 * p.r + 2;
    unsigned fits = 0u;
#pragma unroll
    for (int k = 0; k < kPxTilesPerCta; ++k) {
        const int* r = red + k * 8;
        const int lx = min(r[0], r[4]), hx = max(r[1], r[5]), ly = min(r[2], r[6]), hy = max(r[3], r[7]);
        if (lx == INT_MAX || ((long long)hx - lx + W2 <= pw && (long long)hy - ly + W2 <= pw)) fits |= 1u << k;
    }
    return fits;
}

__device__ __forceinline__ void tmem_ld4_nowait(uint32_t taddr, float* v) {
    uint32_t* r = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x4.b32 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(tc::smem_u32(bar)) : "memory");
}
__host__ __device__ constexpr uint32_t instr_desc_n(uint32_t fmt, uint32_t n) {
    return (1u << 4) | (fmt << 7) | (fmt << 10) | ((n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

// One CTA of 64 threads per tile: the verdict (tile map byte).  Thousands of these CTAs are resident at once, so the verdicts
// of an all-incoherent field (random-init weights) cost ~7 us; the tile kernel's CTAs then exit on their first (one-byte) load
// instead of each loading its tile's 64 flow vectors and reducing them under a 100-KB-shared-memory residency of two per SM.
// (Tried: a compact work list appended with atomicAdd — the list comes out in the arrival order of ~3000 concurrently
// resident verdict CTAs, i.e. scrambled over all batch elements, and the tile kernel lost its L2 locality: 0.52 -> 0.75 ms.)
__global__ void __launch_bounds__(64)
tile_verdict_kernel(const GatherParams p, int pw) {
    __shared__ int red[8];
    bool valid; int x0, y0, X0, Y0; float wx0, wx1, wy0, wy1, gx, gy;
    const bool fits = tile_window_fits(p, blockIdx.z, blockIdx.y * kTileSide, blockIdx.x * kTileSide, pw, red,
                                       valid, x0, y0, wx0, wx1, wy0, wy1, gx, gy, X0, Y0);
    if (threadIdx.x == 0) {
        const int tile = (blockIdx.z * gridDim.y + blockIdx.y) * gridDim.x + blockIdx.x;
        p.tile_flags[tile] = fits ? 1 : 0;
    }
}

template <typename T, int R>
__global__ void __launch_bounds__(256, 2)
local_corr_tile_kernel(const GatherParams p, const __grid_constant__ TileMaps maps, int tiles_x, int tiles_y) {
    using G = TileGeom<R>;
    static_assert(sizeof(T) == 2, "16-bit features");
    extern __shared__ unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[2], empty_bar[2], done_bar, free_bar;
    __shared__ uint32_t tmem_slot;
    __shared__ int red[8];
    __shared__ int s_x0[kTilePx], s_y0[kTilePx];
    __shared__ float s_w[kTilePx][4];
    __shared__ float2 s_g[kTilePx];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    if (!p.tile_flags[blockIdx.x]) return;                        // declined by tile_verdict_kernel: the per-pixel kernel's
    unsigned char* smem = smem_raw + ((1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u);
    float* Dwin = reinterpret_cast<float*>(smem + 2 * (size_t)G::STAGE);
    const int nkb = p.C / 64;                                     // even (C = 256, 512): every tile uses each ring stage an even
                                                                  // number of times, so the full / empty parities restart at 0
    constexpr uint32_t kFmt = std::is_same<T, __half>::value ? 0u : 1u;
    constexpr uint32_t idesc = instr_desc_n(kFmt, (uint32_t)G::NPAD);
    constexpr uint32_t kTmemCols = 256;

    if (warp == 0 && !(p.dbg & 32)) tc::tmem_alloc(&tmem_slot, kTmemCols);
    if (tid == 0) {
        for (int s = 0; s < 2; ++s) { tc::mbar_init(&full_bar[s], 1); tc::mbar_init(&empty_bar[s], 1); }
        tc::mbar_init(&done_bar, 1);
        tc::mbar_init(&free_bar, 64);
        tc::fence_barrier_init();
        tma::prefetch_map(&maps.f0);
        tma::prefetch_map(&maps.f1);
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    const uint32_t tmem_base = tmem_slot;
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1 = reinterpret_cast<const T*>(p.f1);
    T* dout = reinterpret_cast<T*>(p.d);
    uint32_t done_base = 0, free_base = 0;                        // phases of done_bar / free_bar consumed by earlier tiles

    // one tile per CTA, raster order (x fastest, then y, then batch element): neighbouring tiles run together and share
    // their overlapping patches in the L2.  (The loop form is kept for experiments with persistent grids.)
    for (int tile = blockIdx.x; tile == (int)blockIdx.x; tile += gridDim.x) {
    const int b = tile / (tiles_x * tiles_y), trem = tile - b * tiles_x * tiles_y;
    const int ty0 = (trem / tiles_x) * kTileSide, tx0 = (trem % tiles_x) * kTileSide;
    bool valid; int x0, y0, X0, Y0; float wx0, wx1, wy0, wy1, gx, gy;
    int rows_needed = 0;
    tile_window_fits(p, b, ty0, tx0, G::PW, red, valid, x0, y0, wx0, wx1, wy0, wy1, gx, gy, X0, Y0, &rows_needed);
    // row bands of the patch actually touched by the tile's windows: a smooth flow needs 2 of the 3
    const int npass = min(G::NPASS, (rows_needed + G::ROWS - 1) / G::ROWS);
    if (tid < kTilePx) {
        s_x0[tid] = valid ? x0 : INT_MIN; s_y0[tid] = y0;
        s_w[tid][0] = wx0; s_w[tid][1] = wx1; s_w[tid][2] = wy0; s_w[tid][3] = wy1;
        s_g[tid] = make_float2(gx, gy);
    }
    __syncthreads();

    if (p.dbg & 16) {
    } else if (warp == 0) {
        if (lane == 0) {                                         // ---- TMA producer ----
            for (int it = 0; it < ((p.dbg & 2) ? 0 : npass * nkb); ++it) {
                const int s = it & 1, pass = it / nkb, kb = it - pass * nkb;
                if (it >= 2) tc::mbar_wait(&empty_bar[s], (uint32_t)(((it >> 1) - 1) & 1));
                unsigned char* sa = smem + (size_t)s * G::STAGE;
                tma::mbar_expect_tx(&full_bar[s], (uint32_t)(kTilePx * 128 + G::NCOL * 128));
                tma::load_4d(sa, &maps.f0, &full_bar[s], kb * 64, tx0, ty0, b);
                tma::load_4d(sa + G::AOFF, &maps.f1, &full_bar[s], kb * 64, X0, Y0 + pass * G::ROWS, b);
            }
        }
        __syncwarp();
    } else if (warp == 1) {
        if (lane == 0) {                                         // ---- MMA issuer ----
            if (p.dbg & 2) {                                     // timing aid: no main loop (the products are garbage)
                for (int pass = 0; pass < npass; ++pass) {
                    if (pass) tc::mbar_wait(&free_bar, (free_base + (uint32_t)(pass - 1)) & 1u);
                    tc::umma_commit(&done_bar);
                }
            }
            for (int it = 0; it < ((p.dbg & 2) ? 0 : npass * nkb); ++it) {
                const int s = it & 1, pass = it / nkb, kb = it - pass * nkb;
                if (pass > 0 && kb == 0) { tc::mbar_wait(&free_bar, (free_base + (uint32_t)(pass - 1)) & 1u); tc::tc_fence_after(); }   // previous band drained from TMEM
                tc::mbar_wait(&full_bar[s], (uint32_t)((it >> 1) & 1));
                tc::tc_fence_after();
                const uint32_t a0 = tc::smem_u32(smem + (size_t)s * G::STAGE), b0 = a0 + G::AOFF;
#pragma unroll
                for (int kk = 0; kk < 4; ++kk)
                    tc::umma_f16(tmem_base, tc::smem_desc_sw128(a0 + kk * 32), tc::smem_desc_sw128(b0 + kk * 32), idesc, (kb | kk) ? 1u : 0u);
                tc::umma_commit(&empty_bar[s]);
                if (kb == nkb - 1) tc::umma_commit(&done_bar);
            }
        }
        __syncwarp();
    }
    if (!(p.dbg & 16) && warp >= 2) {                             // ---- streaming segments: warps 2..7 (4, 5: before their extraction) ----
        // 16 pixels per warp, items = (pixel, 16-byte vector) pairs spread over the lanes and processed FOUR AT A TIME: all
        // the loads of a group (4 x f0 or 16 x f1 taps) are issued before any is used — one L2 round trip per group instead of
        // one per pixel (the first version handled a pixel at a time: 2/3 of the CTA's life was these four warps waiting)
        const int sw = warp - 2;                                    // 0..5: pixels [m_lo, m_hi): 11, 11, 11, 11, 10, 10
        const int m_lo = sw * 11 - (sw > 4 ? 1 : 0), m_hi = m_lo + (sw < 4 ? 11 : 10);
        const int nvec = p.C / 8;
        const int nitems = (p.dbg & 1) ? 0 : (m_hi - m_lo) * nvec;
        const T* f1b = f1 + (long long)b * p.h1 * p.w1 * p.C;
        const long long rowpix0 = (long long)b * p.h;
        auto pixel_of = [&](int m, long long& pix) -> bool {
            if (s_x0[m] == INT_MIN) return false;
            pix = (rowpix0 + ty0 + (m >> 3)) * p.w + tx0 + (m & 7);
            return true;
        };
        if (p.off_x >= 0) {
            for (int q0 = lane; q0 < nitems; q0 += 128) {
                uint4 v[4]; T* dst[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int q = q0 + 32 * u;
                    dst[u] = nullptr;
                    long long pix;
                    if (q < nitems && pixel_of(m_lo + q / nvec, pix)) {
                        const int vv = q % nvec;
                        v[u] = ldg16(f0 + pix * p.C + vv * 8);
                        dst[u] = dout + pix * p.dch + p.off_x + vv * 8;
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) if (dst[u]) *reinterpret_cast<uint4*>(dst[u]) = v[u];
            }
        }
        if (p.off_xhat >= 0) {
            for (int q0 = lane; q0 < nitems; q0 += 128) {
                uint4 tap[4][4]; float wgt[4][4]; T* dst[4];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int q = q0 + 32 * u;
                    dst[u] = nullptr;
                    long long pix;
                    const int m = m_lo + (q < nitems ? q / nvec : 0);
                    if (q < nitems && pixel_of(m, pix)) {
                        const int vv = q % nvec, px0 = s_x0[m], py0 = s_y0[m];
                        dst[u] = dout + pix * p.dch + p.off_xhat + vv * 8;
#pragma unroll
                        for (int t = 0; t < 4; ++t) {
                            const int yy = py0 + (t >> 1), xx = px0 + (t & 1);
                            const bool in = yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1;
                            wgt[u][t] = in ? s_w[m][2 + (t >> 1)] * s_w[m][t & 1] : 0.f;
                            tap[u][t] = in ? ldg16(f1b + ((long long)yy * p.w1 + xx) * p.C + vv * 8) : make_uint4(0u, 0u, 0u, 0u);
                        }
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    if (!dst[u]) continue;
                    float xh[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) xh[e] = 0.f;
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        float f[8];                       // out-of-image taps: weight 0 x value 0 leaves the sum unchanged
                        unpack16<T>(tap[u][t], f);
#pragma unroll
                        for (int e = 0; e < 8; ++e) xh[e] = fmaf(wgt[u][t], f[e], xh[e]);
                    }
                    *reinterpret_cast<uint4*>(dst[u]) = pack16<T>(xh);
                }
            }
        }
        for (int m = m_lo; m < m_hi; ++m) {
            long long pix;
            if (!pixel_of(m, pix)) continue;
            const int y = ty0 + (m >> 3), x = tx0 + (m & 7);
            T* row = dout + pix * p.dch;
            if (p.off_emb >= 0 && p.E > 0) {
                const float dx = p.disp_scale * (s_g[m].x - pixel_center(x, p.w));
                const float dy = p.disp_scale * (s_g[m].y - pixel_center(y, p.h));
                for (int e = lane; e < p.E; e += 32)
                    row[p.off_emb + e] = Elem<T>::from_f(fmaf(__ldg(p.disp_w + 2 * e), dx,
                                                              fmaf(__ldg(p.disp_w + 2 * e + 1), dy, __ldg(p.disp_b + e))));
            }
            if (lane < p.dch - p.covered) row[p.covered + lane] = Elem<T>::from_f(0.f);
        }
    }
    if (!(p.dbg & 16) && (warp == 4 || warp == 5)) {              // ---- TMEM lanes 0..63 -> window products ----
        const int m = (warp - 4) * 32 + lane;
        const int ox = s_x0[m] == INT_MIN ? 0 : s_x0[m] - p.r - X0, oy = s_y0[m] - p.r - Y0;
        const bool live = s_x0[m] != INT_MIN;
        float* Dm = Dwin + m * G::DSTRIDE;
        const uint32_t tlane = tmem_base + ((uint32_t)((warp - 4) * 32) << 16);
        for (int pass = 0; pass < npass; ++pass) {
            tc::mbar_wait(&done_bar, (done_base + (uint32_t)pass) & 1u);
            tc::tc_fence_after();
            for (int qy = 0; qy < ((p.dbg & 4) ? 0 : G::ROWS); ++qy) {
                constexpr int NL = (G::PW + 3) / 4;
                float v[NL * 4];
#pragma unroll
                for (int q = 0; q < NL; ++q) tmem_ld4_nowait(tlane + (uint32_t)(qy * G::PW + 4 * q), v + 4 * q);
                tmem_wait_ld();
                const int j = pass * G::ROWS + qy - oy;
                if (live && (unsigned)j < (unsigned)G::W2) {
                    float* dst = Dm + j * G::W2 - ox;
#pragma unroll
                    for (int t = 0; t < G::PW; ++t)
                        if ((unsigned)(t - ox) < (unsigned)G::W2) dst[t] = v[t];
                }
            }
            tc::tc_fence_before();
            if (pass + 1 < npass) mbar_arrive(&free_bar);
        }
    }
    tc::tc_fence_before();
    __syncthreads();
    tc::tc_fence_after();
    done_base += (uint32_t)npass; free_base += (uint32_t)(npass - 1);
    // ---- blend: corr[ky][kx] = wy0 (wx0 D[ky][kx] + wx1 D[ky][kx+1]) + wy1 (wx0 D[ky+1][kx] + wx1 D[ky+1][kx+1]) ----
    if (p.off_corr >= 0 && !(p.dbg & 8)) {
        const float inv_sqrt_c = rsqrtf((float)p.C);
        for (int idx = tid; idx < kTilePx * G::K; idx += 256) {
            const int m = idx / G::K, k = idx - m * G::K;
            if (s_x0[m] == INT_MIN) continue;
            const int ky = k / G::K1, kx = k - ky * G::K1;
            const float* D0 = Dwin + m * G::DSTRIDE + ky * G::W2 + kx;
            const float a0 = s_w[m][0], a1 = s_w[m][1], c0 = s_w[m][2], c1 = s_w[m][3];
            const float c = c0 * (a0 * D0[0] + a1 * D0[1]) + c1 * (a0 * D0[G::W2] + a1 * D0[G::W2 + 1]);
            const int y = ty0 + (m >> 3), x = tx0 + (m & 7);
            dout[(((long long)b * p.h + y) * p.w + x) * p.dch + p.off_corr + k] = Elem<T>::from_f(c * inv_sqrt_c);
        }
    }
    __syncthreads();                                              // Dwin / s_* are rewritten by the next tile
    }   // tile
    tc::tc_fence_before();
    __syncthreads();
    if (warp == 0 && !(p.dbg & 32)) tc::tmem_dealloc(tmem_base, kTmemCols);
}

}  // namespace roma
