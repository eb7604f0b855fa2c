// K2: fused bilinear warp (grid_sample) + local correlation + displacement embedding, written
// directly into the channels-last ConvRefiner input buffer.
//
// Reference behaviour being replaced:
//   romatch/models/matcher.py:109-132 and romatch/utils/local_correlation.py:4-48.
//
// Algorithm (local correlation): the (2r+1)^2 window offsets are exact integer pixel steps
// (local_correlation.py:30-39: linspace(-2r/h, 2r/h, 2r+1) has step 2/h == 1 px), so every window
// tap shares the centre's bilinear fractional weights.  Per output pixel we therefore compute the
// (2r+2)^2 integer-neighbour dot products D[j][i] = f0 . f1[y0+j-r, x0+i-r] once and blend
//   corr[ky][kx] = wy0*(wx0*D[ky][kx] + wx1*D[ky][kx+1]) + wy1*(wx0*D[ky+1][kx] + wx1*D[ky+1][kx+1]).
// x_hat is the same 4-tap blend of the four centre neighbours' channel vectors.  Out-of-image
// neighbours contribute 0 (padding_mode="zeros").
//
// Data layout: features channels-last so one pixel's C-vector is a single contiguous burst; one
// warp per output pixel, lanes over 16-byte channel vectors, fp32 accumulation, butterfly
// transpose-reduction (1 shuffle per neighbour), results staged in shared memory per CTA tile and
// written back as ONE contiguous, 16-byte-vectorised span (the rows of `d` are odd-sized, e.g.
// 1377 channels, so per-row vector stores are impossible but the tile span is contiguous).
#include <atomic>
#include "common.cuh"

namespace roma {

constexpr int kTileSide = 8, kTilePx = 64;      // shared-patch tiles of local_corr_tile.cuh
constexpr int kScoresMaxTokens = 4096;          // r = 7: full-correlation GEMM up to 64 x 64 maps (workspace B * n^2 * 4 bytes)

struct GatherParams {
    const void* f0; const void* f1; const float* flow; roma_strides4 fs;
    const float* disp_w; const float* disp_b; float disp_scale;
    void* d; int dch; int off_x, off_xhat, off_emb, off_corr;
    int B, C, h, w, h1, w1, E, r;
    long long npix;
    int tile_px;     // pixels per CTA tile (local-corr kernel)
    int covered;     // channels written by the enabled segments; [covered, dch) is zero-filled pitch padding
    int dbg;         // timing aid for local_corr_tile_kernel (bit 0: no streaming segments, 1: no main loop, 2: no extraction, 3: no blend)
    const float* scores;         // r = 7 on small maps: the FULL correlation f0 . f1^T (B, h*w, h1*w1) fp32 from one tensor-core GEMM; the
                                 // per-pixel kernel then reads its (2r+2)^2 products from its row of it instead of computing them
    unsigned char* tile_flags;   // per 8 x 8 tile: 1 = computed by local_corr_tile_kernel (the per-pixel kernel skips its pixels); or null
    int tile_pw;     // > 0: 8 x 8 pixel tiles (grid = tiles_x, tiles_y, B); the per-pixel kernel SKIPS tiles whose windows fit
                     // a tile_pw x tile_pw patch (they are computed by local_corr_tile_kernel, local_corr_tile.cuh)
};

__device__ __forceinline__ void pixel_coords(const GatherParams& p, long long pix, int& b, int& y, int& x,
                                             float& gx, float& gy) {
    const int hw = p.h * p.w;
    b = (int)(pix / hw);
    const int rem = (int)(pix - (long long)b * hw);
    y = rem / p.w;
    x = rem - y * p.w;
    const float* f = p.flow + (long long)b * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
    gx = __ldg(f);
    gy = __ldg(f + p.fs.c);
}

__device__ __forceinline__ void sample_coords(float gx, float gy, int w1, int h1, int& x0, int& y0,
                                              float& wx0, float& wx1, float& wy0, float& wy1) {
    float ix = unnormalize(gx, w1), iy = unnormalize(gy, h1);
    // keep floor() in int range; anything this far out is all-zero padding anyway
    ix = fminf(fmaxf(ix, -1048576.f), 1048576.f);
    iy = fminf(fmaxf(iy, -1048576.f), 1048576.f);
    const float fx = floorf(ix), fy = floorf(iy);
    x0 = (int)fx; y0 = (int)fy;
    wx0 = (fx + 1.f) - ix; wx1 = ix - fx;
    wy0 = (fy + 1.f) - iy; wy1 = iy - fy;
}

// ------------------------------------------------------------------------------------------------
// Kernel A: r > 0.  One warp per pixel; CTA tile = tile_px consecutive (b,y,x)-flattened pixels.
// ------------------------------------------------------------------------------------------------
template <typename T, int VPL>
__global__ void __launch_bounds__(256)
local_corr_kernel(const GatherParams p) {
    constexpr int kVec = Elem<T>::kVec;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int nwarps = blockDim.x >> 5;
    const int W2 = 2 * p.r + 2, NN = W2 * W2, K1 = 2 * p.r + 1, K = K1 * K1;
    const int nvec = p.C / kVec;

    const long long tile_start = (long long)blockIdx.x * p.tile_px;
    int tile_n = p.tile_px;
    if (tile_start + tile_n > p.npix) tile_n = (int)(p.npix - tile_start);

    // staging laid out exactly like the global span, with the same 16-byte misalignment
    T* gspan = reinterpret_cast<T*>(p.d) + tile_start * p.dch;
    const unsigned mis = (unsigned)(reinterpret_cast<uintptr_t>(gspan) & 15u);
    T* stage = reinterpret_cast<T*>(smem_raw + mis);
    const size_t stage_bytes = ((size_t)p.tile_px * p.dch * sizeof(T) + 16 + 15) & ~(size_t)15;
    float* Dsm = reinterpret_cast<float*>(smem_raw + stage_bytes) + warp * (NN + 1);

    const float inv_sqrt_c = rsqrtf((float)p.C);
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1 = reinterpret_cast<const T*>(p.f1);

    for (int lp = warp; lp < tile_n; lp += nwarps) {
        const long long pix = tile_start + lp;
        int b, y, x; float gx, gy;
        pixel_coords(p, pix, b, y, x, gx, gy);
        int x0, y0; float wx0, wx1, wy0, wy1;
        sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
        T* row = stage + (size_t)lp * p.dch;

        // f0 channel vectors of this pixel, kept in registers as fp32
        float a[VPL][kVec];
        const T* f0p = f0 + pix * p.C;
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
            const int v = lane + 32 * q;
            if (v < nvec) {
                const uint4 raw = ldg16(f0p + v * kVec);
                unpack16<T>(raw, a[q]);
                if (p.off_x >= 0) {
                    const T* e = reinterpret_cast<const T*>(&raw);
#pragma unroll
                    for (int t = 0; t < kVec; ++t) row[p.off_x + v * kVec + t] = e[t];
                }
            } else {
#pragma unroll
                for (int t = 0; t < kVec; ++t) a[q][t] = 0.f;
            }
        }
        float xh[VPL][kVec];
#pragma unroll
        for (int q = 0; q < VPL; ++q)
#pragma unroll
            for (int t = 0; t < kVec; ++t) xh[q][t] = 0.f;

        const T* f1b = f1 + (long long)b * p.h1 * p.w1 * p.C;
        int j = 0, i = 0;   // neighbour (row, col) in the (2r+2)^2 grid
        for (int n0 = 0; n0 < NN; n0 += 32) {
            float part[32];
#pragma unroll
            for (int t = 0; t < 32; ++t) {
                float acc = 0.f;
                if (n0 + t < NN) {
                    const int yy = y0 + j - p.r, xx = x0 + i - p.r;
                    if (yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1) {
                        const T* src = f1b + ((long long)yy * p.w1 + xx) * p.C;
                        const bool cy = (j == p.r) | (j == p.r + 1), cx = (i == p.r) | (i == p.r + 1);
                        const float wgt = (j == p.r ? wy0 : wy1) * (i == p.r ? wx0 : wx1);
#pragma unroll
                        for (int q = 0; q < VPL; ++q) {
                            const int v = lane + 32 * q;
                            if (v < nvec) {
                                float u[kVec];
                                unpack16<T>(ldg16(src + v * kVec), u);
#pragma unroll
                                for (int e = 0; e < kVec; ++e) acc = fmaf(a[q][e], u[e], acc);
                                if (cy && cx) {
#pragma unroll
                                    for (int e = 0; e < kVec; ++e) xh[q][e] = fmaf(wgt, u[e], xh[q][e]);
                                }
                            }
                        }
                    }
                    if (++i == W2) { i = 0; ++j; }
                }
                part[t] = acc;
            }
            // butterfly transpose-reduce: afterwards lane L holds the full sum of neighbour n0+L
#pragma unroll
            for (int s = 16; s >= 1; s >>= 1) {
                const bool up = (lane & s) != 0;
#pragma unroll
                for (int t = 0; t < s; ++t) {
                    const float send = up ? part[t] : part[t + s];
                    const float keep = up ? part[t + s] : part[t];
                    part[t] = keep + __shfl_xor_sync(0xffffffffu, send, s);
                }
            }
            if (n0 + lane < NN) Dsm[n0 + lane] = part[0];
        }
        __syncwarp();

        if (p.off_xhat >= 0) {
#pragma unroll
            for (int q = 0; q < VPL; ++q) {
                const int v = lane + 32 * q;
                if (v < nvec) {
#pragma unroll
                    for (int t = 0; t < kVec; ++t) row[p.off_xhat + v * kVec + t] = Elem<T>::from_f(xh[q][t]);
                }
            }
        }
        if (p.off_corr >= 0) {
            for (int k = lane; k < K; k += 32) {
                const int ky = k / K1, kx = k - ky * K1;
                const float* D0 = Dsm + ky * W2 + kx;
                const float c = wy0 * (wx0 * D0[0] + wx1 * D0[1]) + wy1 * (wx0 * D0[W2] + wx1 * D0[W2 + 1]);
                row[p.off_corr + k] = Elem<T>::from_f(c * inv_sqrt_c);
            }
        }
        if (p.off_emb >= 0 && p.E > 0) {
            const float dx = p.disp_scale * (gx - pixel_center(x, p.w));
            const float dy = p.disp_scale * (gy - pixel_center(y, p.h));
            for (int e = lane; e < p.E; e += 32)
                row[p.off_emb + e] = Elem<T>::from_f(fmaf(__ldg(p.disp_w + 2 * e), dx,
                                                          fmaf(__ldg(p.disp_w + 2 * e + 1), dy, __ldg(p.disp_b + e))));
        }
        if (lane < p.dch - p.covered) row[p.covered + lane] = Elem<T>::from_f(0.f);
        __syncwarp();
    }
    __syncthreads();

    // contiguous span write-back: scalar head, 16-byte body, scalar tail
    const size_t total = (size_t)tile_n * p.dch;          // elements
    const size_t head = mis ? (16 - mis) / sizeof(T) : 0;  // elements until 16-byte alignment
    const size_t head_n = head < total ? head : total;
    for (size_t e = threadIdx.x; e < head_n; e += blockDim.x) gspan[e] = stage[e];
    if (total > head_n) {
        const size_t body_elems = ((total - head_n) / kVec) * kVec;
        const uint4* s4 = reinterpret_cast<const uint4*>(stage + head_n);
        uint4* g4 = reinterpret_cast<uint4*>(gspan + head_n);
        const size_t nv = body_elems / kVec;
        for (size_t v = threadIdx.x; v < nv; v += blockDim.x) g4[v] = s4[v];
        for (size_t e = head_n + body_elems + threadIdx.x; e < total; e += blockDim.x) gspan[e] = stage[e];
    }
}

// ------------------------------------------------------------------------------------------------
// Kernel A2: r > 0, 16-byte aligned rows (C = 64*VPL channels, every segment offset and the row pitch a multiple
// of 8 elements — what ops.refiner_inputs produces with its padded pitch).  No output staging: all segments are
// stored straight from registers with 16-byte (features) / 2-byte coalesced (corr, emb) stores, which frees the
// shared memory and the registers the staged version spent on the transpose-reduction.
//
// One warp per pixel, four neighbour rows in flight per warp: lane = (group g = lane/8, gl = lane%8); group g
// takes neighbour n0+g, its 8 lanes cover the row as VPL interleaved 16-byte vectors (8 lanes x 16 B = one
// 128-byte line per load instruction), i.e. VPL independent 16-byte loads per lane per step, then a 3-step
// shuffle reduction inside the group.  f0 stays in registers as fp32.
// ------------------------------------------------------------------------------------------------
template <typename T, int VPL>
__global__ void __launch_bounds__(256, 2)
local_corr_v2_kernel(const GatherParams p) {
    constexpr int kVec = Elem<T>::kVec;          // 8 (16-bit types only)
    static_assert(kVec == 8, "local_corr_v2 is for 16-bit feature types");
    constexpr int kMaxNN = 256;                  // (2r+2)^2 for r <= 7
    __shared__ float Dall[8][kMaxNN + 4];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int grp = lane >> 3, gl = lane & 7;
    const int W2 = 2 * p.r + 2, NN = W2 * W2, K1 = 2 * p.r + 1, K = K1 * K1;
    float* Dsm = Dall[warp];
    const float inv_sqrt_c = rsqrtf((float)p.C);
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1 = reinterpret_cast<const T*>(p.f1);
    T* dout = reinterpret_cast<T*>(p.d);
    const int covered = p.covered;

    const long long tile_start = (long long)blockIdx.x * p.tile_px;
    int tile_n = p.tile_px;
    if (tile_start + tile_n > p.npix) tile_n = (int)(p.npix - tile_start);

    for (int lp = warp; lp < tile_n; lp += 8) {
        const long long pix = tile_start + lp;
        int b, y, x; float gx, gy;
        pixel_coords(p, pix, b, y, x, gx, gy);
        int x0, y0; float wx0, wx1, wy0, wy1;
        sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
        T* row = dout + pix * p.dch;

        // f0 slice of this lane: vectors gl + 8*q
        float a[VPL][8];
        const T* f0p = f0 + pix * p.C;
#pragma unroll
        for (int q = 0; q < VPL; ++q) {
            const uint4 raw = ldg16(f0p + (gl + 8 * q) * 8);
            unpack16<T>(raw, a[q]);
            if (p.off_x >= 0 && (q & 3) == grp) *reinterpret_cast<uint4*>(row + p.off_x + (gl + 8 * q) * 8) = raw;
        }
        const T* f1b = f1 + (long long)b * p.h1 * p.w1 * p.C;
        for (int n0 = 0; n0 < NN; n0 += 4) {
            const int n = n0 + grp;
            const int j = n / W2, i = n - j * W2;
            const int yy = y0 + j - p.r, xx = x0 + i - p.r;
            float acc = 0.f;
            if (n < NN && yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1) {
                const T* src = f1b + ((long long)yy * p.w1 + xx) * p.C + gl * 8;
                uint4 u[VPL];
#pragma unroll
                for (int q = 0; q < VPL; ++q) u[q] = ldg16(src + 64 * q);
#pragma unroll
                for (int q = 0; q < VPL; ++q) {
                    float f[8];
                    unpack16<T>(u[q], f);
#pragma unroll
                    for (int e = 0; e < 8; ++e) acc = fmaf(a[q][e], f[e], acc);
                }
            }
            acc += __shfl_xor_sync(0xffffffffu, acc, 1);
            acc += __shfl_xor_sync(0xffffffffu, acc, 2);
            acc += __shfl_xor_sync(0xffffffffu, acc, 4);
            if (gl == 0 && n < NN) Dsm[n] = acc;
        }
        __syncwarp();
        // x_hat: 4-tap blend of the centre neighbours, lanes over the row's 16-byte vectors (lines just loaded: L1 hits)
        if (p.off_xhat >= 0) {
#pragma unroll
            for (int q = 0; q < (VPL + 3) / 4; ++q) {
                const int v = lane + 32 * q;
                if (v < 8 * VPL) {
                    float xh[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) xh[e] = 0.f;
#pragma unroll
                    for (int t = 0; t < 4; ++t) {
                        const int yy = y0 + (t >> 1), xx = x0 + (t & 1);
                        if (yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1) {
                            const float wgt = ((t >> 1) ? wy1 : wy0) * ((t & 1) ? wx1 : wx0);
                            float f[8];
                            unpack16<T>(ldg16(f1b + ((long long)yy * p.w1 + xx) * p.C + v * 8), f);
#pragma unroll
                            for (int e = 0; e < 8; ++e) xh[e] = fmaf(wgt, f[e], xh[e]);
                        }
                    }
                    *reinterpret_cast<uint4*>(row + p.off_xhat + v * 8) = pack16<T>(xh);
                }
            }
        }
        if (p.off_corr >= 0) {
            for (int k = lane; k < K; k += 32) {
                const int ky = k / K1, kx = k - ky * K1;
                const float* D0 = Dsm + ky * W2 + kx;
                const float c = wy0 * (wx0 * D0[0] + wx1 * D0[1]) + wy1 * (wx0 * D0[W2] + wx1 * D0[W2 + 1]);
                row[p.off_corr + k] = Elem<T>::from_f(c * inv_sqrt_c);
            }
        }
        if (p.off_emb >= 0 && p.E > 0) {
            const float dx = p.disp_scale * (gx - pixel_center(x, p.w));
            const float dy = p.disp_scale * (gy - pixel_center(y, p.h));
            for (int e = lane; e < p.E; e += 32)
                row[p.off_emb + e] = Elem<T>::from_f(fmaf(__ldg(p.disp_w + 2 * e), dx,
                                                          fmaf(__ldg(p.disp_w + 2 * e + 1), dy, __ldg(p.disp_b + e))));
        }
        if (lane < p.dch - covered) row[covered + lane] = Elem<T>::from_f(0.f);      // zero the pitch padding (< 8 channels)
        __syncwarp();
    }
}

// ------------------------------------------------------------------------------------------------
// Kernel A3 (default for the aligned case): the neighbour dot products on warp-level tensor-core MMAs.
//
// local_corr_v2 is bound by memory-level parallelism, not by bytes: a lane keeps 8 independent 16-byte loads in
// flight and then spends ~130 issue slots per step on bf16 unpacks, FMAs and shuffles, with f0 held as 64 fp32
// registers (coherent and incoherent flows take the same time, tools/kernel_bench.py).  Here one
// mma.sync.m16n8k16 computes 16 rows x 16 channels with fp32 accumulation, fragments filled straight from 16-byte
// global loads (the K dimension is a sum, so its order is free: the same channel permutation is applied to f0,
// which sits in shared memory, 1 KB per warp, one LDS.128 per two MMAs).  Per 8 neighbours x C channels a lane
// issues 2 C/64 independent 16-byte loads (16 for C = 512: twice v2's bytes in flight at half its registers) and
// as many HMMAs; no unpack, no FMA, one shuffle.  Row layout chosen for the L1: see the loop.  Out-of-image
// neighbours load zeros.
// ------------------------------------------------------------------------------------------------
template <typename T> struct Mma16816;
template <> struct Mma16816<__nv_bfloat16> {
    static __device__ __forceinline__ void run(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
};
template <> struct Mma16816<__half> {
    static __device__ __forceinline__ void run(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                     : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    }
};
template <> struct Mma16816<float> {
    static __device__ __forceinline__ void run(float (&)[4], uint32_t, uint32_t, uint32_t, uint32_t, uint32_t, uint32_t) {}
};

// One pixel by one warp: the body shared by the flat per-pixel kernel and the incoherent tiles of local_corr_tile_kernel.
// Dsm: (2r+2)^2 floats, f0w: 64 * VPL elements (16-byte aligned), both private to the warp.
// R > 0: the radius is a compile-time constant (2, 3: the loop over the (2r+2)^2 neighbours is unrolled, so the loads of
// several groups of 8 neighbours are in flight together and the / W2, % W2 index arithmetic folds away); R = 0: p.r at run time.
template <typename T, int VPL, int R = 0>
__device__ __forceinline__ void local_corr_px(const GatherParams& p, long long pix, int lane, float* Dsm, T* f0w) {
    static_assert(sizeof(T) == 2, "local_corr_mma is for 16-bit feature types");
    const int g = lane >> 2, t4 = lane & 3;
    const int r_ = R > 0 ? R : p.r;
    const int W2 = 2 * r_ + 2, NN = W2 * W2, K1 = 2 * r_ + 1, K = K1 * K1;
    const float inv_sqrt_c = rsqrtf((float)p.C);
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1 = reinterpret_cast<const T*>(p.f1);
    T* dout = reinterpret_cast<T*>(p.d);
    const int covered = p.covered;
    int b, y, x; float gx, gy;
    pixel_coords(p, pix, b, y, x, gx, gy);
    int x0, y0; float wx0, wx1, wy0, wy1;
    sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
    T* row = dout + pix * p.dch;

    // f0 row -> shared (B operand of every MMA of this pixel) and -> the x segment of d
    const T* f0p = f0 + pix * p.C;
#pragma unroll
    for (int q = 0; q < (8 * VPL + 31) / 32; ++q) {
        const int v = lane + 32 * q;
        if (v < 8 * VPL) {
            const uint4 raw = ldg16(f0p + v * 8);
            *reinterpret_cast<uint4*>(f0w + v * 8) = raw;
            if (p.off_x >= 0) *reinterpret_cast<uint4*>(row + p.off_x + v * 8) = raw;
        }
    }
    __syncwarp();
    const T* f1b = f1 + (long long)b * p.h1 * p.w1 * p.C;
    if (p.scores) {
        const int n1 = p.h1 * p.w1;
        const float* srow = p.scores + ((long long)b * p.h * p.w + (long long)y * p.w + x) * n1;
        for (int n = lane; n < NN; n += 32) {
            const int j = n / W2, i = n - j * W2;
            const int yy = y0 + j - r_, xx = x0 + i - r_;
            Dsm[n] = (yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1) ? __ldg(srow + yy * p.w1 + xx) : 0.f;
        }
    } else
#pragma unroll
    for (int n0 = 0; n0 < (R > 0 ? (2 * R + 2) * (2 * R + 2) : NN); n0 += 8) {
        // A-fragment rows 2j and 2j+1 are the two 64-byte halves of a 128-byte piece of neighbour n0 + j (rows
        // 8.. : neighbour n0 + 4 + j), so the 8 lanes of a quarter-warp read ONE full line per 16-byte load
        // (4 L1 wavefronts per instruction, not 8).  The B fragment carries the matching f0 channels in the
        // column of the same parity: D[g][g & 1] is the partial dot product of row g, and the two halves meet
        // with one shuffle.
        // Addresses are formed unconditionally from CLAMPED coordinates (always inside the map) and only the loads are
        // predicated: a conditional 64-bit address computation compiles to a branch per neighbour, which cut every group of
        // 8 neighbours into separate basic blocks and kept the next group's loads from being scheduled under this group's MMAs.
        const T* src[2];
        bool okn[2];
#pragma unroll
        for (int hr = 0; hr < 2; ++hr) {
            const int n = n0 + (g >> 1) + 4 * hr;
            const int j = n / W2, i = n - j * W2;
            const int yy = y0 + j - r_, xx = x0 + i - r_;
            okn[hr] = n < NN && yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1;
            const int yc = min(max(yy, 0), p.h1 - 1), xc = min(max(xx, 0), p.w1 - 1);
            src[hr] = f1b + ((long long)yc * p.w1 + xc) * p.C + 32 * (g & 1) + t4 * 8;
        }
        float dA[4] = {0.f, 0.f, 0.f, 0.f}, dB[4] = {0.f, 0.f, 0.f, 0.f};
        constexpr int KCH = VPL;      // (C = 512 in two chunks of 256 at 3 CTAs / SM: 0.74 -> 0.78 ms at s8' — that shape is L2-bound)
#pragma unroll
        for (int k0 = 0; k0 < VPL; k0 += KCH) {
            uint4 u[2][KCH];
#pragma unroll
            for (int kb = 0; kb < KCH; ++kb) {
#pragma unroll
                for (int hr = 0; hr < 2; ++hr)
                    u[hr][kb] = okn[hr] ? ldg16(src[hr] + 64 * (k0 + kb)) : make_uint4(0u, 0u, 0u, 0u);
            }
#pragma unroll
            for (int kb = 0; kb < KCH; ++kb) {
                const uint4 fb = *reinterpret_cast<const uint4*>(f0w + 64 * (k0 + kb) + 32 * (g & 1) + t4 * 8);
                Mma16816<T>::run(dA, u[0][kb].x, u[1][kb].x, u[0][kb].y, u[1][kb].y, fb.x, fb.y);
                Mma16816<T>::run(dB, u[0][kb].z, u[1][kb].z, u[0][kb].w, u[1][kb].w, fb.z, fb.w);
            }
        }
        float vlo = (g & 1) ? dA[1] + dB[1] : dA[0] + dB[0];
        float vhi = (g & 1) ? dA[3] + dB[3] : dA[2] + dB[2];
        vlo += __shfl_xor_sync(0xffffffffu, vlo, 4);
        vhi += __shfl_xor_sync(0xffffffffu, vhi, 4);
        if (t4 == 0 && (g & 1) == 0) {
            if (n0 + (g >> 1) < NN) Dsm[n0 + (g >> 1)] = vlo;
            if (n0 + 4 + (g >> 1) < NN) Dsm[n0 + 4 + (g >> 1)] = vhi;
        }
    }
    __syncwarp();
    // x_hat: 4-tap blend of the centre neighbours, lanes over the row's 16-byte vectors (lines just loaded: L1 hits)
    if (p.off_xhat >= 0) {
#pragma unroll
        for (int q = 0; q < (VPL + 3) / 4; ++q) {
            const int v = lane + 32 * q;
            if (v < 8 * VPL) {
                float xh[8];
#pragma unroll
                for (int e = 0; e < 8; ++e) xh[e] = 0.f;
                // the four taps are loaded together from clamped (always valid) addresses; a tap outside the map is left out
                // by predicating its FMAs — same values as skipping it, without a branch (and a load latency) per tap
                uint4 raw[4];
                bool okt[4];
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const int yy = y0 + (t >> 1), xx = x0 + (t & 1);
                    okt[t] = yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1;
                    const int yc = min(max(yy, 0), p.h1 - 1), xc = min(max(xx, 0), p.w1 - 1);
                    raw[t] = ldg16(f1b + ((long long)yc * p.w1 + xc) * p.C + v * 8);
                }
#pragma unroll
                for (int t = 0; t < 4; ++t) {
                    const float wgt = ((t >> 1) ? wy1 : wy0) * ((t & 1) ? wx1 : wx0);
                    float f[8];
                    unpack16<T>(raw[t], f);
#pragma unroll
                    for (int e = 0; e < 8; ++e) xh[e] = okt[t] ? fmaf(wgt, f[e], xh[e]) : xh[e];
                }
                *reinterpret_cast<uint4*>(row + p.off_xhat + v * 8) = pack16<T>(xh);
            }
        }
    }
    if (p.off_corr >= 0) {
        for (int k = lane; k < K; k += 32) {
            const int ky = k / K1, kx = k - ky * K1;
            const float* D0 = Dsm + ky * W2 + kx;
            const float c = wy0 * (wx0 * D0[0] + wx1 * D0[1]) + wy1 * (wx0 * D0[W2] + wx1 * D0[W2 + 1]);
            row[p.off_corr + k] = Elem<T>::from_f(c * inv_sqrt_c);
        }
    }
    if (p.off_emb >= 0 && p.E > 0) {
        const float dx = p.disp_scale * (gx - pixel_center(x, p.w));
        const float dy = p.disp_scale * (gy - pixel_center(y, p.h));
        for (int e = lane; e < p.E; e += 32)
            row[p.off_emb + e] = Elem<T>::from_f(fmaf(__ldg(p.disp_w + 2 * e), dx,
                                                      fmaf(__ldg(p.disp_w + 2 * e + 1), dy, __ldg(p.disp_b + e))));
    }
    if (lane < p.dch - covered) row[covered + lane] = Elem<T>::from_f(0.f);      // zero the pitch padding (< 8 channels)
    __syncwarp();
}

template <typename T, int VPL, int R = 0>
// C = 256, r = 2 (scale 4) is latency-bound: 4 CTAs / SM at 64 registers beat 2 at 115 (1.53 -> 1.28 ms at s4', incoherent)
__global__ void __launch_bounds__(256, (R == 2 && VPL <= 4) ? 4 : 2)
local_corr_mma_kernel(const GatherParams p) {
    constexpr int kMaxNN = 256;                  // (2r+2)^2 for r <= 7
    __shared__ float Dall[8][kMaxNN + 4];
    __shared__ __align__(16) T f0s[8][64 * VPL];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long tile_start = (long long)blockIdx.x * p.tile_px;
    int tile_n = p.tile_px;
    if (tile_start + tile_n > p.npix) tile_n = (int)(p.npix - tile_start);
    // pixels of the tiles the shared-patch kernel took are skipped: the (up to 4) tile-map bytes of this warp's pixels are
    // fetched together up front (lane i <-> pixel warp + 8 i), not one dependent load per pixel
    unsigned skip = 0u;
    if (p.tile_flags) {
        const int tiles_x = (p.w + kTileSide - 1) / kTileSide, tiles_y = (p.h + kTileSide - 1) / kTileSide;
        const int lp = warp + 8 * lane;
        bool taken = false;
        if (lp < tile_n) {
            const long long pix = tile_start + lp;
            const int hw = p.h * p.w;
            const int b = (int)(pix / hw), rem = (int)(pix - (long long)b * hw);
            const int y = rem / p.w, x = rem - y * p.w;
            taken = p.tile_flags[((long long)b * tiles_y + (y >> 3)) * tiles_x + (x >> 3)] != 0;
        }
        skip = __ballot_sync(0xffffffffu, taken);
    }
    for (int lp = warp; lp < tile_n; lp += 8) {
        const long long pix = tile_start + lp;
        if ((skip >> (lp >> 3)) & 1u) continue;
        local_corr_px<T, VPL, R>(p, pix, lane, Dall[warp], f0s[warp]);
    }
}

}  // namespace roma
#include "local_corr_tile.cuh"
namespace roma {

// ------------------------------------------------------------------------------------------------
// Kernel C: r == 0 with rows that cannot be vectorised (C*sizeof(T) not a multiple of 16, e.g. the 9-channel
// finest scale).  One thread per pixel computes its whole output row (x copy, x_hat, emb) into a shared-memory
// stage; the CTA then writes the contiguous span of its 256 rows with 16-byte stores.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
gather_rowstage_kernel(const GatherParams p) {
    extern __shared__ __align__(16) unsigned char smem_raw[];
    constexpr int kVec = Elem<T>::kVec;
    const long long tile_start = (long long)blockIdx.x * 256;
    int tile_n = 256;
    if (tile_start + tile_n > p.npix) tile_n = (int)(p.npix - tile_start);
    T* gspan = reinterpret_cast<T*>(p.d) + tile_start * p.dch;
    const unsigned mis = (unsigned)(reinterpret_cast<uintptr_t>(gspan) & 15u);
    T* stage = reinterpret_cast<T*>(smem_raw + mis);
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1 = reinterpret_cast<const T*>(p.f1);
    const int lp = threadIdx.x;
    if (lp < tile_n) {
        const long long pix = tile_start + lp;
        int b, y, x; float gx, gy;
        pixel_coords(p, pix, b, y, x, gx, gy);
        T* row = stage + (size_t)lp * p.dch;
        if (p.off_x >= 0) {
            const T* src = f0 + pix * p.C;
            for (int c = 0; c < p.C; ++c) row[p.off_x + c] = src[c];
        }
        if (p.off_xhat >= 0) {
            int x0, y0; float wx0, wx1, wy0, wy1;
            sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
            const T* f1b = f1 + (long long)b * p.h1 * p.w1 * p.C;
            const T* tap[4]; float wgt[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
                const int yy = y0 + (t >> 1), xx = x0 + (t & 1);
                const bool ok = yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1;
                tap[t] = ok ? f1b + ((long long)yy * p.w1 + xx) * p.C : f1b;
                wgt[t] = ok ? ((t >> 1) ? wy1 : wy0) * ((t & 1) ? wx1 : wx0) : 0.f;
            }
            for (int c = 0; c < p.C; ++c) {
                float acc = 0.f;
#pragma unroll
                for (int t = 0; t < 4; ++t) acc = fmaf(wgt[t], Elem<T>::to_f(tap[t][c]), acc);
                row[p.off_xhat + c] = Elem<T>::from_f(acc);
            }
        }
        if (p.off_emb >= 0 && p.E > 0) {
            const float dx = p.disp_scale * (gx - pixel_center(x, p.w));
            const float dy = p.disp_scale * (gy - pixel_center(y, p.h));
            for (int e = 0; e < p.E; ++e)
                row[p.off_emb + e] = Elem<T>::from_f(fmaf(__ldg(p.disp_w + 2 * e), dx, fmaf(__ldg(p.disp_w + 2 * e + 1), dy, __ldg(p.disp_b + e))));
        }
        for (int c = p.covered; c < p.dch; ++c) row[c] = Elem<T>::from_f(0.f);
    }
    __syncthreads();
    const size_t total = (size_t)tile_n * p.dch;
    const size_t head = mis ? (16 - mis) / sizeof(T) : 0;
    const size_t head_n = head < total ? head : total;
    for (size_t e = threadIdx.x; e < head_n; e += blockDim.x) gspan[e] = stage[e];
    if (total > head_n) {
        const size_t body_elems = ((total - head_n) / kVec) * kVec;
        const uint4* s4 = reinterpret_cast<const uint4*>(stage + head_n);
        uint4* g4 = reinterpret_cast<uint4*>(gspan + head_n);
        const size_t nv = body_elems / kVec;
        for (size_t v = threadIdx.x; v < nv; v += blockDim.x) g4[v] = s4[v];
        for (size_t e = head_n + body_elems + threadIdx.x; e < total; e += blockDim.x) gspan[e] = stage[e];
    }
}

// ------------------------------------------------------------------------------------------------
// Kernel B: r == 0 (gather-only scales 2 and 1, Tiny-RoMa).  One thread per (pixel, VEC-channel chunk).
// ------------------------------------------------------------------------------------------------
// IDX32: every index fits 32 bits (nthreads < 2^31): unsigned 32-bit divisions instead of the emulated 64-bit ones.
template <typename T, int VEC, bool IDX32>
__global__ void __launch_bounds__(256)
gather_only_kernel(const GatherParams p, long long nthreads, int cpp) {
    long long pix;
    int ch, b, y, x; float gx, gy;
    if (IDX32) {
        const unsigned idx = blockIdx.x * blockDim.x + threadIdx.x;
        if (idx >= (unsigned)nthreads) return;
        const unsigned pix32 = idx / (unsigned)cpp;
        ch = (int)(idx - pix32 * (unsigned)cpp) * VEC;
        const unsigned hw = (unsigned)(p.h * p.w);
        const unsigned bb = pix32 / hw, rem = pix32 - bb * hw;
        const unsigned yy = rem / (unsigned)p.w;
        b = (int)bb; y = (int)yy; x = (int)(rem - yy * (unsigned)p.w);
        const float* f = p.flow + (long long)b * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
        gx = __ldg(f);
        gy = __ldg(f + p.fs.c);
        pix = pix32;
    } else {
        const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
        if (idx >= nthreads) return;
        pix = idx / cpp;
        ch = (int)(idx - pix * cpp) * VEC;
        pixel_coords(p, pix, b, y, x, gx, gy);
    }
    T* out = reinterpret_cast<T*>(p.d) + pix * p.dch + ch;
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1 = reinterpret_cast<const T*>(p.f1);

    if (p.off_x >= 0 && ch >= p.off_x && ch < p.off_x + p.C) {
        const T* src = f0 + pix * p.C + (ch - p.off_x);
        if (VEC > 1) *reinterpret_cast<uint4*>(out) = ldg16(src);
        else *out = *src;
        return;
    }
    if (p.off_xhat >= 0 && ch >= p.off_xhat && ch < p.off_xhat + p.C) {
        const int c = ch - p.off_xhat;
        int x0, y0; float wx0, wx1, wy0, wy1;
        sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
        float acc[VEC];
#pragma unroll
        for (int e = 0; e < VEC; ++e) acc[e] = 0.f;
        const T* f1b = f1 + (long long)b * p.h1 * p.w1 * p.C + c;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int yy = y0 + (t >> 1), xx = x0 + (t & 1);
            if (yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1) {
                const float wgt = ((t >> 1) ? wy1 : wy0) * ((t & 1) ? wx1 : wx0);
                const T* src = f1b + ((long long)yy * p.w1 + xx) * p.C;
                if (VEC > 1) {
                    float u[VEC > 1 ? VEC : 8];
                    unpack16<T>(ldg16(src), u);
#pragma unroll
                    for (int e = 0; e < VEC; ++e) acc[e] = fmaf(wgt, u[e], acc[e]);
                } else {
                    acc[0] = fmaf(wgt, Elem<T>::to_f(*src), acc[0]);
                }
            }
        }
        if (VEC > 1) *reinterpret_cast<uint4*>(out) = pack16<T>(acc);
        else *out = Elem<T>::from_f(acc[0]);
        return;
    }
    if (p.off_emb >= 0 && ch >= p.off_emb && ch < p.off_emb + p.E) {
        const int e0 = ch - p.off_emb;
        const float dx = p.disp_scale * (gx - pixel_center(x, p.w));
        const float dy = p.disp_scale * (gy - pixel_center(y, p.h));
        float v[VEC];
#pragma unroll
        for (int e = 0; e < VEC; ++e)
            v[e] = fmaf(__ldg(p.disp_w + 2 * (e0 + e)), dx, fmaf(__ldg(p.disp_w + 2 * (e0 + e) + 1), dy, __ldg(p.disp_b + e0 + e)));
        if (VEC > 1) *reinterpret_cast<uint4*>(out) = pack16<T>(v);
        else *out = Elem<T>::from_f(v[0]);
        return;
    }
    if (ch >= p.covered) {      // pitch padding
        if (VEC > 1) *reinterpret_cast<uint4*>(out) = make_uint4(0u, 0u, 0u, 0u);
        else *out = Elem<T>::from_f(0.f);
    }
}

// 0 (default): shared-patch tcgen05 tiles for coherent tiles (r = 2, 3) + per-pixel tensor-core kernel for the rest;
// 1: SIMT FMA dot products (v2); 2: per-pixel tensor-core kernel for every pixel (the round-1 default)
static std::atomic<int> g_local_corr_path{0};
static std::atomic<int> g_tile_dbg{0};          // roma_local_corr_set_path(path | dbg << 8): phases of the tile kernel switched off (timing only)

// Shared-patch path: both kernels over the same (tiles_x, tiles_y, B) grid; every tile is taken by exactly one of them.
// ROMA_ERR_UNSUPPORTED (silently, no message) when the feature maps cannot be described to TMA -> caller falls back.
template <typename T, int R>
static int launch_tiled_r(GatherParams p, cudaStream_t st, int dtype) {
    using G = TileGeom<R>;
    TileMaps maps;
    const uint64_t e = sizeof(T);
    const uint64_t d0[4] = {(uint64_t)p.C, (uint64_t)p.w, (uint64_t)p.h, (uint64_t)p.B};
    const uint64_t s0[3] = {(uint64_t)p.C * e, (uint64_t)p.w * p.C * e, (uint64_t)p.h * p.w * p.C * e};
    const uint32_t b0[4] = {64u, (uint32_t)kTileSide, (uint32_t)kTileSide, 1u};
    const uint64_t d1[4] = {(uint64_t)p.C, (uint64_t)p.w1, (uint64_t)p.h1, (uint64_t)p.B};
    const uint64_t s1[3] = {(uint64_t)p.C * e, (uint64_t)p.w1 * p.C * e, (uint64_t)p.h1 * p.w1 * p.C * e};
    const uint32_t b1[4] = {64u, (uint32_t)G::PW, (uint32_t)G::ROWS, 1u};
    if (!tma::encode_fn()) return ROMA_ERR_UNSUPPORTED;
    if (tma::encode_tiled(&maps.f0, dtype, 4, p.f0, d0, s0, b0, CU_TENSOR_MAP_SWIZZLE_128B) != 0 ||
        tma::encode_tiled(&maps.f1, dtype, 4, p.f1, d1, s1, b1, CU_TENSOR_MAP_SWIZZLE_128B) != 0) return ROMA_ERR_UNSUPPORTED;
    static DeviceOnce::Flags configured;
    if (DeviceOnce::pending(configured)) {
        cudaError_t err = cudaFuncSetAttribute(local_corr_tile_kernel<T, R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)G::SMEM);
        if (err != cudaSuccess) return fail((int)err, "roma_refiner_gather: smem opt-in (%zu B): %s", G::SMEM, cudaGetErrorString(err));
        DeviceOnce::done(configured);
    }
    p.tile_pw = G::PW;
    p.dbg = g_tile_dbg.load();
    const int tiles_x = cdiv(p.w, kTileSide), tiles_y = cdiv(p.h, kTileSide);
    const long long ntiles = (long long)p.B * tiles_x * tiles_y;
    if (ntiles >= (1ll << 31)) return ROMA_ERR_UNSUPPORTED;
    const dim3 grid((unsigned)tiles_x, (unsigned)tiles_y, (unsigned)p.B);
    tile_verdict_kernel<<<grid, 64, 0, st>>>(p, G::PW);
    local_corr_tile_kernel<T, R><<<(unsigned)ntiles, 256, G::SMEM, st>>>(p, maps, tiles_x, tiles_y);
    p.tile_px = 32;
    const int gpx = cdiv(p.npix, p.tile_px);
    switch (p.C) {
        case 512: local_corr_mma_kernel<T, 8, R><<<gpx, 256, 0, st>>>(p); break;
        default:  local_corr_mma_kernel<T, 4, R><<<gpx, 256, 0, st>>>(p); break;
    }
    return check_launch("roma_refiner_gather(tile_verdict + local_corr_tile + local_corr_mma)", 3);
}

// ------------------------------------------------------------------------------------------------
// Kernel B2: r == 0, 16-bit features, C = 8 * NJ with NJ a power of two (the 64-channel scale-2 visits, Tiny-RoMa's
// coarse features).  Kernel B spends one thread per 16-byte chunk of d and repeats the pixel decode, the two flow loads
// and sample_coords for each of a pixel's 18 chunks: measured issue-bound (issue slots 69 % busy at 21 % of the DRAM
// peak).  Here a thread owns chunk j of the x segment AND of the x_hat segment of its pixel (+ a share of the embedding /
// padding chunks), so the per-pixel work is done C / 8 times instead of dch / 8 times, the four taps are loaded before
// any is used, and the blend runs as packed fp32 FMAs (fma.rn.f32x2: the two halves of a 32-bit word ride one
// instruction).  Same products and order per element as Kernel B: bit-identical output.
// ------------------------------------------------------------------------------------------------
template <typename T> __device__ __forceinline__ float2 unpack_pair(uint32_t w);
template <> __device__ __forceinline__ float2 unpack_pair<__nv_bfloat16>(uint32_t w) {
    return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}
template <> __device__ __forceinline__ float2 unpack_pair<__half>(uint32_t w) {
    return __half22float2(*reinterpret_cast<const __half2*>(&w));
}
template <> __device__ __forceinline__ float2 unpack_pair<float>(uint32_t) { return make_float2(0.f, 0.f); }

template <typename T, int NJ>
__global__ void __launch_bounds__(256)
gather_only_v2_kernel(const GatherParams p, unsigned nthreads) {
    static_assert(sizeof(T) == 2, "16-bit features");
    const unsigned idx = blockIdx.x * 256u + threadIdx.x;
    if (idx >= nthreads) return;
    const unsigned pix = idx / NJ;
    const int j = (int)(idx % NJ);
    const unsigned hw = (unsigned)(p.h * p.w);
    const unsigned bb = pix / hw, rem = pix - bb * hw;
    const unsigned yy0 = rem / (unsigned)p.w;
    const int b = (int)bb, y = (int)yy0, x = (int)(rem - yy0 * (unsigned)p.w);
    const float* fl = p.flow + (long long)b * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
    const float gx = __ldg(fl), gy = __ldg(fl + p.fs.c);
    T* row = reinterpret_cast<T*>(p.d) + (long long)pix * p.dch;
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1 = reinterpret_cast<const T*>(p.f1);

    uint4 xv = make_uint4(0u, 0u, 0u, 0u);
    if (p.off_x >= 0) xv = ldg16(f0 + (long long)pix * p.C + j * 8);
    if (p.off_xhat >= 0) {
        int x0, y0; float wx0, wx1, wy0, wy1;
        sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
        const T* f1b = f1 + (long long)b * p.h1 * p.w1 * p.C + j * 8;
        uint4 tap[4]; float wgt[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int yy = y0 + (t >> 1), xx = x0 + (t & 1);
            const bool in = yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1;
            wgt[t] = in ? ((t >> 1) ? wy1 : wy0) * ((t & 1) ? wx1 : wx0) : 0.f;
            tap[t] = in ? ldg16(f1b + ((long long)yy * p.w1 + xx) * p.C) : make_uint4(0u, 0u, 0u, 0u);
        }
        float2 acc[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[i] = make_float2(0.f, 0.f);
#pragma unroll
        for (int t = 0; t < 4; ++t) {                       // out-of-image taps: weight 0 x value 0 leaves the sums unchanged
            const uint32_t w[4] = {tap[t].x, tap[t].y, tap[t].z, tap[t].w};
            const float2 ww = make_float2(wgt[t], wgt[t]);
#pragma unroll
            for (int i = 0; i < 4; ++i) acc[i] = __ffma2_rn(ww, unpack_pair<T>(w[i]), acc[i]);
        }
        const float o[8] = {acc[0].x, acc[0].y, acc[1].x, acc[1].y, acc[2].x, acc[2].y, acc[3].x, acc[3].y};
        *reinterpret_cast<uint4*>(row + p.off_xhat + j * 8) = pack16<T>(o);
    }
    if (p.off_x >= 0) *reinterpret_cast<uint4*>(row + p.off_x + j * 8) = xv;
    const int nE = (p.off_emb >= 0) ? p.E >> 3 : 0, nP = (p.dch - p.covered) >> 3;
    for (int e = j; e < nE + nP; e += NJ) {
        if (e < nE) {
            const float dx = p.disp_scale * (gx - pixel_center(x, p.w));
            const float dy = p.disp_scale * (gy - pixel_center(y, p.h));
            float v[8];
#pragma unroll
            for (int k = 0; k < 8; ++k)
                v[k] = fmaf(__ldg(p.disp_w + 2 * (8 * e + k)), dx, fmaf(__ldg(p.disp_w + 2 * (8 * e + k) + 1), dy, __ldg(p.disp_b + 8 * e + k)));
            *reinterpret_cast<uint4*>(row + p.off_emb + 8 * e) = pack16<T>(v);
        } else {
            *reinterpret_cast<uint4*>(row + p.covered + 8 * (e - nE)) = make_uint4(0u, 0u, 0u, 0u);
        }
    }
}

// ------------------------------------------------------------------------------------------------
// Kernel C2: the finest visits of roma_outdoor — C = 9 channels (proj["1"], 18-byte rows: only 2-byte aligned), E = 6,
// d row = [x(9) | x_hat(9) | emb(6)] = 24 elements = 48 bytes.  Kernel C issues 45 two-byte loads per pixel and stages
// rows through shared memory; here a thread reads each 18-byte row as FIVE aligned 32-bit words (funnel-shifted when the
// row starts on an odd half-word), blends channel pairs with packed fp32 FMAs and writes its 48-byte row with three
// 16-byte stores.  Same products and order per element: bit-identical output.
// ------------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ void load_row9(const T* ptr, uint32_t (&w)[5]) {
    const uintptr_t a = reinterpret_cast<uintptr_t>(ptr);
    const uint32_t* q = reinterpret_cast<const uint32_t*>(a & ~(uintptr_t)3);
    uint32_t r[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) r[i] = __ldg(q + i);
    if (a & 2) {
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = __funnelshift_r(r[i], r[i + 1], 16);
        w[4] = r[4] >> 16;
    } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) w[i] = r[i];
        w[4] = r[4] & 0xffffu;
    }
}

template <typename T>
__global__ void __launch_bounds__(256)
gather_c9_kernel(const GatherParams p, unsigned npix) {
    static_assert(sizeof(T) == 2, "16-bit features");
    const unsigned pix = blockIdx.x * 256u + threadIdx.x;
    if (pix >= npix) return;
    const unsigned hw = (unsigned)(p.h * p.w);
    const unsigned bb = pix / hw, rem = pix - bb * hw;
    const unsigned yy0 = rem / (unsigned)p.w;
    const int b = (int)bb, y = (int)yy0, x = (int)(rem - yy0 * (unsigned)p.w);
    const float* fl = p.flow + (long long)b * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
    const float gx = __ldg(fl), gy = __ldg(fl + p.fs.c);
    const T* f0 = reinterpret_cast<const T*>(p.f0);
    const T* f1b = reinterpret_cast<const T*>(p.f1) + (long long)b * p.h1 * p.w1 * 9;
    int x0, y0; float wx0, wx1, wy0, wy1;
    sample_coords(gx, gy, p.w1, p.h1, x0, y0, wx0, wx1, wy0, wy1);
    uint32_t xw[5], tap[4][5]; float wgt[4];
    load_row9<T>(f0 + (long long)pix * 9, xw);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int yy = y0 + (t >> 1), xx = x0 + (t & 1);
        const bool in = yy >= 0 && yy < p.h1 && xx >= 0 && xx < p.w1;
        wgt[t] = in ? ((t >> 1) ? wy1 : wy0) * ((t & 1) ? wx1 : wx0) : 0.f;
        if (in) load_row9<T>(f1b + ((long long)yy * p.w1 + xx) * 9, tap[t]);
        else {
#pragma unroll
            for (int i = 0; i < 5; ++i) tap[t][i] = 0u;
        }
    }
    float2 acc[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) acc[i] = make_float2(0.f, 0.f);
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const float2 ww = make_float2(wgt[t], wgt[t]);
#pragma unroll
        for (int i = 0; i < 5; ++i) acc[i] = __ffma2_rn(ww, unpack_pair<T>(tap[t][i]), acc[i]);
    }
    const float dx = p.disp_scale * (gx - pixel_center(x, p.w));
    const float dy = p.disp_scale * (gy - pixel_center(y, p.h));
    unsigned short o[24];
#pragma unroll
    for (int i = 0; i < 9; ++i) o[i] = (unsigned short)((i & 1) ? (xw[i >> 1] >> 16) : (xw[i >> 1] & 0xffffu));
#pragma unroll
    for (int i = 0; i < 9; ++i) {
        const T v = Elem<T>::from_f((i & 1) ? acc[i >> 1].y : acc[i >> 1].x);
        o[9 + i] = *reinterpret_cast<const unsigned short*>(&v);
    }
#pragma unroll
    for (int e = 0; e < 6; ++e) {
        const T v = Elem<T>::from_f(fmaf(__ldg(p.disp_w + 2 * e), dx, fmaf(__ldg(p.disp_w + 2 * e + 1), dy, __ldg(p.disp_b + e))));
        o[18 + e] = *reinterpret_cast<const unsigned short*>(&v);
    }
    uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.d) + (long long)pix * 24);
#pragma unroll
    for (int k = 0; k < 3; ++k)
        dst[k] = make_uint4((uint32_t)o[8 * k] | ((uint32_t)o[8 * k + 1] << 16), (uint32_t)o[8 * k + 2] | ((uint32_t)o[8 * k + 3] << 16),
                            (uint32_t)o[8 * k + 4] | ((uint32_t)o[8 * k + 5] << 16), (uint32_t)o[8 * k + 6] | ((uint32_t)o[8 * k + 7] << 16));
}

template <typename T>
static bool launch_gather_v2(const GatherParams& p, cudaStream_t st) {
    if constexpr (sizeof(T) == 2) {
        const int nj = p.C / 8;
        if (p.C % 8 || (nj & (nj - 1)) || nj > 64 || p.npix * nj >= (1ll << 32) || p.npix >= (1ll << 31)) return false;
        const unsigned n = (unsigned)(p.npix * nj);
        const unsigned grid = (n + 255u) / 256u;
        switch (nj) {
            case 1:  gather_only_v2_kernel<T, 1><<<grid, 256, 0, st>>>(p, n); break;
            case 2:  gather_only_v2_kernel<T, 2><<<grid, 256, 0, st>>>(p, n); break;
            case 4:  gather_only_v2_kernel<T, 4><<<grid, 256, 0, st>>>(p, n); break;
            case 8:  gather_only_v2_kernel<T, 8><<<grid, 256, 0, st>>>(p, n); break;
            case 16: gather_only_v2_kernel<T, 16><<<grid, 256, 0, st>>>(p, n); break;
            case 32: gather_only_v2_kernel<T, 32><<<grid, 256, 0, st>>>(p, n); break;
            default: gather_only_v2_kernel<T, 64><<<grid, 256, 0, st>>>(p, n); break;
        }
        return true;
    }
    return false;
}

template <typename T>
static int launch_tiled(GatherParams p, cudaStream_t st) {
    if constexpr (sizeof(T) == 2) {
        const int dtype = std::is_same<T, __half>::value ? ROMA_F16 : ROMA_BF16;
        if (p.B > 65535 || cdiv(p.h, kTileSide) > 65535) return ROMA_ERR_UNSUPPORTED;
        return p.r == 2 ? launch_tiled_r<T, 2>(p, st, dtype) : launch_tiled_r<T, 3>(p, st, dtype);
    }
    return ROMA_ERR_UNSUPPORTED;
}

template <typename T>
static int launch_gather(GatherParams p, cudaStream_t st) {
    constexpr int kVec = Elem<T>::kVec;
    const bool ptr16 = ((reinterpret_cast<uintptr_t>(p.f0) | reinterpret_cast<uintptr_t>(p.f1)) & 15u) == 0;
    if (p.r > 0) {
        ROMA_REQUIRE(p.C % kVec == 0 && ptr16, ROMA_ERR_ALIGNMENT,
                     "roma_refiner_gather: local correlation needs C*sizeof(T) %% 16 == 0 and 16-byte aligned features (C=%d)", p.C);
        const int nvec = p.C / kVec;
        ROMA_REQUIRE(nvec <= 128, ROMA_ERR_UNSUPPORTED, "roma_refiner_gather: C=%d too large for the local-correlation kernel", p.C);
        ROMA_REQUIRE(p.r <= 15, ROMA_ERR_UNSUPPORTED, "roma_refiner_gather: radius %d > 15", p.r);
        // aligned fast path: 16-bit features, C = 64*VPL, every segment and the pitch 16-byte aligned
        auto al8 = [](int off) { return off < 0 || (off & 7) == 0; };
        const bool d16 = (reinterpret_cast<uintptr_t>(p.d) & 15u) == 0;
        if (sizeof(T) == 2 && d16 && (p.dch & 7) == 0 && al8(p.off_x) && al8(p.off_xhat) && p.r <= 7 &&
            (p.C == 512 || p.C == 256 || p.C == 128 || p.C == 64)) {
            p.tile_px = 32;
            const int grid = cdiv(p.npix, p.tile_px);
            if constexpr (sizeof(T) == 2) {
                if (p.scores && g_local_corr_path.load() == 0) {
                    // r = 7 on a small map: all (h w) x (h1 w1) products as ONE batched tcgen05 GEMM (TMA-fed, fp32 accumulate of the same
                    // 16-bit products), whatever the flow; the per-pixel kernel below only gathers and blends
                    tc::Launch L{};
                    const int n0 = p.h * p.w, n1 = p.h1 * p.w1;
                    L.p[0] = tc::make_problem(p.f0, p.f1, const_cast<float*>(p.scores), n0, n1, p.C, p.C, p.C, n1,
                                              (long long)n0 * p.C, (long long)n1 * p.C, (long long)n0 * n1, tc::EPI_STORE);
                    L.nprob = 1; L.tiles0 = L.p[0].tiles_m * L.p[0].tiles_n;
                    const int rc = tc::launch_gemm_nt_16<T>(L, p.B, st, "roma_refiner_gather(full correlation)",
                                                            std::is_same<T, __half>::value ? ROMA_F16 : ROMA_BF16);
                    if (rc) return rc;
                } else p.scores = nullptr;
                if (g_local_corr_path.load() == 0 && p.tile_flags && (p.r == 2 || p.r == 3) && (p.C == 512 || p.C == 256) && p.h == p.h1 && p.w == p.w1 && p.off_corr >= 0) {
                    // shared-patch tiles (coherent flow) + per-pixel kernel for the tiles that decline (local_corr_tile.cuh)
                    const int rc = launch_tiled<T>(p, st);
                    if (rc != ROMA_ERR_UNSUPPORTED) return rc;
                }
                p.tile_flags = nullptr;
                if (g_local_corr_path.load() == 0 || g_local_corr_path.load() == 2) {
                    if (p.r == 3 && p.C == 512 && !p.scores) { local_corr_mma_kernel<T, 8, 3><<<grid, 256, 0, st>>>(p); return check_launch("roma_refiner_gather(local_corr_mma)"); }
                    if (p.r == 2 && p.C == 256 && !p.scores) { local_corr_mma_kernel<T, 4, 2><<<grid, 256, 0, st>>>(p); return check_launch("roma_refiner_gather(local_corr_mma)"); }
                    switch (p.C) {
                        case 512: local_corr_mma_kernel<T, 8><<<grid, 256, 0, st>>>(p); break;
                        case 256: local_corr_mma_kernel<T, 4><<<grid, 256, 0, st>>>(p); break;
                        case 128: local_corr_mma_kernel<T, 2><<<grid, 256, 0, st>>>(p); break;
                        default:  local_corr_mma_kernel<T, 1><<<grid, 256, 0, st>>>(p); break;
                    }
                    return check_launch("roma_refiner_gather(local_corr_mma)");
                }
                switch (p.C) {
                    case 512: local_corr_v2_kernel<T, 8><<<grid, 256, 0, st>>>(p); break;
                    case 256: local_corr_v2_kernel<T, 4><<<grid, 256, 0, st>>>(p); break;
                    case 128: local_corr_v2_kernel<T, 2><<<grid, 256, 0, st>>>(p); break;
                    default:  local_corr_v2_kernel<T, 1><<<grid, 256, 0, st>>>(p); break;
                }
            }
            return check_launch("roma_refiner_gather(local_corr_v2)");
        }
        const int W2 = 2 * p.r + 2, NN = W2 * W2;
        const int nwarps = 8;
        // tile size: keep the staging buffer under ~40 KB so several CTAs fit per SM
        int tile_px = (int)((40 * 1024) / ((size_t)p.dch * sizeof(T)));
        tile_px = tile_px < nwarps ? nwarps : (tile_px > 32 ? 32 : tile_px);
        p.tile_px = tile_px;
        const size_t stage_bytes = ((size_t)tile_px * p.dch * sizeof(T) + 16 + 15) & ~(size_t)15;
        const size_t smem = stage_bytes + (size_t)nwarps * (NN + 1) * sizeof(float);
        const int grid = cdiv(p.npix, tile_px);
        auto launch = [&](auto kern) -> int {
            if (smem > 48 * 1024) {
                cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
                if (e != cudaSuccess) return fail((int)e, "roma_refiner_gather: smem opt-in (%zu B): %s", smem, cudaGetErrorString(e));
            }
            kern<<<grid, nwarps * 32, smem, st>>>(p);
            return check_launch("roma_refiner_gather(local_corr)");
        };
        if (nvec <= 32) return launch(local_corr_kernel<T, 1>);
        if (nvec <= 64) return launch(local_corr_kernel<T, 2>);
        return launch(local_corr_kernel<T, 4>);
    }
    // gather-only
    const bool d16 = (reinterpret_cast<uintptr_t>(p.d) & 15u) == 0;
    auto seg_ok = [&](int off, int len) { return off < 0 || (off % kVec == 0 && len % kVec == 0); };
    const bool vec = ptr16 && d16 && p.C % kVec == 0 && p.dch % kVec == 0 &&
                     seg_ok(p.off_x, p.C) && seg_ok(p.off_xhat, p.C) && seg_ok(p.off_emb, p.E);
    auto seg_pad_ok = p.covered % kVec == 0;
    if (vec && seg_pad_ok && g_local_corr_path.load() != 1 && launch_gather_v2<T>(p, st)) {
        return check_launch("roma_refiner_gather(gather_only_v2)");
    } else if (vec && seg_pad_ok) {
        const int cpp = p.dch / kVec;
        const long long n = p.npix * cpp;
        if (n < (1ll << 31)) gather_only_kernel<T, kVec, true><<<cdiv(n, 256), 256, 0, st>>>(p, n, cpp);
        else gather_only_kernel<T, kVec, false><<<cdiv(n, 256), 256, 0, st>>>(p, n, cpp);
    } else if (sizeof(T) == 2 && p.C == 9 && p.E == 6 && p.dch == 24 && p.off_x == 0 && p.off_xhat == 9 && p.off_emb == 18 && d16 &&
               ((reinterpret_cast<uintptr_t>(p.f0) | reinterpret_cast<uintptr_t>(p.f1)) & 3u) == 0 && p.npix < (1ll << 31) &&
               g_local_corr_path.load() != 1) {
        // the last two bytes of the last aligned word of a row may lie past an odd-sized tensor: stay inside the allocation
        // (feature tensors here are batch * h * w * 9 halves: the total is even exactly when the pixel count is even)
        if constexpr (sizeof(T) == 2) {
            const bool even0 = ((p.npix) & 1) == 0, even1 = (((long long)p.B * p.h1 * p.w1) & 1) == 0;
            if (even0 && even1) {
                gather_c9_kernel<T><<<cdiv(p.npix, 256), 256, 0, st>>>(p, (unsigned)p.npix);
                return check_launch("roma_refiner_gather(gather_c9)");
            }
        }
        const size_t smem = (size_t)256 * p.dch * sizeof(T) + 32;
        gather_rowstage_kernel<T><<<cdiv(p.npix, 256), 256, smem, st>>>(p);
    } else if ((size_t)256 * p.dch * sizeof(T) + 32 <= 96 * 1024 && p.C <= 64) {
        // small, unalignable rows (the 9-channel finest scale): one thread per pixel + staged span write
        const size_t smem = (size_t)256 * p.dch * sizeof(T) + 32;
        if (smem > 48 * 1024) {
            cudaError_t e = cudaFuncSetAttribute(gather_rowstage_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
            if (e != cudaSuccess) return fail((int)e, "roma_refiner_gather: smem opt-in (%zu B): %s", smem, cudaGetErrorString(e));
        }
        gather_rowstage_kernel<T><<<cdiv(p.npix, 256), 256, smem, st>>>(p);
    } else {
        const long long n = p.npix * p.dch;
        if (n < (1ll << 31)) gather_only_kernel<T, 1, true><<<cdiv(n, 256), 256, 0, st>>>(p, n, p.dch);
        else gather_only_kernel<T, 1, false><<<cdiv(n, 256), 256, 0, st>>>(p, n, p.dch);
    }
    return check_launch("roma_refiner_gather(gather_only)");
}

}  // namespace roma

extern "C" int roma_local_corr_set_path(int path) {
    if (path < 0 || (path & 0xff) > 2 || (path >> 8) > 63) return roma::fail(ROMA_ERR_BAD_ARG, "roma_local_corr_set_path: path %d (0 default, 1 SIMT, 2 per-pixel tensor-core only)", path);
    roma::g_local_corr_path.store(path & 0xff);
    roma::g_tile_dbg.store((path >> 8) & 63);
    return ROMA_OK;
}

extern "C" size_t roma_refiner_gather_workspace_bytes(int B, int h, int w, int r) {
    if (B > 0 && h > 0 && w > 0 && r == 7 && (long long)h * w <= roma::kScoresMaxTokens)          // full correlation (B, hw, hw) fp32
        return ((size_t)B * h * w * h * w * sizeof(float) + 255) & ~(size_t)255;
    if (B <= 0 || h <= 0 || w <= 0 || (r != 2 && r != 3)) return 0;
    const size_t ntiles = (size_t)B * ((h + roma::kTileSide - 1) / roma::kTileSide) * ((w + roma::kTileSide - 1) / roma::kTileSide);
    return (ntiles + 255) & ~(size_t)255;                                                            // tile map
}

extern "C" int roma_refiner_gather(const void* f0, const void* f1, const float* flow, roma_strides4 fs,
                                   const float* disp_w, const float* disp_b, float disp_scale,
                                   void* d, int d_channels, int off_x, int off_xhat, int off_emb, int off_corr,
                                   int B, int C, int h, int w, int h1, int w1, int E, int r, int dtype,
                                   void* stream) {
    return roma_refiner_gather_ws(f0, f1, flow, fs, disp_w, disp_b, disp_scale, d, d_channels, off_x, off_xhat, off_emb, off_corr,
                                  B, C, h, w, h1, w1, E, r, dtype, nullptr, 0, stream);
}

extern "C" int roma_refiner_gather_ws(const void* f0, const void* f1, const float* flow, roma_strides4 fs,
                                      const float* disp_w, const float* disp_b, float disp_scale,
                                      void* d, int d_channels, int off_x, int off_xhat, int off_emb, int off_corr,
                                      int B, int C, int h, int w, int h1, int w1, int E, int r, int dtype,
                                      void* workspace, size_t workspace_bytes, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(f1 && flow && d, ROMA_ERR_BAD_ARG, "roma_refiner_gather: null pointer");
    ROMA_REQUIRE(B > 0 && C > 0 && h > 0 && w > 0 && h1 > 0 && w1 > 0 && r >= 0 && E >= 0 && d_channels > 0,
                 ROMA_ERR_BAD_ARG, "roma_refiner_gather: bad sizes B=%d C=%d h=%d w=%d h1=%d w1=%d E=%d r=%d", B, C, h, w, h1, w1, E, r);
    if (E == 0) off_emb = -1;
    if (r == 0) off_corr = -1;
    // the window offsets are linspace(-2r/h, 2r/h) in feature0's size (local_correlation.py:30-39): exact one-pixel steps in
    // feature1 only when the two maps have the same size — the shared-weight form below relies on that
    ROMA_REQUIRE(off_corr < 0 || (h == h1 && w == w1), ROMA_ERR_UNSUPPORTED,
                 "roma_refiner_gather: local correlation needs feature maps of one size (f0 %dx%d, f1 %dx%d)", h, w, h1, w1);
    ROMA_REQUIRE(!(off_emb >= 0) || (disp_w && disp_b), ROMA_ERR_BAD_ARG, "roma_refiner_gather: embedding requested without weights");
    ROMA_REQUIRE(f0 || (off_x < 0 && off_corr < 0), ROMA_ERR_BAD_ARG, "roma_refiner_gather: f0 is required for the copy / correlation segments");
    // the enabled segments must tile [0, d_channels) exactly (the tile span is written as a whole)
    const int K = (2 * r + 1) * (2 * r + 1);
    int offs[4] = {off_x, off_xhat, off_emb, off_corr};
    int lens[4] = {C, C, E, K};
    int covered = 0;
    for (int guard = 0; guard < 5; ++guard) {
        bool found = false;
        for (int s = 0; s < 4; ++s)
            if (offs[s] == covered) { covered += lens[s]; offs[s] = -2; found = true; break; }
        if (!found) break;
    }
    bool leftover = false;
    for (int s = 0; s < 4; ++s) leftover |= (offs[s] >= 0);
    ROMA_REQUIRE(covered <= d_channels && d_channels - covered < 32 && !leftover, ROMA_ERR_BAD_ARG,
                 "roma_refiner_gather: enabled segments must be contiguous from 0 and fill [0,%d) up to < 32 padding channels (covered %d)", d_channels, covered);

    GatherParams p;
    p.f0 = f0; p.f1 = f1; p.flow = flow; p.fs = fs;
    p.disp_w = disp_w; p.disp_b = disp_b; p.disp_scale = disp_scale;
    p.d = d; p.dch = d_channels; p.off_x = off_x; p.off_xhat = off_xhat; p.off_emb = off_emb; p.off_corr = off_corr;
    p.B = B; p.C = C; p.h = h; p.w = w; p.h1 = h1; p.w1 = w1; p.E = E; p.r = r;
    p.npix = (long long)B * h * w; p.tile_px = 0; p.covered = covered; p.tile_pw = 0; p.dbg = 0;
    const size_t need_ws = roma_refiner_gather_workspace_bytes(B, h, w, r);
    const bool have_ws = workspace && need_ws > 0 && workspace_bytes >= need_ws && (reinterpret_cast<uintptr_t>(workspace) & 15u) == 0;
    p.tile_flags = (have_ws && r != 7) ? reinterpret_cast<unsigned char*>(workspace) : nullptr;
    p.scores = (have_ws && r == 7 && h == h1 && w == w1 && off_corr >= 0 && dtype != ROMA_F32 && C % 8 == 0) ? reinterpret_cast<const float*>(workspace) : nullptr;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (dtype) {
        case ROMA_F32:  return launch_gather<float>(p, st);
        case ROMA_F16:  return launch_gather<__half>(p, st);
        case ROMA_BF16: return launch_gather<__nv_bfloat16>(p, st);
        default: return fail(ROMA_ERR_BAD_ARG, "roma_refiner_gather: unknown dtype %d", dtype);
    }
}
