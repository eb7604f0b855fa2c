// (f-2) ConvRefiner body: depthwise 5x5 conv + folded BatchNorm + ReLU in one pass, channels-last.
//
// Replaces (reference): the conv1 -> norm -> relu triple of every ConvRefiner block
//   romatch/models/matcher.py:77-103 (create_block: depthwise kernel_size=5, padding=2, BatchNorm2d, ReLU)
// as it runs at inference (BatchNorm in eval mode is an affine map; the host folds it into the weights/bias).
// In the reference this is four ATen launches per block (cudnn depthwise conv, bias add_, batch_norm, relu_),
// 81 blocks per 560->864 match; here it is one launch per block and one read + one write of the activation.
//
// Layout: x, out (B, h, w, Cp) channels-last, Cp even; weights fp32 (25, Cp) tap-major so that one tap of
// consecutive channels is contiguous; bias fp32 (Cp).  fp32 accumulation.
// Mapping: a thread owns one channel PAIR (one 32-bit word for 16-bit types), a run of 4 output columns and a
// strip of RY output rows.  It walks the RY+4 input rows of the strip once: each row (8 pixels, unpacked once)
// is scattered into the accumulators of the five output rows it touches (ring of 5 x 4 float2 accumulators,
// static slots, nothing is shifted).  The two channels of the pair ride one packed fma.rn.f32x2 (FFMA2): scalar
// 3-register FFMA issues at half rate on sm_100, so FFMA2 is what reaches the FP32 pipe's peak.  The kernel sits
// where the FP32 floor (25 MAC per output element) and the HBM floor (one read + one write of the activation)
// meet; lanes run over channel pairs, so a warp touches 128 contiguous bytes per pixel.
//
// Two input paths:
//   * TMA (default when the pixel stride is a multiple of 16 bytes — every padded refiner width): the block's
//     whole input tile is fetched by cp.async.bulk.tensor boxes, borders and channel tail zero-filled by the copy
//     engine;
//   * per-thread cp.async staging (any even channel count).
#include <atomic>
#include "common.cuh"
#include "tma.cuh"

namespace roma {

template <typename T> struct Pair;
template <> struct Pair<float> {
    using Raw = float2;
    __device__ static __forceinline__ Raw zero() { return make_float2(0.f, 0.f); }
    __device__ static __forceinline__ Raw load(const float* p) { return *reinterpret_cast<const float2*>(p); }
    __device__ static __forceinline__ float2 unpack(Raw r) { return r; }
    __device__ static __forceinline__ void store(float* p, float a, float b) { *reinterpret_cast<float2*>(p) = make_float2(a, b); }
    __device__ static __forceinline__ void store_relu(float* p, float a, float b) { store(p, fmaxf(a, 0.f), fmaxf(b, 0.f)); }
};
template <> struct Pair<__half> {
    using Raw = uint32_t;
    __device__ static __forceinline__ Raw zero() { return 0u; }
    __device__ static __forceinline__ Raw load(const __half* p) { return *reinterpret_cast<const uint32_t*>(p); }
    __device__ static __forceinline__ float2 unpack(Raw r) { return __half22float2(*reinterpret_cast<const __half2*>(&r)); }
    __device__ static __forceinline__ void store(__half* p, float a, float b) { *reinterpret_cast<__half2*>(p) = __floats2half2_rn(a, b); }
    __device__ static __forceinline__ void store_relu(__half* p, float a, float b) {        // ReLU rides the conversion
        uint32_t r;
        asm("cvt.rn.relu.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
        *reinterpret_cast<uint32_t*>(p) = r;
    }
};
template <> struct Pair<__nv_bfloat16> {
    using Raw = uint32_t;
    __device__ static __forceinline__ Raw zero() { return 0u; }
    __device__ static __forceinline__ Raw load(const __nv_bfloat16* p) { return *reinterpret_cast<const uint32_t*>(p); }
    __device__ static __forceinline__ float2 unpack(Raw r) { return make_float2(__uint_as_float(r << 16), __uint_as_float(r & 0xffff0000u)); }
    __device__ static __forceinline__ void store(__nv_bfloat16* p, float a, float b) { *reinterpret_cast<__nv_bfloat162*>(p) = __floats2bfloat162_rn(a, b); }
    __device__ static __forceinline__ void store_relu(__nv_bfloat16* p, float a, float b) {
        uint32_t r;
        asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(b), "f"(a));
        *reinterpret_cast<uint32_t*>(p) = r;
    }
};

constexpr int kDwRY = 16;     // output rows per thread strip; kDwRY + 4 input rows = a whole number of 5-row chunks
constexpr int kDwTW = 4;      // output columns per thread
constexpr int kDwNV = kDwTW + 4;   // input columns per thread and row
constexpr int kDwThreads = 128;
constexpr int kDwAhead = 4;   // input rows in flight per thread (cp.async groups)
constexpr int kDwChunks = (kDwRY + 4) / 5;
static_assert((kDwRY + 4) % 5 == 0 && kDwChunks >= 2, "strip height + halo must be a multiple of the accumulator ring");
// TMA-tiled kernel: strips of 36 output rows = 8 five-row input chunks streamed through a ring of 4 shared-memory
// slots (halo overhead 40/36 instead of 20/16, half as many cold starts per SM, same shared memory per block)
constexpr int kDwRYL = 36;
constexpr int kDwChunksL = (kDwRYL + 4) / 5;
constexpr int kDwRing = 4;
static_assert((kDwRYL + 4) % 5 == 0 && kDwChunksL > kDwRing, "long strips: whole chunks, more chunks than ring slots");

// One input row (8 unpacked pixels of the thread's channel pair) scattered into the output rows it feeds.
// Input row r = r0 + I feeds the five output rows r-2..r+2 (tap row dy = r - o + 2); the accumulators of those
// rows live in a ring of 5 whose slot indices are compile-time constants inside a 5-row chunk.  FIRST / LAST
// prune the contributions to rows outside the strip.  Afterwards output row r - 2 is complete and is stored.
// Output addressing: `orow` points at (row o, column x0) of the thread's channel pair and advances by one image
// row after every completed row; `full` (all four columns inside the image) removes the per-column predicates.
template <typename T>
struct DwOut {
    T* orow;
    long long row_stride;      // elements between image rows: w * Cp
    int col_stride;            // elements between columns: Cp
    unsigned storemask;
    bool full, relu;
};

// FAST (compile time): the strip lies inside the image, all four columns exist and the ReLU is on — the completed row is
// stored unconditionally, so a whole 5-row chunk is ONE basic block and ptxas overlaps the next row's LDS / unpack with
// this row's FFMA2s; CP > 0 additionally makes the column offsets immediates of the stores.
template <typename T, bool FIRST, bool LAST, int I, bool FAST = false, int CP = 0>
__device__ __forceinline__ void dw_row(const float2 (&v)[kDwNV], DwOut<T>& out, int r, int h, const float2 (&wv)[25], float2 bv,
                                       float2 (&acc)[5][kDwTW]) {
    using P = Pair<T>;
#pragma unroll
    for (int d = 4; d >= 0; --d) {
        if (FIRST && d < 4 - I) continue;              // output row above the strip
        if (LAST && d > 4 - I) continue;               // output row below the strip
        const int slot = (I + 1 + d) % 5, dy = 4 - d;
        if (d == 4) {
#pragma unroll
            for (int ox = 0; ox < kDwTW; ++ox) acc[slot][ox] = bv;
        }
#pragma unroll
        for (int dx = 0; dx < 5; ++dx)
#pragma unroll
            for (int ox = 0; ox < kDwTW; ++ox) acc[slot][ox] = __ffma2_rn(v[ox + dx], wv[dy * 5 + dx], acc[slot][ox]);
    }
    if (FAST) {
        if (!(FIRST && I < 4)) {
            const int slot = (I + 1) % 5;
            T* op = out.orow;
#pragma unroll
            for (int ox = 0; ox < kDwTW; ++ox)
                P::store_relu(op + ox * (CP > 0 ? CP : out.col_stride), acc[slot][ox].x, acc[slot][ox].y);
            out.orow += out.row_stride;
        }
        return;
    }
    if (!(FIRST && I < 4)) {
        const int slot = (I + 1) % 5;
        if (r - 2 < h) {
            T* op = out.orow;
            if (out.full && out.relu) {
#pragma unroll
                for (int ox = 0; ox < kDwTW; ++ox) P::store_relu(op + ox * out.col_stride, acc[slot][ox].x, acc[slot][ox].y);
            } else {
#pragma unroll
                for (int ox = 0; ox < kDwTW; ++ox) {
                    if ((out.storemask >> ox) & 1u) {
                        float a = acc[slot][ox].x, b = acc[slot][ox].y;
                        if (out.relu) { a = fmaxf(a, 0.f); b = fmaxf(b, 0.f); }
                        P::store(op + ox * out.col_stride, a, b);
                    }
                }
            }
        }
        out.orow += out.row_stride;
    }
}

// ---------------------------------------------------------------------------------------------- TMA-tiled variant
// The block's whole input tile — (kDwRY + 4) rows x (4 * rpb + 4) columns x 2 * ppb channels — is fetched by
// cp.async.bulk.tensor in 5-row boxes (one mbarrier per box, matching the 5-row compute chunks), so the image
// border, the channel tail and every address computation of the input side are done by the copy engine; the
// compute threads only issue LDS, FFMA2 and the output stores.  4 blocks per SM keep three tiles in flight
// while one computes.
struct DwTmaParams {
    int h, w, Cp, relu;
    int ppb, rpb, pair_groups;
    int fast;                        // w % 4 == 0 and ReLU on: strips inside the image take the branch-free store path
    int row_bytes, pix_bytes;        // of the shared tile
    int chunk_stride, chunk_bytes;   // 5-row box: distance between boxes in shared memory (128-aligned), bytes transferred
};

template <typename T, bool FIRST, bool LAST, bool FAST = false, int CP = 0>
__device__ __forceinline__ void dw_chunk_smem(const unsigned char* __restrict__ rows, int row_bytes, int pix_bytes, DwOut<T>& out,
                                              int r0, int h, const float2 (&wv)[25], float2 bv, float2 (&acc)[5][kDwTW]) {
    using P = Pair<T>;
    using Raw = typename P::Raw;
    auto step = [&](auto Ic) {
        constexpr int I = decltype(Ic)::value;
        const unsigned char* rp = rows + I * row_bytes;
        float2 v[kDwNV];
#pragma unroll
        for (int j = 0; j < kDwNV; ++j) v[j] = P::unpack(*reinterpret_cast<const Raw*>(rp + j * pix_bytes));
        dw_row<T, FIRST, LAST, I, FAST, CP>(v, out, r0 + I, h, wv, bv, acc);
    };
    step(std::integral_constant<int, 0>{}); step(std::integral_constant<int, 1>{}); step(std::integral_constant<int, 2>{});
    step(std::integral_constant<int, 3>{}); step(std::integral_constant<int, 4>{});
}

// PPB > 0 fixes the pairs per block at compile time (tile strides become immediates); PPB == 0 reads them from pr.
template <typename T, int PPB, int CP = 0>
__global__ void __launch_bounds__(kDwThreads, 4)
dwconv5_tma_kernel(const __grid_constant__ CUtensorMap map, const float* __restrict__ wgt, const float* __restrict__ bias,
                   T* __restrict__ out, const DwTmaParams pr) {
    struct Geo {
        int ppb, rpb, pix_bytes, row_bytes, chunk_bytes, chunk_stride;
    };
    Geo p;
    if (PPB > 0) {
        p.ppb = PPB; p.rpb = kDwThreads / (PPB > 0 ? PPB : 1); p.pix_bytes = 2 * PPB * (int)sizeof(T);
        p.row_bytes = (p.rpb * kDwTW + 4) * p.pix_bytes; p.chunk_bytes = 5 * p.row_bytes;
        p.chunk_stride = (p.chunk_bytes + 127) / 128 * 128;
    } else {
        p.ppb = pr.ppb; p.rpb = pr.rpb; p.pix_bytes = pr.pix_bytes; p.row_bytes = pr.row_bytes;
        p.chunk_bytes = pr.chunk_bytes; p.chunk_stride = pr.chunk_stride;
    }
    extern __shared__ unsigned char dw_smem_raw[];
    unsigned char* tile = dw_smem_raw + ((128u - (tma::saddr(dw_smem_raw) & 127u)) & 127u);   // 128-byte aligned box destinations
    uint64_t* bars = reinterpret_cast<uint64_t*>(tile + kDwRing * p.chunk_stride);

    const int tid = threadIdx.x;
    const int pg = blockIdx.x % pr.pair_groups, rg = blockIdx.x / pr.pair_groups;
    const int b = blockIdx.z, y0 = blockIdx.y * kDwRYL;
    const int h = pr.h, w = pr.w, Cp = pr.Cp;
    // chunk ch completes output rows y0 + 5 ch - 4 ...: it is needed only if the first of them is inside the image
    int nch = kDwChunksL;
    while (nch > 1 && y0 + 5 * (nch - 1) - 4 >= h) --nch;

    if (tid == 0) {
        tma::prefetch_map(&map);
        for (int k = 0; k < kDwRing; ++k) tma::mbar_init(&bars[k], 1);
        tma::fence_barrier_init();
    }
    __syncthreads();
    auto issue = [&](int ch) {      // chunk ch -> ring slot ch % kDwRing (thread 0 only)
        const int slot = ch % kDwRing;
        tma::mbar_expect_tx(&bars[slot], (uint32_t)p.chunk_bytes);
        tma::load_4d(tile + slot * p.chunk_stride, &map, &bars[slot], pg * p.ppb * 2, rg * p.rpb * kDwTW - 2, y0 - 2 + 5 * ch, b);
    };
    if (tid == 0) {
        for (int k = 0; k < kDwRing && k < nch; ++k) issue(k);
    }

    const int npairs = Cp >> 1, runs = (w + kDwTW - 1) / kDwTW;
    const int pair_l = tid % p.ppb, run_l = tid / p.ppb;
    const int pair = pg * p.ppb + pair_l, run = rg * p.rpb + run_l;
    const bool active = run_l < p.rpb && pair < npairs && run < runs;   // thread 0 (the issuer) is always active
    if (!active) return;                                                // exited threads no longer count at the barriers below
    const int c = pair * 2, x0 = run * kDwTW;
    DwOut<T> o;
    o.row_stride = (long long)w * Cp;
    o.col_stride = Cp;
    o.orow = out + (long long)b * h * o.row_stride + ((long long)y0 * w + x0) * Cp + c;
    o.storemask = 0;
#pragma unroll
    for (int ox = 0; ox < kDwTW; ++ox) o.storemask |= (x0 + ox < w) ? (1u << ox) : 0u;
    o.full = o.storemask == (1u << kDwTW) - 1u;
    o.relu = pr.relu != 0;

    float2 wv[25];
#pragma unroll
    for (int k = 0; k < 25; ++k) wv[k] = __ldg(reinterpret_cast<const float2*>(wgt + (long long)k * Cp + c));
    const float2 bv = __ldg(reinterpret_cast<const float2*>(bias + c));

    const unsigned char* tb = tile + run_l * kDwTW * p.pix_bytes + pair_l * (int)sizeof(typename Pair<T>::Raw);
    float2 acc[5][kDwTW];
    // after a chunk has been consumed by every thread of the block its slot is refilled with the chunk kDwRing ahead
    auto refill = [&](int ch) {
        if (ch + kDwRing < nch) {
            __syncthreads();
            if (tid == 0) issue(ch + kDwRing);
        }
    };
    if (pr.fast && y0 + kDwRYL <= h) {       // block-uniform: whole strip inside the image, w % 4 == 0, ReLU on
        tma::mbar_wait(&bars[0], 0);
        dw_chunk_smem<T, true, false, true, CP>(tb, p.row_bytes, p.pix_bytes, o, y0 - 2, h, wv, bv, acc);
        refill(0);
#pragma unroll 1
        for (int ch = 1; ch < kDwChunksL - 1; ++ch) {
            tma::mbar_wait(&bars[ch % kDwRing], (uint32_t)((ch / kDwRing) & 1));
            dw_chunk_smem<T, false, false, true, CP>(tb + (ch % kDwRing) * p.chunk_stride, p.row_bytes, p.pix_bytes, o, y0 - 2 + 5 * ch, h,
                                                     wv, bv, acc);
            refill(ch);
        }
        tma::mbar_wait(&bars[(kDwChunksL - 1) % kDwRing], (uint32_t)(((kDwChunksL - 1) / kDwRing) & 1));
        dw_chunk_smem<T, false, true, true, CP>(tb + ((kDwChunksL - 1) % kDwRing) * p.chunk_stride, p.row_bytes, p.pix_bytes, o,
                                                y0 - 2 + 5 * (kDwChunksL - 1), h, wv, bv, acc);
        return;
    }
    tma::mbar_wait(&bars[0], 0);
    dw_chunk_smem<T, true, false>(tb, p.row_bytes, p.pix_bytes, o, y0 - 2, h, wv, bv, acc);
    refill(0);
#pragma unroll 1
    for (int ch = 1; ch < kDwChunksL - 1; ++ch) {
        if (ch >= nch) return;
        tma::mbar_wait(&bars[ch % kDwRing], (uint32_t)((ch / kDwRing) & 1));
        dw_chunk_smem<T, false, false>(tb + (ch % kDwRing) * p.chunk_stride, p.row_bytes, p.pix_bytes, o, y0 - 2 + 5 * ch, h, wv, bv, acc);
        refill(ch);
    }
    if (nch < kDwChunksL) return;
    tma::mbar_wait(&bars[(kDwChunksL - 1) % kDwRing], (uint32_t)(((kDwChunksL - 1) / kDwRing) & 1));
    dw_chunk_smem<T, false, true>(tb + ((kDwChunksL - 1) % kDwRing) * p.chunk_stride, p.row_bytes, p.pix_bytes, o,
                                  y0 - 2 + 5 * (kDwChunksL - 1), h, wv, bv, acc);
}

template <typename T>
static int launch_dw_tma(const void* x, const float* weight, const float* bias, void* out, int B, int h, int w, int C, int relu,
                         int dtype, cudaStream_t st) {
    const int esz = (int)sizeof(T);
    const int npairs = C / 2, runs = (w + kDwTW - 1) / kDwTW, strips = (h + kDwRYL - 1) / kDwRYL;
    const int pair_groups = cdiv(npairs, 32);
    const int gran = 16 / (2 * esz);                               // pairs per 16 bytes
    int ppb = cdiv(cdiv(npairs, pair_groups), gran) * gran;        // <= 32; box rows are whole 16-byte units
    if (ppb > 32) ppb = 32;
    const int rpb = kDwThreads / ppb;
    const int tcols = rpb * kDwTW + 4;
    DwTmaParams p;
    p.h = h; p.w = w; p.Cp = C; p.relu = relu; p.ppb = ppb; p.rpb = rpb; p.pair_groups = pair_groups;
    p.fast = (relu && w % kDwTW == 0) ? 1 : 0;
    p.pix_bytes = 2 * ppb * esz;
    p.row_bytes = tcols * p.pix_bytes;
    p.chunk_bytes = 5 * p.row_bytes;
    p.chunk_stride = (p.chunk_bytes + 127) / 128 * 128;
    const size_t smem = (size_t)kDwRing * p.chunk_stride + kDwRing * sizeof(uint64_t) + 128;

    CUtensorMap map;
    const uint64_t dims[4] = {(uint64_t)C, (uint64_t)w, (uint64_t)h, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)C * esz, (uint64_t)w * C * esz, (uint64_t)h * w * C * esz};
    const uint32_t box[4] = {(uint32_t)(2 * ppb), (uint32_t)tcols, 5u, 1u};
    if (int rc = tma::encode_tiled(&map, dtype, 4, x, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;

    const dim3 grid((unsigned)(pair_groups * cdiv(runs, rpb)), (unsigned)strips, (unsigned)B);
    using Kernel = void (*)(const CUtensorMap, const float*, const float*, T*, const DwTmaParams);
    // the four refiner widths above the finest scale get their channel count at compile time (store offsets become immediates)
    Kernel kernel = (ppb == 24 && C == 144) ? dwconv5_tma_kernel<T, 24, 144>
                  : (ppb == 32 && C == 576) ? dwconv5_tma_kernel<T, 32, 576>
                  : (ppb == 32 && C == 1144) ? dwconv5_tma_kernel<T, 32, 1144>
                  : (ppb == 32 && C == 1384) ? dwconv5_tma_kernel<T, 32, 1384>
                  : ppb == 32 ? dwconv5_tma_kernel<T, 32> : ppb == 24 ? dwconv5_tma_kernel<T, 24>
                  : ppb == 12 ? dwconv5_tma_kernel<T, 12> : dwconv5_tma_kernel<T, 0>;
    // opt in to > 48 KB of dynamic shared memory (idempotent, cheap; done per call to stay per-instantiation correct)
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024);
    if (e != cudaSuccess) return fail((int)e, "roma_dwconv5x5_bias_act: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    kernel<<<grid, kDwThreads, smem, st>>>(map, weight, bias, (T*)out, p);
    return 0;
}

// ---------------------------------------------------------------------------------------------- fused block, Cp = 24
// One ConvRefiner block of the finest scale in one pass: depthwise 5x5 + folded BN + ReLU (as above) followed by the
// block's 1x1 convolution 24 -> 24 (matcher.py:77-103: conv1, norm, relu, conv2).  At 24 channels both halves are
// memory-bound when run separately (four passes over a 573 MB activation at 864 x 864, B = 16); fused, the depthwise
// result never leaves the SM: the 16 x 40-pixel tile is written as 16-bit pairs into shared memory, and after one
// barrier the four warps run the 1x1 convolution as warp-level tensor-core MMAs (mma.sync m16n8k16, fp32 accumulate;
// A fragments by conflict-free LDS.32 from the tile, the 24 x 24 weights held as B fragments in registers since
// kernel start) and store the packed 16-bit result: two passes instead of four.  The rounding points are those of
// the two-kernel path (16-bit activation between the convolutions, fp32 accumulation in both).
constexpr int kPwC = 24, kPwPPB = 12, kPwRPB = 10;                 // channels, pairs and 4-column runs per block
constexpr int kPwCols = kPwRPB * kDwTW;                            // 40 output columns per tile
constexpr int kPwSub = 3;                                          // consecutive 16-row strips per block
constexpr int kPwPixBytes = 2 * kPwPPB * 2, kPwRowBytes = (kPwCols + 4) * kPwPixBytes, kPwChunkBytes = 5 * kPwRowBytes;
constexpr int kPwChunkStride = (kPwChunkBytes + 127) / 128 * 128;
constexpr int kPwObufBytes = kDwRY * kPwCols * kPwC * 2;
constexpr size_t kPwSmem = (size_t)kDwChunks * kPwChunkStride + 128 + kPwObufBytes + 128;

template <typename T> __device__ __forceinline__ void dw_mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1);
template <> __device__ __forceinline__ void dw_mma16816<__nv_bfloat16>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <> __device__ __forceinline__ void dw_mma16816<__half>(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <typename T> __device__ __forceinline__ uint32_t dw_pack2(float lo, float hi);
template <> __device__ __forceinline__ uint32_t dw_pack2<__nv_bfloat16>(float lo, float hi) {
    const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}
template <> __device__ __forceinline__ uint32_t dw_pack2<__half>(float lo, float hi) {
    const __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<const uint32_t*>(&v);
}

template <typename T>
__global__ void __launch_bounds__(kDwThreads, 3)
dwconv5_pw24_tma_kernel(const __grid_constant__ CUtensorMap map, const float* __restrict__ wgt, const float* __restrict__ bias,
                        const float* __restrict__ pw_w, const float* __restrict__ pw_b, T* __restrict__ out, int h, int w, int relu) {
    static_assert(sizeof(T) == 2, "fused block: 16-bit activations");
    extern __shared__ unsigned char dw_smem_raw[];
    unsigned char* tile = dw_smem_raw + ((128u - (tma::saddr(dw_smem_raw) & 127u)) & 127u);
    uint64_t* bars = reinterpret_cast<uint64_t*>(tile + kDwChunks * kPwChunkStride);
    T* obuf = reinterpret_cast<T*>(tile + kDwChunks * kPwChunkStride + 128);          // [16 rows][40 cols][24 ch]

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int rg = blockIdx.x, b = blockIdx.z, ybase = blockIdx.y * (kDwRY * kPwSub);
    // chunks of the 16-row strip starting at row y that complete at least one output row inside the image
    auto chunks_of = [&](int y) { int n = kDwChunks; while (n > 1 && y + 5 * (n - 1) - 4 >= h) --n; return n; };
    auto issue_strip = [&](int y) {                                  // thread 0: the strip's 5-row boxes, one mbarrier each
        const int n = chunks_of(y);
        for (int k = 0; k < n; ++k) {
            tma::mbar_expect_tx(&bars[k], (uint32_t)kPwChunkBytes);
            tma::load_4d(tile + k * kPwChunkStride, &map, &bars[k], 0, rg * kPwCols - 2, y - 2 + 5 * k, b);
        }
    };

    if (tid == 0) {
        tma::prefetch_map(&map);
        for (int k = 0; k < kDwChunks; ++k) tma::mbar_init(&bars[k], 1);
        tma::fence_barrier_init();
    }
    __syncthreads();
    if (tid == 0) issue_strip(ybase);

    // B fragments of the 1x1 convolution: B[k][n] = W[n][k]; b0 = (k = 16 ks + 2t, +1; n = 8 nt + g), b1 = 8 further in k
    const int g = lane >> 2, t4 = lane & 3;
    uint32_t bf[2][3][2];
    float bv[3][2];
#pragma unroll
    for (int nt = 0; nt < 3; ++nt) {
        const int n = nt * 8 + g;
#pragma unroll
        for (int ks = 0; ks < 2; ++ks) {
#pragma unroll
            for (int hh = 0; hh < 2; ++hh) {
                const int k = ks * 16 + 2 * t4 + 8 * hh;
                const float w0 = k < kPwC ? __ldg(pw_w + n * kPwC + k) : 0.f;
                const float w1 = k + 1 < kPwC ? __ldg(pw_w + n * kPwC + k + 1) : 0.f;
                bf[ks][nt][hh] = dw_pack2<T>(w0, w1);
            }
        }
        bv[nt][0] = __ldg(pw_b + nt * 8 + 2 * t4);
        bv[nt][1] = __ldg(pw_b + nt * 8 + 2 * t4 + 1);
    }

    // A block walks kPwSub consecutive 16-row strips: weights and fragments are loaded once, and the boxes of the next strip
    // are issued right after the barrier that ends phase 1 (all four slots are free then), so they arrive under phase 2.
    const int runs = (w + kDwTW - 1) / kDwTW;
    const int pair_l = tid % kPwPPB, run_l = tid / kPwPPB;
    const int run = rg * kPwRPB + run_l;
    const bool ph1 = run_l < kPwRPB && run < runs;
    const int c = pair_l * 2;
    float2 wv[25];
#pragma unroll
    for (int k = 0; k < 25; ++k) wv[k] = ph1 ? __ldg(reinterpret_cast<const float2*>(wgt + k * kPwC + c)) : make_float2(0.f, 0.f);
    const float2 bvv = ph1 ? __ldg(reinterpret_cast<const float2*>(bias + c)) : make_float2(0.f, 0.f);
    const unsigned char* tb = tile + run_l * kDwTW * kPwPixBytes + pair_l * 4;
    const uint32_t* ob32 = reinterpret_cast<const uint32_t*>(obuf);
    T* outb = out + (long long)b * h * w * kPwC;
    const int x_base = rg * kPwCols;

#pragma unroll 1
    for (int ss = 0; ss < kPwSub; ++ss) {
        const int y0 = ybase + ss * kDwRY;
        if (y0 >= h) break;                                          // block-uniform
        const int nch = chunks_of(y0);
        const uint32_t par = (uint32_t)(ss & 1);                     // every barrier completes once per strip
        // ---- phase 1: depthwise 5x5 + bias + ReLU into the shared-memory tile ----------------------------------
        if (ph1) {
            DwOut<T> o;
            o.row_stride = kPwCols * kPwC;
            o.col_stride = kPwC;
            o.orow = obuf + run_l * kDwTW * kPwC + c;
            o.storemask = (1u << kDwTW) - 1u;
            o.full = true;
            o.relu = relu != 0;
            float2 acc[5][kDwTW];
            // FAST (ReLU on): rows are written to the tile unconditionally — phase 2 masks what lies outside the image
            auto phase1 = [&](auto fast) {
                constexpr bool F = decltype(fast)::value;
                tma::mbar_wait(&bars[0], par);
                dw_chunk_smem<T, true, false, F, kPwC>(tb, kPwRowBytes, kPwPixBytes, o, y0 - 2, h, wv, bvv, acc);
#pragma unroll 1
                for (int ch = 1; ch < kDwChunks - 1; ++ch) {
                    if (ch < nch) {
                        tma::mbar_wait(&bars[ch], par);
                        dw_chunk_smem<T, false, false, F, kPwC>(tb + ch * kPwChunkStride, kPwRowBytes, kPwPixBytes, o, y0 - 2 + 5 * ch, h, wv,
                                                                bvv, acc);
                    }
                }
                if (nch == kDwChunks) {
                    tma::mbar_wait(&bars[kDwChunks - 1], par);
                    dw_chunk_smem<T, false, true, F, kPwC>(tb + (kDwChunks - 1) * kPwChunkStride, kPwRowBytes, kPwPixBytes, o,
                                                           y0 - 2 + 5 * (kDwChunks - 1), h, wv, bvv, acc);
                }
            };
            if (relu) phase1(std::true_type{}); else phase1(std::false_type{});
        }
        __syncthreads();                                             // tile complete, every input slot consumed
        if (tid == 0 && ss + 1 < kPwSub && y0 + kDwRY < h) issue_strip(y0 + kDwRY);

        // ---- phase 2: 1x1 convolution of the tile on tensor cores, 16 pixels per MMA tile ----------------------
#pragma unroll 2
        for (int mt = warp; mt < kDwRY * kPwCols / 16; mt += kDwThreads / 32) {
            const int qlo = mt * 16 + g, qhi = qlo + 8;
            uint32_t a0[4], a1[4];
            a0[0] = ob32[qlo * 12 + t4];     a0[1] = ob32[qhi * 12 + t4];
            a0[2] = ob32[qlo * 12 + 4 + t4]; a0[3] = ob32[qhi * 12 + 4 + t4];
            a1[0] = ob32[qlo * 12 + 8 + t4]; a1[1] = ob32[qhi * 12 + 8 + t4];
            a1[2] = 0u; a1[3] = 0u;                                   // k = 24..31 do not exist
            const int ylo = y0 + qlo / kPwCols, xlo = x_base + qlo % kPwCols;
            const int yhi = y0 + qhi / kPwCols, xhi = x_base + qhi % kPwCols;
            const bool vlo = ylo < h && xlo < w, vhi = yhi < h && xhi < w;
            T* plo = outb + ((long long)ylo * w + xlo) * kPwC + 2 * t4;
            T* phi = outb + ((long long)yhi * w + xhi) * kPwC + 2 * t4;
#pragma unroll
            for (int nt = 0; nt < 3; ++nt) {
                float cc[4] = {bv[nt][0], bv[nt][1], bv[nt][0], bv[nt][1]};
                dw_mma16816<T>(cc, a0, bf[0][nt][0], bf[0][nt][1]);
                dw_mma16816<T>(cc, a1, bf[1][nt][0], bf[1][nt][1]);
                if (vlo) *reinterpret_cast<uint32_t*>(plo + nt * 8) = dw_pack2<T>(cc[0], cc[1]);
                if (vhi) *reinterpret_cast<uint32_t*>(phi + nt * 8) = dw_pack2<T>(cc[2], cc[3]);
            }
        }
        if (ss + 1 < kPwSub) __syncthreads();                        // the tile is rewritten by the next strip's phase 1
    }
}

template <typename T>
static int launch_dw_pw24(const void* x, const float* weight, const float* bias, const float* pw_w, const float* pw_b, void* out,
                          int B, int h, int w, int relu, int dtype, cudaStream_t st) {
    const int runs = (w + kDwTW - 1) / kDwTW, strips = (h + kDwRY * kPwSub - 1) / (kDwRY * kPwSub);
    CUtensorMap map;
    const uint64_t dims[4] = {(uint64_t)kPwC, (uint64_t)w, (uint64_t)h, (uint64_t)B};
    const uint64_t strides[3] = {(uint64_t)kPwC * 2, (uint64_t)w * kPwC * 2, (uint64_t)h * w * kPwC * 2};
    const uint32_t box[4] = {(uint32_t)kPwC, (uint32_t)(kPwCols + 4), 5u, 1u};
    if (int rc = tma::encode_tiled(&map, dtype, 4, x, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
    cudaError_t e = cudaFuncSetAttribute(dwconv5_pw24_tma_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPwSmem);
    if (e != cudaSuccess) return fail((int)e, "roma_dwconv5x5_pointwise: cudaFuncSetAttribute: %s", cudaGetErrorString(e));
    const dim3 grid((unsigned)cdiv(runs, kPwRPB), (unsigned)strips, (unsigned)B);
    dwconv5_pw24_tma_kernel<T><<<grid, kDwThreads, kPwSmem, st>>>(map, weight, bias, pw_w, pw_b, (T*)out, h, w, relu);
    return 0;
}

// ---------------------------------------------------------------------------------------------- cp.async variant
// Asynchronous global -> shared copy of one channel pair; src_bytes = 0 zero-fills (image border).
template <int BYTES>
__device__ __forceinline__ void cp_async_pair(void* smem_dst, const void* gsrc, bool valid) {
    const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
    const int n = valid ? BYTES : 0;
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2, %3;" ::"r"(d), "l"(gsrc), "n"(BYTES), "r"(n) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// Per-thread private staging ring in shared memory: stage s, column j of thread t lives at [s][j][t]
// (consecutive threads -> consecutive banks).  Nothing is shared between threads, so no block barrier is needed:
// a thread only waits for its own copy groups.
template <typename T>
struct DwStage {
    typename Pair<T>::Raw v[5][kDwNV][kDwThreads];
};

template <typename T>
__device__ __forceinline__ void dw_issue_row(DwStage<T>& st, int stage, const T* __restrict__ xb, int r, int x0, unsigned colmask,
                                             int h, int w, int Cp) {
    using Raw = typename Pair<T>::Raw;
    const bool yok = r >= 0 && r < h;
    const T* rp = xb + ((long long)(yok ? r : 0) * w + (x0 - 2)) * Cp;
#pragma unroll
    for (int j = 0; j < kDwNV; ++j) {
        const bool ok = yok && ((colmask >> j) & 1u);
        cp_async_pair<sizeof(Raw)>(&st.v[stage][j][threadIdx.x], ok ? (const void*)(rp + j * Cp) : (const void*)xb, ok);
    }
}

// Row r0 + I is read from staging slot I (it was requested kDwAhead steps earlier); row r0 + I + kDwAhead is
// requested into slot (I + kDwAhead) % 5, whose previous content (row r0 + I - 1) has already been consumed.
template <typename T, bool FIRST, bool LAST>
__device__ __forceinline__ void dw_chunk(DwStage<T>& st, const T* __restrict__ xb, DwOut<T>& out, int r0, int r_end, int x0,
                                         unsigned colmask, int h, int w, int Cp, const float2 (&wv)[25], float2 bv,
                                         float2 (&acc)[5][kDwTW]) {
    using P = Pair<T>;
    auto step = [&](auto Ic) {
        constexpr int I = decltype(Ic)::value;
        const int r = r0 + I;
        if (r + kDwAhead < r_end) dw_issue_row<T>(st, (I + kDwAhead) % 5, xb, r + kDwAhead, x0, colmask, h, w, Cp);
        cp_async_commit();                             // one group per step, empty or not: keeps the count uniform
        cp_async_wait<kDwAhead>();                     // row r has landed
        float2 v[kDwNV];
#pragma unroll
        for (int j = 0; j < kDwNV; ++j) v[j] = P::unpack(st.v[I][j][threadIdx.x]);
        dw_row<T, FIRST, LAST, I>(v, out, r, h, wv, bv, acc);
    };
    step(std::integral_constant<int, 0>{}); step(std::integral_constant<int, 1>{}); step(std::integral_constant<int, 2>{});
    step(std::integral_constant<int, 3>{}); step(std::integral_constant<int, 4>{});
}

// Grid: x = pair groups * run groups, y = strips, z = batch.  Inside a block, threads run over ppb channel pairs
// (fastest: a warp reads contiguous bytes of one pixel) and rpb = 128 / ppb adjacent 4-column runs, so that the
// 4-column halo two neighbouring runs share is requested from the same SM's L1.
template <typename T>
__global__ void __launch_bounds__(kDwThreads, 4)
dwconv5_kernel(const T* __restrict__ x, const float* __restrict__ wgt, const float* __restrict__ bias, T* __restrict__ out,
               int h, int w, int Cp, int relu, int ppb, int rpb, int pair_groups) {
    __shared__ DwStage<T> st;
    const int npairs = Cp >> 1, runs = (w + kDwTW - 1) / kDwTW;
    const int pg = blockIdx.x % pair_groups, rg = blockIdx.x / pair_groups;
    const int pair = pg * ppb + (int)threadIdx.x % ppb;
    const int run = rg * rpb + (int)threadIdx.x / ppb;
    if ((int)threadIdx.x >= ppb * rpb || pair >= npairs || run >= runs) return;
    const int b = blockIdx.z;
    const int c = pair * 2, x0 = run * kDwTW, y0 = blockIdx.y * kDwRY;

    const T* xb = x + (long long)b * h * w * Cp + c;
    unsigned colmask = 0;                              // bit j: column x0 - 2 + j lies inside the image
#pragma unroll
    for (int j = 0; j < kDwNV; ++j) colmask |= (x0 - 2 + j >= 0 && x0 - 2 + j < w) ? (1u << j) : 0u;
    DwOut<T> o;
    o.row_stride = (long long)w * Cp;
    o.col_stride = Cp;
    o.orow = out + (long long)b * h * o.row_stride + ((long long)y0 * w + x0) * Cp + c;
    o.storemask = (colmask >> 2) & ((1u << kDwTW) - 1u);
    o.full = o.storemask == (1u << kDwTW) - 1u;
    o.relu = relu != 0;

    // rows past the image contribute nothing and no output row >= h exists: stop after input row h + 1
    const int r_begin = y0 - 2;
    const int r_end = min(y0 + kDwRY + 2, h + 2);
#pragma unroll
    for (int i = 0; i < kDwAhead; ++i) {
        if (r_begin + i < r_end) dw_issue_row<T>(st, i, xb, r_begin + i, x0, colmask, h, w, Cp);
        cp_async_commit();
    }

    float2 wv[25];
#pragma unroll
    for (int k = 0; k < 25; ++k) wv[k] = __ldg(reinterpret_cast<const float2*>(wgt + (long long)k * Cp + c));
    const float2 bv = __ldg(reinterpret_cast<const float2*>(bias + c));

    float2 acc[5][kDwTW];
    dw_chunk<T, true, false>(st, xb, o, r_begin, r_end, x0, colmask, h, w, Cp, wv, bv, acc);
#pragma unroll 1
    for (int ch = 1; ch < kDwChunks - 1; ++ch) {
        if (r_begin + 5 * ch - 2 >= h) break;          // every remaining output row is below the image
        dw_chunk<T, false, false>(st, xb, o, r_begin + 5 * ch, r_end, x0, colmask, h, w, Cp, wv, bv, acc);
    }
    if (r_begin + 5 * (kDwChunks - 1) - 2 < h)
        dw_chunk<T, false, true>(st, xb, o, r_begin + 5 * (kDwChunks - 1), r_end, x0, colmask, h, w, Cp, wv, bv, acc);
    cp_async_wait<0>();
}

}  // namespace roma

static std::atomic<int> g_dw_path{0};
extern "C" int roma_dwconv_set_path(int path) {
    if (path < 0 || path > 1) return roma::fail(ROMA_ERR_BAD_ARG, "roma_dwconv_set_path: path %d (0 auto, 1 per-thread cp.async)", path);
    g_dw_path.store(path);
    return 0;
}

extern "C" int roma_dwconv5x5_bias_act(const void* x, const float* weight, const float* bias, void* out,
                                       int B, int h, int w, int C, int relu, int dtype, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && weight && bias && out, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: null pointer");
    ROMA_REQUIRE(B > 0 && h > 0 && w > 0 && C > 0, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: bad sizes");
    ROMA_REQUIRE(C % 2 == 0, ROMA_ERR_UNSUPPORTED, "roma_dwconv5x5_bias_act: channel count %d must be even (pad it)", C);
    ROMA_REQUIRE(x != out, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: in-place operation is not supported");
    const uintptr_t al = dtype == ROMA_F32 ? 7u : 3u;
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & al) == 0 &&
                 ((reinterpret_cast<uintptr_t>(weight) | reinterpret_cast<uintptr_t>(bias)) & 7u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_dwconv5x5_bias_act: misaligned pointer");
    const int npairs = C / 2, runs = (w + kDwTW - 1) / kDwTW, strips = (h + kDwRY - 1) / kDwRY;
    ROMA_REQUIRE(B <= 65535 && strips <= 65535, ROMA_ERR_UNSUPPORTED, "roma_dwconv5x5_bias_act: batch / height too large");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int esz = dtype == ROMA_F32 ? 4 : 2;
    // TMA path: 16-byte aligned base and pixel stride (every padded refiner width); otherwise per-thread cp.async
    const bool tma_ok = g_dw_path.load() != 1 && ((long long)C * esz) % 16 == 0 && (reinterpret_cast<uintptr_t>(x) & 15u) == 0 &&
                        tma::encode_fn() != nullptr;
    if (tma_ok) {
        int rc;
        switch (dtype) {
            case ROMA_F32:  rc = launch_dw_tma<float>(x, weight, bias, out, B, h, w, C, relu, dtype, st); break;
            case ROMA_F16:  rc = launch_dw_tma<__half>(x, weight, bias, out, B, h, w, C, relu, dtype, st); break;
            case ROMA_BF16: rc = launch_dw_tma<__nv_bfloat16>(x, weight, bias, out, B, h, w, C, relu, dtype, st); break;
            default: return fail(ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: unknown dtype %d", dtype);
        }
        if (rc) return rc;
        return check_launch("roma_dwconv5x5_bias_act");
    }
    const int pair_groups = cdiv(npairs, 32);
    const int ppb = cdiv(npairs, pair_groups);                  // <= 32 pairs per block, evenly split
    const int rpb = kDwThreads / ppb;
    const dim3 grid((unsigned)(pair_groups * cdiv(runs, rpb)), (unsigned)strips, (unsigned)B);
    switch (dtype) {
        case ROMA_F32:  dwconv5_kernel<float><<<grid, kDwThreads, 0, st>>>((const float*)x, weight, bias, (float*)out, h, w, C, relu, ppb, rpb, pair_groups); break;
        case ROMA_F16:  dwconv5_kernel<__half><<<grid, kDwThreads, 0, st>>>((const __half*)x, weight, bias, (__half*)out, h, w, C, relu, ppb, rpb, pair_groups); break;
        case ROMA_BF16: dwconv5_kernel<__nv_bfloat16><<<grid, kDwThreads, 0, st>>>((const __nv_bfloat16*)x, weight, bias, (__nv_bfloat16*)out, h, w, C, relu, ppb, rpb, pair_groups); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_dwconv5x5_bias_act: unknown dtype %d", dtype);
    }
    return check_launch("roma_dwconv5x5_bias_act");
}

extern "C" int roma_dwconv5x5_pointwise(const void* x, const float* dw_weight, const float* dw_bias, const float* pw_weight,
                                        const float* pw_bias, void* out, int B, int h, int w, int C, int relu, int dtype,
                                        void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x && dw_weight && dw_bias && pw_weight && pw_bias && out, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_pointwise: null pointer");
    ROMA_REQUIRE(B > 0 && h > 0 && w > 0, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_pointwise: bad sizes");
    ROMA_REQUIRE(C == kPwC && (dtype == ROMA_F16 || dtype == ROMA_BF16), ROMA_ERR_UNSUPPORTED,
                 "roma_dwconv5x5_pointwise: the fused block exists for C = 24 in fp16 / bf16 (got C = %d, dtype %d); "
                 "use roma_dwconv5x5_bias_act + a 1x1 convolution", C, dtype);
    ROMA_REQUIRE(x != out, ROMA_ERR_BAD_ARG, "roma_dwconv5x5_pointwise: in-place operation is not supported");
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(out)) & 15u) == 0 &&
                 ((reinterpret_cast<uintptr_t>(dw_weight) | reinterpret_cast<uintptr_t>(dw_bias)) & 7u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_dwconv5x5_pointwise: misaligned pointer");
    const int strips = (h + kDwRY - 1) / kDwRY;
    ROMA_REQUIRE(B <= 65535 && strips <= 65535, ROMA_ERR_UNSUPPORTED, "roma_dwconv5x5_pointwise: batch / height too large");
    ROMA_REQUIRE(tma::encode_fn() != nullptr, ROMA_ERR_UNSUPPORTED, "roma_dwconv5x5_pointwise: cuTensorMapEncodeTiled is not available");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    int rc = dtype == ROMA_F16 ? launch_dw_pw24<__half>(x, dw_weight, dw_bias, pw_weight, pw_bias, out, B, h, w, relu, dtype, st)
                               : launch_dw_pw24<__nv_bfloat16>(x, dw_weight, dw_bias, pw_weight, pw_bias, out, B, h, w, relu, dtype, st);
    if (rc) return rc;
    return check_launch("roma_dwconv5x5_pointwise");
}
