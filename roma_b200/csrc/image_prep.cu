// (f-3) input path: uint8 RGB batch -> ImageNet-normalised fp32 images at the resolution a pass needs, one kernel.
//
// Replaces (reference): the per-image preprocessing of RegressionMatcher.match
//   romatch/models/matcher.py:606-637, 667-676 and romatch/utils/utils.py:165-261 (get_tuple_transform_ops:
//   resize -> ToTensor (/255) -> Normalize(mean, std)), in the tensor form the drop-in accepts: a batch of uint8
//   images of any resolution is normalised and, where the pass resolution differs, resized with the antialiased
//   bicubic filter of F.interpolate(..., mode="bicubic", antialias=True, align_corners=False).
//   Before: float() / 255, sub, div (three passes over a fp32 copy of the full-resolution batch) and ATen's
//   _upsample_bicubic2d_aa; here one launch per target resolution reads the uint8 pixels and writes the result.
//
// Arithmetic = ATen's (UpSample.cuh, upsample_gen2d_aa_out_frame), in fp32:
//   scale = in / out;  support = scale >= 1 ? 2 * scale : 2;  centre = scale * (i + 0.5);
//   xmin = max(int(centre - support + 0.5), 0);  xsize = min(int(centre + support + 0.5), in) - xmin;
//   w_j = cubic_{a = -0.5}((j + xmin - centre + 0.5) * (scale >= 1 ? 1 / scale : 1)),  normalised by their sum;
//   out = sum_y wy * (sum_x wx * v[y][x])   (inner sum over x first, like ATen).
// v is the NORMALISED pixel ((u8 / 255 - mean) / std with IEEE divisions): taken from a 3 x 256 table built with exactly
// those operations, so the resize sees bit-identical inputs to the elementwise pipeline it replaces, and the
// identity case (same resolution) is bit-identical to it.
#include "common.cuh"

namespace roma {

constexpr int kPrepMaxTaps = 18;      // 2 * ceil(2 * scale) + 1 taps per dimension: scale <= 4 (e.g. 2240 -> 560)

__device__ __forceinline__ float cubic_aa(float x) {
    const float a = -0.5f;
    x = fabsf(x);
    if (x < 1.f) return ((a + 2.f) * x - (a + 3.f)) * x * x + 1.f;
    if (x < 2.f) return ((a * x - 5.f * a) * x + 8.f * a) * x - 4.f * a;
    return 0.f;
}

constexpr int kPrepTile = 16;         // output tile of a block: 16 x 16 pixels, all three channels
constexpr int kPrepThreads = kPrepTile * kPrepTile;

// geometry of output index i along one dimension (the integer arithmetic of ATen's weight computation)
struct AaSpan { float center; int lo, n; };
__device__ __forceinline__ AaSpan aa_span(int i, int in_size, float scale, float support) {
    AaSpan s;
    s.center = scale * ((float)i + 0.5f);
    s.lo = max((int)(s.center - support + 0.5f), 0);
    s.n = min(min((int)(s.center + support + 0.5f), in_size) - s.lo, kPrepMaxTaps);
    return s;
}

// Resize: a block owns a 16 x 16 output tile of all three channels and works in four phases out of shared memory.
//   1. weights: thread (o, j) evaluates the raw cubic taps j and j + 16 of column ox0 + o and of row oy0 + o; after a
//      barrier it sums its vector's taps in ascending order (ATen's `total`) and divides its own taps by it;
//   2. stage: the union of the input spans of the tile (the spans are monotonic in the output index, so the first and
//      last column / row bound it) is loaded once, coalesced along x, as NORMALISED fp32 values (table lookup per byte);
//   3. horizontal: tmp[c][y][ox] = sum_i tile[c][y][x0(ox) + i] * wx[ox][i], ascending i - ATen's inner sum, which
//      does not depend on the output row, so it is evaluated once per input row instead of once per (row, tap);
//   4. vertical: out[c][oy][ox] = sum_j tmp[c][y0(oy) + j][ox] * wy[oy][j], ascending j.
// Same products and the same summation order as the direct double loop, 2.3x fewer of them at scale 1.54.
__global__ void __launch_bounds__(kPrepThreads)
prepare_images_kernel(const unsigned char* __restrict__ src, float* __restrict__ out, int B, int H, int W, int h, int w,
                      float m0, float m1, float m2, float s0, float s1, float s2, int tile_w, int tile_h) {
    extern __shared__ float prep_smem[];
    __shared__ float lut[3][256];
    __shared__ float wraw[2][kPrepTile][kPrepMaxTaps], wn[2][kPrepTile][kPrepMaxTaps];
    __shared__ int first[2][kPrepTile], taps[2][kPrepTile];
    float* tile = prep_smem;                                          // [3][tile_h][tile_w]
    float* tmp = prep_smem + 3 * tile_h * tile_w;                     // [3][tile_h][16]
    const int tid = threadIdx.x, o = tid / kPrepTile, jj = tid % kPrepTile;
    const int b = blockIdx.z;
    const int ox0 = blockIdx.x * kPrepTile, oy0 = blockIdx.y * kPrepTile;
    const float sc[2] = {(float)W / (float)w, (float)H / (float)h};
    const int in_size[2] = {W, H}, out_size[2] = {w, h}, o0[2] = {ox0, oy0};
    {
        const float f = (float)tid / 255.f;
        lut[0][tid] = (f - m0) / s0; lut[1][tid] = (f - m1) / s1; lut[2][tid] = (f - m2) / s2;
    }
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        const float support = sc[d] >= 1.f ? 2.f * sc[d] : 2.f, invscale = sc[d] >= 1.f ? 1.f / sc[d] : 1.f;
        const int i = min(o0[d] + o, out_size[d] - 1);                // columns / rows past the edge repeat the last one
        const AaSpan sp = aa_span(i, in_size[d], sc[d], support);
        const float xmc = (float)sp.lo - sp.center;
        for (int j = jj; j < kPrepMaxTaps; j += kPrepTile)
            wraw[d][o][j] = j < sp.n ? cubic_aa(((float)j + xmc + 0.5f) * invscale) : 0.f;
        if (jj == 0) { first[d][o] = sp.lo; taps[d][o] = sp.n; }
    }
    __syncthreads();
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        float total = 0.f;
        const int n = taps[d][o];
        for (int j = 0; j < n; ++j) total += wraw[d][o][j];
        for (int j = jj; j < kPrepMaxTaps; j += kPrepTile)
            wn[d][o][j] = total != 0.f ? wraw[d][o][j] / total : wraw[d][o][j];
    }
    const int last = kPrepTile - 1;
    const int xlo = first[0][0], ylo = first[1][0];
    const int tw = min(first[0][last] + taps[0][last] - xlo, tile_w), th = min(first[1][last] + taps[1][last] - ylo, tile_h);
    const unsigned char* sb = src + (long long)b * 3 * H * W;
    for (int r = tid / 32; r < 3 * th; r += kPrepThreads / 32) {
        const int c = r / th, y = r - c * th;
        const unsigned char* srow = sb + ((long long)c * H + ylo + y) * W + xlo;
        float* trow = tile + (c * tile_h + y) * tile_w;
        for (int x = tid % 32; x < tw; x += 32) trow[x] = lut[c][srow[x]];
    }
    __syncthreads();
    {
        const int ox = jj, x0 = first[0][ox] - xlo, nx = taps[0][ox];
        const float* wx = wn[0][ox];
        for (int r = o; r < 3 * th; r += kPrepTile) {
            const int c = r / th, y = r - c * th;
            const float* row = tile + (c * tile_h + y) * tile_w + x0;
            float acc = 0.f;
            for (int i = 0; i < nx; ++i) acc += row[i] * wx[i];
            tmp[(c * tile_h + y) * kPrepTile + ox] = acc;
        }
    }
    __syncthreads();
    const int ox = ox0 + jj, oy = oy0 + o;
    if (ox >= w || oy >= h) return;
    const int y0 = first[1][o] - ylo, ny = taps[1][o];
    const float* wy = wn[1][o];
    float* ob = out + (long long)b * 3 * h * w + (long long)oy * w + ox;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float* col = tmp + (c * tile_h + y0) * kPrepTile + jj;
        float acc = 0.f;
        for (int j = 0; j < ny; ++j) acc += col[j * kPrepTile] * wy[j];
        ob[(long long)c * h * w] = acc;
    }
}

// Same resolution: pure table lookup, 4 pixels per thread where the row length allows (uchar4 in, float4 out).
__global__ void __launch_bounds__(256)
normalise_images_kernel(const unsigned char* __restrict__ src, float* __restrict__ out, long long plane, long long total,
                        float m0, float m1, float m2, float s0, float s1, float s2, int vec) {
    __shared__ float lut[3][256];
    {
        const int v = threadIdx.x;
        const float f = (float)v / 255.f;
        lut[0][v] = (f - m0) / s0; lut[1][v] = (f - m1) / s1; lut[2][v] = (f - m2) / s2;
    }
    __syncthreads();
    const long long stride = (long long)gridDim.x * blockDim.x;
    if (vec) {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i * 4 < total; i += stride) {
            const int c = (int)((i * 4 / plane) % 3);
            const uchar4 u = reinterpret_cast<const uchar4*>(src)[i];
            reinterpret_cast<float4*>(out)[i] = make_float4(lut[c][u.x], lut[c][u.y], lut[c][u.z], lut[c][u.w]);
        }
    } else {
        for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += stride)
            out[i] = lut[(int)((i / plane) % 3)][src[i]];
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// PIL semantics: `Image.resize(size, Image.BICUBIC)` of an 8-bit image (Pillow, libImaging/Resample.c), which is what the
// reference applies to every file / PIL input (utils/utils.py:223-232), followed by ToTensor and Normalize.  Integer
// arithmetic, bit-exact: coefficients in double precision (explicit round-to-nearest intrinsics, so that no FMA
// contraction changes a rounding), quantised to 22 fractional bits half away from zero; horizontal pass then vertical
// pass, each producing 8-bit pixels `clip8((2^21 + sum u8 * k) >> 22)`; a dimension that does not change is passed
// through (coefficient 2^22 on the pixel itself gives the pixel back exactly).
// Same block structure as above: 16 x 16 output tile, coefficients once per tile, input span staged once (as bytes),
// horizontal pass into an 8-bit intermediate in shared memory, vertical pass from it.
constexpr int kPilBits = 32 - 8 - 2;

__device__ __forceinline__ double pil_bicubic(double x) {
    if (x < 0.0) x = -x;
    if (x < 1.0) return __dadd_rn(__dmul_rn(__dmul_rn(__dsub_rn(__dmul_rn(1.5, x), 2.5), x), x), 1.0);
    if (x < 2.0) return __dmul_rn(__dsub_rn(__dmul_rn(__dadd_rn(__dmul_rn(__dsub_rn(x, 5.0), x), 8.0), x), 4.0), -0.5);
    return 0.0;
}

__global__ void __launch_bounds__(kPrepThreads)
prepare_images_pil_kernel(const unsigned char* __restrict__ src, float* __restrict__ out, unsigned char* __restrict__ out_u8,
                          int H, int W, int h, int w, long long sb, long long sc, long long sy, long long sx,
                          float m0, float m1, float m2, float s0, float s1, float s2, int tile_w, int tile_h) {
    extern __shared__ unsigned char pil_smem[];
    __shared__ float lut[3][256];
    __shared__ double wraw[2][kPrepTile][kPrepMaxTaps];
    __shared__ int kq[2][kPrepTile][kPrepMaxTaps];
    __shared__ int first[2][kPrepTile], taps[2][kPrepTile];
    unsigned char* tile = pil_smem;                                   // [3][tile_h][tile_w]
    unsigned char* tmp = pil_smem + 3 * tile_h * tile_w;              // [3][tile_h][16]
    const int tid = threadIdx.x, o = tid / kPrepTile, jj = tid % kPrepTile;
    const int b = blockIdx.z;
    const int ox0 = blockIdx.x * kPrepTile, oy0 = blockIdx.y * kPrepTile;
    const int in_size[2] = {W, H}, out_size[2] = {w, h}, o0[2] = {ox0, oy0};
    if (out) {
        const float f = (float)tid / 255.f;
        lut[0][tid] = (f - m0) / s0; lut[1][tid] = (f - m1) / s1; lut[2][tid] = (f - m2) / s2;
    }
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        const int i = min(o0[d] + o, out_size[d] - 1);
        if (in_size[d] == out_size[d]) {                              // pass skipped by PIL: identity coefficient
            for (int j = jj; j < kPrepMaxTaps; j += kPrepTile) wraw[d][o][j] = j == 0 ? 1.0 : 0.0;
            if (jj == 0) { first[d][o] = i; taps[d][o] = 1; }
            continue;
        }
        const double scale = __ddiv_rn((double)in_size[d], (double)out_size[d]);
        const double filterscale = scale < 1.0 ? 1.0 : scale;
        const double support = __dmul_rn(2.0, filterscale), ss = __ddiv_rn(1.0, filterscale);
        const double center = __dmul_rn(__dadd_rn((double)i, 0.5), scale);
        int xmin = (int)__dadd_rn(__dsub_rn(center, support), 0.5);
        if (xmin < 0) xmin = 0;
        int xmax = (int)__dadd_rn(__dadd_rn(center, support), 0.5);
        if (xmax > in_size[d]) xmax = in_size[d];
        const int n = min(xmax - xmin, kPrepMaxTaps);
        for (int j = jj; j < kPrepMaxTaps; j += kPrepTile)
            wraw[d][o][j] = j < n ? pil_bicubic(__dmul_rn(__dadd_rn(__dsub_rn((double)(j + xmin), center), 0.5), ss)) : 0.0;
        if (jj == 0) { first[d][o] = xmin; taps[d][o] = n; }
    }
    __syncthreads();
#pragma unroll
    for (int d = 0; d < 2; ++d) {
        double ww = 0.0;
        const int n = taps[d][o];
        for (int j = 0; j < n; ++j) ww = __dadd_rn(ww, wraw[d][o][j]);
        for (int j = jj; j < kPrepMaxTaps; j += kPrepTile) {
            double wv = wraw[d][o][j];
            if (ww != 0.0) wv = __ddiv_rn(wv, ww);
            const double q = __dmul_rn(wv, (double)(1 << kPilBits));
            kq[d][o][j] = wv < 0.0 ? (int)__dadd_rn(-0.5, q) : (int)__dadd_rn(0.5, q);
        }
    }
    const int last = kPrepTile - 1;
    const int xlo = first[0][0], ylo = first[1][0];
    const int tw = min(first[0][last] + taps[0][last] - xlo, tile_w), th = min(first[1][last] + taps[1][last] - ylo, tile_h);
    const unsigned char* sbp = src + (long long)b * sb;
    for (int r = tid / 32; r < 3 * th; r += kPrepThreads / 32) {
        const int c = r / th, y = r - c * th;
        const unsigned char* srow = sbp + c * sc + (long long)(ylo + y) * sy + (long long)xlo * sx;
        unsigned char* trow = tile + (c * tile_h + y) * tile_w;
        for (int x = tid % 32; x < tw; x += 32) trow[x] = srow[x * sx];
    }
    __syncthreads();
    {
        const int ox = jj, x0 = first[0][ox] - xlo, nx = taps[0][ox];
        const int* kx = kq[0][ox];
        for (int r = o; r < 3 * th; r += kPrepTile) {
            const int c = r / th, y = r - c * th;
            const unsigned char* row = tile + (c * tile_h + y) * tile_w + x0;
            int acc = 1 << (kPilBits - 1);
            for (int i = 0; i < nx; ++i) acc += (int)row[i] * kx[i];
            tmp[(c * tile_h + y) * kPrepTile + ox] = (unsigned char)min(max(acc >> kPilBits, 0), 255);
        }
    }
    __syncthreads();
    const int ox = ox0 + jj, oy = oy0 + o;
    if (ox >= w || oy >= h) return;
    const int y0 = first[1][o] - ylo, ny = taps[1][o];
    const int* ky = kq[1][o];
    const long long obase = (long long)b * 3 * h * w + (long long)oy * w + ox;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const unsigned char* col = tmp + (c * tile_h + y0) * kPrepTile + jj;
        int acc = 1 << (kPilBits - 1);
        for (int j = 0; j < ny; ++j) acc += (int)col[j * kPrepTile] * ky[j];
        const int v = min(max(acc >> kPilBits, 0), 255);
        if (out) out[obase + (long long)c * h * w] = lut[c][v];
        if (out_u8) out_u8[obase + (long long)c * h * w] = (unsigned char)v;
    }
}

}  // namespace roma

extern "C" int roma_prepare_images(const unsigned char* images, float* out, int B, int H, int W, int h, int w,
                                   const float* mean3, const float* std3, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(images && out && mean3 && std3, ROMA_ERR_BAD_ARG, "roma_prepare_images: null pointer");
    ROMA_REQUIRE(B > 0 && H > 0 && W > 0 && h > 0 && w > 0, ROMA_ERR_BAD_ARG, "roma_prepare_images: bad sizes");
    ROMA_REQUIRE(std3[0] != 0.f && std3[1] != 0.f && std3[2] != 0.f, ROMA_ERR_BAD_ARG, "roma_prepare_images: zero std");
    const float sx = (float)W / (float)w, sy = (float)H / (float)h;
    const float smax = sx > sy ? sx : sy;
    ROMA_REQUIRE(smax <= 4.f, ROMA_ERR_UNSUPPORTED,                  // taps per dimension <= 4 * scale + 1 <= 17
                 "roma_prepare_images: downscale factor %.2f > 4 is not supported (resize in two steps)", smax);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    if (h == H && w == W) {
        const long long plane = (long long)H * W, total = 3ll * B * plane;
        const int vec = plane % 4 == 0 && ((reinterpret_cast<uintptr_t>(images) & 3u) | (reinterpret_cast<uintptr_t>(out) & 15u)) == 0;
        long long blocks = ((vec ? total / 4 : total) + 255) / 256;
        const long long cap = (long long)sm_count() * 16;
        if (blocks > cap) blocks = cap;
        normalise_images_kernel<<<(unsigned)blocks, 256, 0, st>>>(images, out, plane, total, mean3[0], mean3[1], mean3[2],
                                                                 std3[0], std3[1], std3[2], vec);
        return check_launch("roma_prepare_images");
    }
    ROMA_REQUIRE(B <= 65535, ROMA_ERR_UNSUPPORTED, "roma_prepare_images: batch too large");
    // input span of 16 consecutive outputs: <= 15 * scale + 2 * support + 2
    auto span = [](float sc) { const float sup = sc >= 1.f ? 2.f * sc : 2.f; return (int)ceilf(15.f * sc + 2.f * sup) + 3; };
    const int tile_w = span(sx), tile_h = span(sy);
    const size_t smem = (size_t)(3 * tile_w * tile_h + 3 * tile_h * kPrepTile) * sizeof(float);
    cudaError_t e = cudaFuncSetAttribute(prepare_images_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "roma_prepare_images: smem opt-in (%zu B): %s", smem, cudaGetErrorString(e));
    const dim3 grid((unsigned)cdiv(w, kPrepTile), (unsigned)cdiv(h, kPrepTile), (unsigned)B);
    prepare_images_kernel<<<grid, kPrepThreads, smem, st>>>(images, out, B, H, W, h, w, mean3[0], mean3[1], mean3[2],
                                                                     std3[0], std3[1], std3[2], tile_w, tile_h);
    return check_launch("roma_prepare_images");
}

extern "C" int roma_prepare_images_pil(const unsigned char* images, float* out, unsigned char* out_u8, int B, int H, int W,
                                       int h, int w, long long stride_b, long long stride_c, long long stride_y,
                                       long long stride_x, const float* mean3, const float* std3, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(images && (out || out_u8), ROMA_ERR_BAD_ARG, "roma_prepare_images_pil: null pointer");
    ROMA_REQUIRE(!out || (mean3 && std3), ROMA_ERR_BAD_ARG, "roma_prepare_images_pil: mean / std required for the fp32 output");
    ROMA_REQUIRE(B > 0 && B <= 65535 && H > 0 && W > 0 && h > 0 && w > 0, ROMA_ERR_BAD_ARG, "roma_prepare_images_pil: bad sizes");
    ROMA_REQUIRE(!out || (std3[0] != 0.f && std3[1] != 0.f && std3[2] != 0.f), ROMA_ERR_BAD_ARG, "roma_prepare_images_pil: zero std");
    const double sx = (double)W / w, sy = (double)H / h;
    ROMA_REQUIRE(sx <= 4.0 && sy <= 4.0, ROMA_ERR_UNSUPPORTED,       // taps per dimension <= 4 * scale + 1 <= 17
                 "roma_prepare_images_pil: downscale factor %.2f > 4 is not supported", sx > sy ? sx : sy);
    // input span of 16 consecutive outputs: <= 15 * scale + 2 * support + 1 (a skipped pass: 16)
    auto span = [](double sc) { const double sup = sc >= 1.0 ? 2.0 * sc : 2.0; return (int)ceil(15.0 * sc + 2.0 * sup) + 3; };
    const int tile_w = W == w ? kPrepTile : span(sx), tile_h = H == h ? kPrepTile : span(sy);
    const size_t smem = (size_t)(3 * tile_w * tile_h + 3 * tile_h * kPrepTile);
    cudaError_t e = cudaFuncSetAttribute(prepare_images_pil_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return fail((int)e, "roma_prepare_images_pil: smem opt-in (%zu B): %s", smem, cudaGetErrorString(e));
    const float zero3[3] = {0.f, 0.f, 0.f}, one3[3] = {1.f, 1.f, 1.f};
    const float* m = out ? mean3 : zero3;
    const float* sd = out ? std3 : one3;
    const dim3 grid((unsigned)cdiv(w, kPrepTile), (unsigned)cdiv(h, kPrepTile), (unsigned)B);
    prepare_images_pil_kernel<<<grid, kPrepThreads, smem, reinterpret_cast<cudaStream_t>(stream)>>>(
        images, out, out_u8, H, W, h, w, stride_b, stride_c, stride_y, stride_x, m[0], m[1], m[2], sd[0], sd[1], sd[2], tile_w, tile_h);
    return check_launch("roma_prepare_images_pil");
}
