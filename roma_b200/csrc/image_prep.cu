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

// weights of output index i along one dimension; returns the first input index, n = number of taps
__device__ __forceinline__ int aa_weights(int i, int in_size, float scale, float support, float (&wt)[kPrepMaxTaps], int& n) {
    const float center = scale * ((float)i + 0.5f);
    const int xmin = max((int)(center - support + 0.5f), 0);
    n = min((int)(center + support + 0.5f), in_size) - xmin;
    n = min(n, kPrepMaxTaps);
    const float invscale = scale >= 1.f ? 1.f / scale : 1.f;
    const float xmc = (float)xmin - center;
    float total = 0.f;
#pragma unroll
    for (int j = 0; j < kPrepMaxTaps; ++j) {
        float wv = 0.f;
        if (j < n) { wv = cubic_aa(((float)j + xmc + 0.5f) * invscale); total += wv; }
        wt[j] = wv;
    }
    if (total != 0.f) {
#pragma unroll
        for (int j = 0; j < kPrepMaxTaps; ++j) wt[j] = wt[j] / total;
    }
    return xmin;
}

__global__ void __launch_bounds__(256)
prepare_images_kernel(const unsigned char* __restrict__ src, float* __restrict__ out, int B, int H, int W, int h, int w,
                      float m0, float m1, float m2, float s0, float s1, float s2) {
    __shared__ float lut[3][256];
    {
        const int v = threadIdx.x;                                    // blockDim.x == 256
        const float f = (float)v / 255.f;
        lut[0][v] = (f - m0) / s0; lut[1][v] = (f - m1) / s1; lut[2][v] = (f - m2) / s2;
    }
    __syncthreads();
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)B * h * w) return;
    const int ox = (int)(idx % w);
    const int oy = (int)((idx / w) % h);
    const int b = (int)(idx / ((long long)w * h));
    const unsigned char* sb = src + (long long)b * 3 * H * W;
    float* ob = out + (long long)b * 3 * h * w + (long long)oy * w + ox;
    if (h == H && w == W) {
#pragma unroll
        for (int c = 0; c < 3; ++c) ob[(long long)c * h * w] = lut[c][sb[((long long)c * H + oy) * W + ox]];
        return;
    }
    const float sx = (float)W / (float)w, sy = (float)H / (float)h;
    const float supx = sx >= 1.f ? 2.f * sx : 2.f, supy = sy >= 1.f ? 2.f * sy : 2.f;
    float wx[kPrepMaxTaps], wy[kPrepMaxTaps];
    int nx, ny;
    const int x0 = aa_weights(ox, W, sx, supx, wx, nx);
    const int y0 = aa_weights(oy, H, sy, supy, wy, ny);
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const unsigned char* sc = sb + (long long)c * H * W;
        float acc = 0.f;
        for (int j = 0; j < ny; ++j) {
            const unsigned char* row = sc + (long long)(y0 + j) * W + x0;
            float r = 0.f;
#pragma unroll
            for (int i = 0; i < kPrepMaxTaps; ++i)
                if (i < nx) r += lut[c][row[i]] * wx[i];
            acc += r * wy[j];
        }
        ob[(long long)c * h * w] = acc;
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
    const long long n = (long long)B * h * w;
    prepare_images_kernel<<<cdiv(n, 256), 256, 0, st>>>(images, out, B, H, W, h, w, mean3[0], mean3[1], mean3[2],
                                                        std3[0], std3[1], std3[2]);
    return check_launch("roma_prepare_images");
}
