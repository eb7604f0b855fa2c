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

__global__ void kde_finalize_kernel(const float* __restrict__ partial, int N, int nsplit, int half,
                                    float* __restrict__ density, float* __restrict__ p_out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    float s = 0.f;
    for (int k = 0; k < nsplit; ++k) s += partial[(long long)k * N + i];
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
static int kde_nsplit(int N) {
    // enough CTAs for >= 2 waves of 148 SMs x 4 resident CTAs
    const int qblocks = (N + kKdeThreads - 1) / kKdeThreads;
    int ns = (2 * 4 * sm_count() + qblocks - 1) / qblocks;
    if (ns < 1) ns = 1;
    if (ns > 64) ns = 64;
    return ns;
}
}  // namespace roma

extern "C" size_t roma_kde_workspace_bytes(int N) {
    if (N <= 0) return 0;
    return (size_t)64 * (size_t)N * sizeof(float);
}

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
    float* partial = reinterpret_cast<float*>(workspace);
    dim3 grid((N + kKdeThreads - 1) / kKdeThreads, nsplit);
    if (half) kde_partial_kernel<true><<<grid, kKdeThreads, 0, st>>>(x, N, M, down, inv_two_var, partial, nsplit);
    else      kde_partial_kernel<false><<<grid, kKdeThreads, 0, st>>>(x, N, M, down, inv_two_var, partial, nsplit);
    int rc = check_launch("roma_kde(partial)");
    if (rc) return rc;
    kde_finalize_kernel<<<(N + 255) / 256, 256, 0, st>>>(partial, N, nsplit, half, density, p_out);
    return check_launch("roma_kde(finalize)");
}
