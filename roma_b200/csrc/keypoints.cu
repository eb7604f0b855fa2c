// K3c: match_keypoints (romatch/models/matcher.py:576-591) — warp the keypoints of A into B by bilinear
// sampling of the dense warp, then mutual nearest neighbours against the keypoints of B.
// Small problem (Na, Nb ~ 1e3..1e4): three O(Na*Nb) passes that all evaluate the SAME fp32 distance
// expression, so the equality tests D == rowmin and D == colmin of the reference are exact here too.
#include "common.cuh"

#include <mutex>

namespace roma {
namespace kp {

struct Ws { float* xab; float* certa; float* rowmin; float* colmin; int* cnt; size_t bytes; };

static Ws carve(void* base, int Na, int Nb) {
    Ws w; size_t off = 0;
    auto take = [&](size_t bytes) { char* p = base ? reinterpret_cast<char*>(base) + off : nullptr; off = (off + bytes + 255) / 256 * 256; return p; };
    w.xab = reinterpret_cast<float*>(take((size_t)Na * 2 * sizeof(float)));
    w.certa = reinterpret_cast<float*>(take((size_t)Na * sizeof(float)));
    w.rowmin = reinterpret_cast<float*>(take((size_t)Na * sizeof(float)));
    w.colmin = reinterpret_cast<float*>(take((size_t)Nb * sizeof(float)));
    w.cnt = reinterpret_cast<int*>(take((size_t)(Na + 1) * sizeof(int)));
    w.bytes = off;
    return w;
}

__device__ __forceinline__ float dist(float ax, float ay, float bx, float by) {
    const float dx = ax - bx, dy = ay - by;
    return sqrtf(__fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy)));
}

// bilinear sample (zeros padding, align_corners=False) of warp[..., 2:4] and certainty at x_A
__global__ void warp_points_kernel(const float* __restrict__ xA, int Na, const float* __restrict__ warp,
                                   const float* __restrict__ cert, int H, int Wp, float* __restrict__ xab, float* __restrict__ certa) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= Na) return;
    const float ix = unnormalize(xA[2 * a], Wp), iy = unnormalize(xA[2 * a + 1], H);
    const float fx = floorf(fminf(fmaxf(ix, -1048576.f), 1048576.f)), fy = floorf(fminf(fmaxf(iy, -1048576.f), 1048576.f));
    const int x0 = (int)fx, y0 = (int)fy;
    const float wx1 = ix - fx, wx0 = (fx + 1.f) - ix, wy1 = iy - fy, wy0 = (fy + 1.f) - iy;
    float ox = 0.f, oy = 0.f, oc = 0.f;
#pragma unroll
    for (int t = 0; t < 4; ++t) {
        const int yy = y0 + (t >> 1), xx = x0 + (t & 1);
        if (yy >= 0 && yy < H && xx >= 0 && xx < Wp) {
            const float wgt = ((t & 1) ? wx1 : wx0) * ((t >> 1) ? wy1 : wy0);
            const float4 v = reinterpret_cast<const float4*>(warp)[(long long)yy * Wp + xx];
            ox = fmaf(v.z, wgt, ox); oy = fmaf(v.w, wgt, oy);
            oc = fmaf(cert[(long long)yy * Wp + xx], wgt, oc);
        }
    }
    xab[2 * a] = ox; xab[2 * a + 1] = oy; certa[a] = oc;
}

// out[i] = min_j dist(P[i], Q[j])   (one thread per i, Q streamed through shared memory)
__global__ void __launch_bounds__(256) min_dist_kernel(const float* __restrict__ P, int Np, const float* __restrict__ Q, int Nq,
                                                       float* __restrict__ out) {
    __shared__ float2 tile[1024];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float2 p = make_float2(0.f, 0.f);
    if (i < Np) p = reinterpret_cast<const float2*>(P)[i];
    float m = INFINITY;
    for (int j0 = 0; j0 < Nq; j0 += 1024) {
        const int n = min(1024, Nq - j0);
        __syncthreads();
        for (int t = threadIdx.x; t < n; t += blockDim.x) tile[t] = reinterpret_cast<const float2*>(Q)[j0 + t];
        __syncthreads();
        if (i < Np)
            for (int t = 0; t < n; ++t) m = fminf(m, dist(p.x, p.y, tile[t].x, tile[t].y));
    }
    if (i < Np) out[i] = m;
}

// pass 1 (emit == false): cnt[a] = #b with D==rowmin[a] && D==colmin[b] && cert>thresh ; pass 2: write pairs
template <bool kEmit>
__global__ void __launch_bounds__(256) mutual_kernel(const float* __restrict__ xab, const float* __restrict__ certa, int Na,
                                                     const float* __restrict__ xB, int Nb, const float* __restrict__ rowmin,
                                                     const float* __restrict__ colmin, float thresh, int* __restrict__ cnt,
                                                     int64_t* __restrict__ indsA, int64_t* __restrict__ indsB, int capacity) {
    const int a = blockIdx.x * blockDim.x + threadIdx.x;
    if (a >= Na) return;
    int n = 0;
    if (certa[a] > thresh) {
        const float ax = xab[2 * a], ay = xab[2 * a + 1], rm = rowmin[a];
        const int base = kEmit ? cnt[a] : 0;
        for (int b = 0; b < Nb; ++b) {
            const float d = dist(ax, ay, __ldg(xB + 2 * b), __ldg(xB + 2 * b + 1));
            if (d == rm && d == __ldg(colmin + b)) {
                if (kEmit && base + n < capacity) { indsA[base + n] = a; indsB[base + n] = b; }
                ++n;
            }
        }
    }
    if (!kEmit) cnt[a] = n;
}

// exclusive scan of cnt[0..Na) in place by one CTA; cnt[Na] and *count receive the total
__global__ void __launch_bounds__(1024) scan_kernel(int* __restrict__ cnt, int Na, int* __restrict__ count) {
    __shared__ int warp_tot[32];
    __shared__ int carry;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < Na; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < Na ? cnt[i] : 0;
        int s = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, s, o); if ((threadIdx.x & 31) >= o) s += t; }
        if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = s;
        __syncthreads();
        if (threadIdx.x < 32) {
            int w = warp_tot[threadIdx.x];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, w, o); if (threadIdx.x >= o) w += t; }
            warp_tot[threadIdx.x] = w;
        }
        __syncthreads();
        const int prefix = carry + ((threadIdx.x >> 5) ? warp_tot[(threadIdx.x >> 5) - 1] : 0) + s - v;
        if (i < Na) cnt[i] = prefix;
        __syncthreads();
        if (threadIdx.x == 0) carry += warp_tot[31];
        __syncthreads();
    }
    if (threadIdx.x == 0) { cnt[Na] = carry; *count = carry; }
}

}  // namespace kp
}  // namespace roma

extern "C" size_t roma_match_keypoints_workspace_bytes(int Na, int Nb) {
    if (Na <= 0 || Nb <= 0) return 0;
    return roma::kp::carve(nullptr, Na, Nb).bytes;
}

extern "C" int roma_match_keypoints(const float* x_A, int Na, const float* x_B, int Nb,
                                    const float* warp, const float* certainty, int H, int Wp,
                                    float thresh, int64_t* inds_A, int64_t* inds_B, int capacity, int* count,
                                    void* workspace, size_t workspace_bytes, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x_A && x_B && warp && certainty && inds_A && inds_B && count && workspace, ROMA_ERR_BAD_ARG, "roma_match_keypoints: null pointer");
    ROMA_REQUIRE(Na > 0 && Nb > 0 && H > 0 && Wp > 0 && capacity >= 0, ROMA_ERR_BAD_ARG, "roma_match_keypoints: bad sizes");
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(warp) & 15u) == 0 && (reinterpret_cast<uintptr_t>(x_A) & 7u) == 0 &&
                 (reinterpret_cast<uintptr_t>(x_B) & 7u) == 0 && (reinterpret_cast<uintptr_t>(workspace) & 255u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_match_keypoints: warp needs 16-byte, keypoints 8-byte, workspace 256-byte alignment");
    kp::Ws w = kp::carve(workspace, Na, Nb);
    ROMA_REQUIRE(workspace_bytes >= w.bytes, ROMA_ERR_WORKSPACE, "roma_match_keypoints: workspace %zu < required %zu", workspace_bytes, w.bytes);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    kp::warp_points_kernel<<<cdiv(Na, 256), 256, 0, st>>>(x_A, Na, warp, certainty, H, Wp, w.xab, w.certa);
    kp::min_dist_kernel<<<cdiv(Na, 256), 256, 0, st>>>(w.xab, Na, x_B, Nb, w.rowmin);
    kp::min_dist_kernel<<<cdiv(Nb, 256), 256, 0, st>>>(x_B, Nb, w.xab, Na, w.colmin);
    kp::mutual_kernel<false><<<cdiv(Na, 256), 256, 0, st>>>(w.xab, w.certa, Na, x_B, Nb, w.rowmin, w.colmin, thresh, w.cnt, nullptr, nullptr, 0);
    kp::scan_kernel<<<1, 1024, 0, st>>>(w.cnt, Na, count);
    kp::mutual_kernel<true><<<cdiv(Na, 256), 256, 0, st>>>(w.xab, w.certa, Na, x_B, Nb, w.rowmin, w.colmin, thresh, w.cnt, inds_A, inds_B, capacity);
    return check_launch("roma_match_keypoints", 6);
}

// ---------------------------------------------------------------------------------------------------------------
// roma_match_keypoints_static: the same six launches with the scratch in a file-scope device array — for callers
// that may not allocate (CUDA graph capture, serving loops).  No counterpart in the reference.
// The array is per device (every GPU that loads the library holds ROMA_KP_STATIC_BYTES until exit); its device
// address is resolved ONCE per device, with that device current, by cudaGetSymbolAddress: in host code the
// symbol's name is a host-side stand-in, not a device pointer.
// ---------------------------------------------------------------------------------------------------------------
#define ROMA_KP_STATIC_MAX_POINTS 16384
#define ROMA_KP_STATIC_BYTES (512 * 1024)

static __device__ __align__(256) unsigned char g_kp_static_scratch[ROMA_KP_STATIC_BYTES];

namespace roma {
namespace kp {
static void* g_static_addr[64] = {};
static std::mutex g_static_mu;

// device address of g_kp_static_scratch on the CURRENT device (resolved once per device); nullptr + *err on failure
static void* static_scratch(int* device, cudaError_t* err) {
    *err = cudaGetDevice(device);
    if (*err != cudaSuccess) return nullptr;
    if (*device < 0 || *device >= 64) { *err = cudaErrorInvalidDevice; return nullptr; }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_addr[*device]) {
        void* p = nullptr;
        *err = cudaGetSymbolAddress(&p, g_kp_static_scratch);
        if (*err != cudaSuccess) return nullptr;
        g_static_addr[*device] = p;
    }
    return g_static_addr[*device];
}
}  // namespace kp
}  // namespace roma

extern "C" int roma_match_keypoints_static_max_points(void) { return ROMA_KP_STATIC_MAX_POINTS; }

extern "C" int roma_match_keypoints_static(const float* x_A, int Na, const float* x_B, int Nb,
                                           const float* warp, const float* certainty, int H, int Wp,
                                           float thresh, int64_t* inds_A, int64_t* inds_B, int capacity, int* count,
                                           void* stream) {
    using namespace roma;
    ROMA_REQUIRE(x_A && x_B && warp && certainty && inds_A && inds_B && count, ROMA_ERR_BAD_ARG, "roma_match_keypoints_static: null pointer");
    ROMA_REQUIRE(Na > 0 && Nb > 0 && H > 0 && Wp > 0 && capacity >= 0, ROMA_ERR_BAD_ARG, "roma_match_keypoints_static: bad sizes");
    ROMA_REQUIRE(Na <= ROMA_KP_STATIC_MAX_POINTS && Nb <= ROMA_KP_STATIC_MAX_POINTS, ROMA_ERR_UNSUPPORTED,
                 "roma_match_keypoints_static: Na = %d / Nb = %d exceed the static scratch limit of %d keypoints per image "
                 "(use roma_match_keypoints with a caller workspace)", Na, Nb, ROMA_KP_STATIC_MAX_POINTS);
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(warp) & 15u) == 0 && (reinterpret_cast<uintptr_t>(x_A) & 7u) == 0 &&
                 (reinterpret_cast<uintptr_t>(x_B) & 7u) == 0,
                 ROMA_ERR_ALIGNMENT, "roma_match_keypoints_static: warp needs 16-byte, keypoints 8-byte alignment");
    int device = -1;
    cudaError_t err = cudaSuccess;
    void* scratch = kp::static_scratch(&device, &err);
    ROMA_REQUIRE(scratch != nullptr, ROMA_ERR_UNSUPPORTED, "roma_match_keypoints_static: cudaGetSymbolAddress failed on device %d: %s",
                 device, cudaGetErrorString(err));
    kp::Ws w = kp::carve(scratch, Na, Nb);
    ROMA_REQUIRE(w.bytes <= (size_t)ROMA_KP_STATIC_BYTES, ROMA_ERR_WORKSPACE, "roma_match_keypoints_static: %zu scratch bytes > %d", w.bytes, ROMA_KP_STATIC_BYTES);
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    // every scratch entry read below is written first by an earlier launch of THIS call (xab, certa, rowmin, colmin,
    // cnt[0..Na]); nothing relies on the array's previous contents, so no clearing is needed between calls.
    kp::warp_points_kernel<<<cdiv(Na, 256), 256, 0, st>>>(x_A, Na, warp, certainty, H, Wp, w.xab, w.certa);
    kp::min_dist_kernel<<<cdiv(Na, 256), 256, 0, st>>>(w.xab, Na, x_B, Nb, w.rowmin);
    kp::min_dist_kernel<<<cdiv(Nb, 256), 256, 0, st>>>(x_B, Nb, w.xab, Na, w.colmin);
    kp::mutual_kernel<false><<<cdiv(Na, 256), 256, 0, st>>>(w.xab, w.certa, Na, x_B, Nb, w.rowmin, w.colmin, thresh, w.cnt, nullptr, nullptr, 0);
    kp::scan_kernel<<<1, 1024, 0, st>>>(w.cnt, Na, count);
    kp::mutual_kernel<true><<<cdiv(Na, 256), 256, 0, st>>>(w.xab, w.certa, Na, x_B, Nb, w.rowmin, w.colmin, thresh, w.cnt, inds_A, inds_B, capacity);
    return check_launch("roma_match_keypoints_static", 6);
}
