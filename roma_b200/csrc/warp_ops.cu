// K3a: streaming kernels on the warp / certainty fields.
//   roma_flow_update_resample — matcher.py:397-402 (update) fused with :349-360 / :407-417 (bilinear resample)
//   roma_match_epilogue       — matcher.py:656-662 and :684-730 (attenuation, sigmoid, range mask, clamp, layout)
// Both are pure HBM streaming: every tensor is read once and written once; the resample phase
// re-derives the updated source values from (flow_in, delta) instead of re-reading flow_out, so the
// update and the resample run as ONE launch with no dependency between its two index ranges.
#include "common.cuh"

namespace roma {

struct UpdParams {
    const float* flow_in; roma_strides4 fs;
    const float* cert_in; roma_strides4 cs;
    const float* delta; roma_strides4 ds;
    float ins, den_x, den_y;
    float* flow_out; float* cert_out; float* flow_next; float* cert_next;
    int B, h, w, h2, w2;
    int from_out;           // phase 2 reads the ALREADY UPDATED fields (flow_out / cert_out of an earlier launch, L2-resident)
                            // instead of re-deriving them from (flow_in, delta): 12 loads per output instead of 24 loads + 12 divisions
    float rh, rw;           // in/out ratios for the resample
    long long n_src, n_dst;
};

__device__ __forceinline__ float3 updated_at(const UpdParams& p, int b, int y, int x) {
    if (p.from_out) {
        const long long i = ((long long)b * p.h + y) * p.w + x;
        const float2 f2 = reinterpret_cast<const float2*>(p.flow_out)[i];
        return make_float3(f2.x, f2.y, p.cert_out[i]);
    }
    const float* f = p.flow_in + (long long)b * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
    float fx = __ldg(f), fy = __ldg(f + p.fs.c);
    float c = p.cert_in ? __ldg(p.cert_in + (long long)b * p.cs.b + (long long)y * p.cs.y + (long long)x * p.cs.x) : 0.f;
    if (p.delta) {
        const float* d = p.delta + (long long)b * p.ds.b + (long long)y * p.ds.y + (long long)x * p.ds.x;
        // matcher.py:397-399: displacement = ins*(d/(4*w)); flow = flow + displacement.  Same op order
        // (div, mul, add; no fma contraction) so the update is bit-identical to the reference sequence.
        fx = __fadd_rn(fx, __fmul_rn(p.ins, __fdiv_rn(__ldg(d), p.den_x)));
        fy = __fadd_rn(fy, __fmul_rn(p.ins, __fdiv_rn(__ldg(d + p.ds.c), p.den_y)));
        c = __fadd_rn(c, __ldg(d + 2 * p.ds.c));
    }
    return make_float3(fx, fy, c);
}

// IDX = unsigned when every index fits 32 bits (the emulated 64-bit divisions were a third of the instructions), else long long
template <typename IDX>
__global__ void __launch_bounds__(256) flow_update_resample_kernel(const UpdParams p) {
    const IDX stride = (IDX)gridDim.x * blockDim.x;
    const IDX n_src = (IDX)p.n_src, n_all = (IDX)(p.n_src + p.n_dst);
    for (IDX idx = (IDX)blockIdx.x * blockDim.x + threadIdx.x; idx < n_all; idx += stride) {
        if (idx < n_src) {
            // phase 1: updated values at this scale
            const IDX hw = (IDX)(p.h * p.w);
            const int b = (int)(idx / hw);
            const int rem = (int)(idx - (IDX)b * hw);
            const int y = rem / p.w, x = rem - y * p.w;
            const float3 v = updated_at(p, b, y, x);
            if (p.flow_out) reinterpret_cast<float2*>(p.flow_out)[idx] = make_float2(v.x, v.y);
            if (p.cert_out) p.cert_out[idx] = v.z;
        } else {
            // phase 2: bilinear resample (ATen upsample_bilinear2d, align_corners=False) of the updated fields
            const IDX o = idx - n_src;
            const IDX hw2 = (IDX)(p.h2 * p.w2);
            const int b = (int)(o / hw2);
            const int rem = (int)(o - (IDX)b * hw2);
            const int y2 = rem / p.w2, x2 = rem - y2 * p.w2;
            int ya, yb, xa, xb; float ly, lx;
            bilinear_src(y2, p.rh, p.h, ya, yb, ly);
            bilinear_src(x2, p.rw, p.w, xa, xb, lx);
            const float hy = 1.f - ly, hx = 1.f - lx;
            const float3 v00 = updated_at(p, b, ya, xa), v01 = updated_at(p, b, ya, xb);
            const float3 v10 = updated_at(p, b, yb, xa), v11 = updated_at(p, b, yb, xb);
            const float ox = hy * (hx * v00.x + lx * v01.x) + ly * (hx * v10.x + lx * v11.x);
            const float oy = hy * (hx * v00.y + lx * v01.y) + ly * (hx * v10.y + lx * v11.y);
            const float oc = hy * (hx * v00.z + lx * v01.z) + ly * (hx * v10.z + lx * v11.z);
            if (p.flow_next) reinterpret_cast<float2*>(p.flow_next)[o] = make_float2(ox, oy);
            if (p.cert_next) p.cert_next[o] = oc;
        }
    }
}

struct EpiParams {
    const float* flow; roma_strides4 fs;
    const float* cert; roma_strides4 cs;
    const float* cert16; roma_strides4 c16s; int h16, w16;
    float* warp; float* certainty; float* logits;
    int B, H, W, symmetric;
    float rh16, rw16;
};

template <typename IDX>
__global__ void __launch_bounds__(256) match_epilogue_kernel(const EpiParams p) {
    const int Wout = p.symmetric ? 2 * p.W : p.W;
    const IDX n = (IDX)((long long)p.B * p.H * Wout);
    const IDX stride = (IDX)gridDim.x * blockDim.x;
    for (IDX idx = (IDX)blockIdx.x * blockDim.x + threadIdx.x; idx < n; idx += stride) {
        const IDX hw = (IDX)(p.H * Wout);
        const int b = (int)(idx / hw);
        const int rem = (int)(idx - (IDX)b * hw);
        const int y = rem / Wout, xo = rem - y * Wout;
        const bool right = xo >= p.W;               // symmetric: right half holds B->A of decoder element b+B
        const int x = right ? xo - p.W : xo;
        const int e = right ? b + p.B : b;
        const float* f = p.flow + (long long)e * p.fs.b + (long long)y * p.fs.y + (long long)x * p.fs.x;
        float fx = __ldg(f), fy = __ldg(f + p.fs.c);
        float c = __ldg(p.cert + (long long)e * p.cs.b + (long long)y * p.cs.y + (long long)x * p.cs.x);
        if (p.cert16) {
            // low = 0.5 * up(cert16) * (up(cert16) < 0)   (matcher.py:657-662)
            int ya, yb, xa, xb; float ly, lx;
            bilinear_src(y, p.rh16, p.h16, ya, yb, ly);
            bilinear_src(x, p.rw16, p.w16, xa, xb, lx);
            const float* c16 = p.cert16 + (long long)e * p.c16s.b;
            const float v00 = __ldg(c16 + ya * p.c16s.y + xa * p.c16s.x), v01 = __ldg(c16 + ya * p.c16s.y + xb * p.c16s.x);
            const float v10 = __ldg(c16 + yb * p.c16s.y + xa * p.c16s.x), v11 = __ldg(c16 + yb * p.c16s.y + xb * p.c16s.x);
            const float up = (1.f - ly) * ((1.f - lx) * v00 + lx * v01) + ly * ((1.f - lx) * v10 + lx * v11);
            const float low = up < 0.f ? 0.5f * up : 0.f;
            c = c - low;
        }
        if (p.logits) p.logits[idx] = c;
        float prob = 1.f / (1.f + expf(-c));
        if (fabsf(fx) > 1.f || fabsf(fy) > 1.f) prob = 0.f;      // matcher.py:708-710
        fx = fminf(fmaxf(fx, -1.f), 1.f);                        // matcher.py:711
        fy = fminf(fmaxf(fy, -1.f), 1.f);
        const float gx = pixel_center(x, p.W), gy = pixel_center(y, p.H);
        // matcher.py:714-717: left half (grid, A->B); right half (B->A, grid)
        const float4 o = right ? make_float4(fx, fy, gx, gy) : make_float4(gx, gy, fx, fy);
        reinterpret_cast<float4*>(p.warp)[idx] = o;
        p.certainty[idx] = prob;
    }
}

static inline int stream_grid(long long n) {
    long long blocks = (n + 255) / 256;
    const long long cap = (long long)sm_count() * 8;   // 8 resident CTAs of 256 threads per SM
    return (int)(blocks < cap ? (blocks < 1 ? 1 : blocks) : cap);
}

}  // namespace roma

extern "C" int roma_flow_update_resample(const float* flow_in, roma_strides4 fs,
                                         const float* cert_in, roma_strides4 cs,
                                         const float* delta, roma_strides4 ds,
                                         float ins, float den_x, float den_y,
                                         float* flow_out, float* cert_out,
                                         float* flow_next, float* cert_next,
                                         int B, int h, int w, int h2, int w2, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(flow_in, ROMA_ERR_BAD_ARG, "roma_flow_update_resample: flow_in is null");
    ROMA_REQUIRE(B > 0 && h > 0 && w > 0 && h2 >= 0 && w2 >= 0, ROMA_ERR_BAD_ARG, "roma_flow_update_resample: bad sizes");
    const bool resample = (flow_next || cert_next);
    ROMA_REQUIRE(!resample || (h2 > 0 && w2 > 0), ROMA_ERR_BAD_ARG, "roma_flow_update_resample: resample requested with empty target size");
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(flow_out) | reinterpret_cast<uintptr_t>(flow_next)) & 7u) == 0, ROMA_ERR_ALIGNMENT,
                 "roma_flow_update_resample: flow outputs must be 8-byte aligned");
    UpdParams p;
    p.flow_in = flow_in; p.fs = fs; p.cert_in = cert_in; p.cs = cs; p.delta = delta; p.ds = ds;
    p.ins = ins; p.den_x = den_x; p.den_y = den_y; p.flow_out = flow_out; p.cert_out = cert_out; p.flow_next = flow_next; p.cert_next = cert_next;
    p.B = B; p.h = h; p.w = w; p.h2 = h2; p.w2 = w2;
    p.rh = h2 > 0 ? (float)h / (float)h2 : 1.f;
    p.rw = w2 > 0 ? (float)w / (float)w2 : 1.f;
    p.n_src = (flow_out || cert_out) ? (long long)B * h * w : 0;
    p.n_dst = resample ? (long long)B * h2 * w2 : 0;
    if (p.n_src + p.n_dst == 0) return ROMA_OK;
    p.from_out = 0;
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    auto launch = [&](const UpdParams& q) {
        if (q.n_src + q.n_dst < (1ll << 31)) flow_update_resample_kernel<unsigned><<<stream_grid(q.n_src + q.n_dst), 256, 0, st>>>(q);
        else flow_update_resample_kernel<long long><<<stream_grid(q.n_src + q.n_dst), 256, 0, st>>>(q);
    };
    if (delta && flow_out && cert_out && resample && p.n_dst >= p.n_src) {
        // upsampling with both current outputs requested: update first, then resample FROM the updated fields (same values,
        // bit for bit: phase 2 of the fused form recomputes exactly what phase 1 stores)
        UpdParams a = p; a.n_dst = 0;
        UpdParams b2 = p; b2.n_src = 0; b2.from_out = 1;
        launch(a);
        launch(b2);
        return check_launch("roma_flow_update_resample", 2);
    }
    launch(p);
    return check_launch("roma_flow_update_resample");
}

extern "C" int roma_match_epilogue(const float* flow, roma_strides4 fs,
                                   const float* cert, roma_strides4 cs,
                                   const float* cert16, roma_strides4 c16s, int h16, int w16,
                                   float* warp, float* certainty, float* logits_out,
                                   int B, int H, int W, int symmetric, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(flow && cert && warp && certainty, ROMA_ERR_BAD_ARG, "roma_match_epilogue: null pointer");
    ROMA_REQUIRE(B > 0 && H > 0 && W > 0, ROMA_ERR_BAD_ARG, "roma_match_epilogue: bad sizes");
    ROMA_REQUIRE(!cert16 || (h16 > 0 && w16 > 0), ROMA_ERR_BAD_ARG, "roma_match_epilogue: cert16 given with empty size");
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(warp) & 15u) == 0, ROMA_ERR_ALIGNMENT, "roma_match_epilogue: warp must be 16-byte aligned");
    EpiParams p;
    p.flow = flow; p.fs = fs; p.cert = cert; p.cs = cs; p.cert16 = cert16; p.c16s = c16s; p.h16 = h16; p.w16 = w16;
    p.warp = warp; p.certainty = certainty; p.logits = logits_out;
    p.B = B; p.H = H; p.W = W; p.symmetric = symmetric ? 1 : 0;
    p.rh16 = cert16 ? (float)h16 / (float)H : 1.f;
    p.rw16 = cert16 ? (float)w16 / (float)W : 1.f;
    const long long n = (long long)B * H * (symmetric ? 2 * W : W);
    if (n < (1ll << 31)) match_epilogue_kernel<unsigned><<<stream_grid(n), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p);
    else match_epilogue_kernel<long long><<<stream_grid(n), 256, 0, reinterpret_cast<cudaStream_t>(stream)>>>(p);
    return check_launch("roma_match_epilogue");
}
