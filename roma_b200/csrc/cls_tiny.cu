// (f-1) cls_to_flow_refine (romatch/utils/utils.py:301-323) and
// KT  Tiny-RoMa fused corr_volume + pos_embed (romatch/models/tiny.py:241-254, 178-199).
#include "common.cuh"
#include "tc_gemm.cuh"

namespace roma {

// ------------------------------------------------------------------------------------------------
// cls_to_flow_refine: one warp per pixel.  softmax -> mode -> 5-tap weighted mean where the softmax
// normaliser cancels: flow = sum_i e_i G[idx_i] / sum_i e_i with e_i = exp(l_i - l_max), so a single
// streaming argmax pass over the res^2 logits is all that touches memory (52 MB/pair at 560^2 would
// otherwise be read ~4 times by softmax + max + gather).
// Ties in the argmax resolve to the lowest index.
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
cls_to_flow_kernel(const T* __restrict__ cls, roma_strides4 s, float* __restrict__ out, int B, int C, int H, int W, int res) {
    const long long pix = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    const long long npix = (long long)B * H * W;
    if (pix >= npix) return;
    const int b = (int)(pix / (H * W));
    const int rem = (int)(pix - (long long)b * H * W);
    const int y = rem / W, x = rem - y * W;
    const T* p = cls + (long long)b * s.b + (long long)y * s.y + (long long)x * s.x;
    float best = -INFINITY; int bi = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
        const float v = Elem<T>::to_f(p[(long long)c * s.c]);
        if (v > best || bi == 0x7fffffff) { best = v; bi = c; }   // c increases per lane: first max kept
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
    }
    if (lane == 0) {
        const int idx[5] = {bi - 1, bi, bi + 1, bi - res, bi + res};
        float sx = 0.f, sy = 0.f, st = 0.f;
#pragma unroll
        for (int i = 0; i < 5; ++i) {
            const int c = min(max(idx[i], 0), C - 1);
            const float e = expf(Elem<T>::to_f(p[(long long)c * s.c]) - best);
            sx = fmaf(e, pixel_center(c % res, res), sx);
            sy = fmaf(e, pixel_center(c / res, res), sy);
            st += e;
        }
        reinterpret_cast<float2*>(out)[pix] = make_float2(sx / st, sy / st);
    }
}

// ------------------------------------------------------------------------------------------------
// Tiny-RoMa: one warp per image-0 pixel q.  Lanes stride over the image-1 positions j, each computing the
// C-channel dot product s_j = f0_q . f1_j / sqrt(C) (f0_q broadcast from shared memory), keeping
//   (a) the running argmax over ALL j (first index on ties), and
//   (b) an online softmax-expectation over the 4x4-strided subset (running max, sum, sum*grid_lr),
// then merges the extra logit float(argmax) (tiny.py:196) and writes E[pos].  The (H1*W1) x (H0*W0)
// volume (92 MB per 480x640 pair) is never materialised.
// ------------------------------------------------------------------------------------------------
template <int kC>
__global__ void __launch_bounds__(256)
tiny_corr_posembed_kernel(const float* __restrict__ f0, const float* __restrict__ f1, float* __restrict__ out,
                          int B, int H0, int W0, int H1, int W1) {
    __shared__ float q_s[8][kC];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long q = (long long)blockIdx.x * 8 + warp;
    const int n0 = H0 * W0, n1 = H1 * W1;
    if (q >= (long long)B * n0) return;
    const int b = (int)(q / n0);
    const int qi = (int)(q - (long long)b * n0);
    for (int c = lane; c < kC; c += 32) q_s[warp][c] = f0[q * kC + c];
    __syncwarp();
    const float inv_sqrt_c = rsqrtf((float)kC);
    const float* f1b = f1 + (long long)b * n1 * kC;
    const int down = 4;
    const int wl = W1 / down, hl = H1 / down;
    const float lx0 = -1.f + (float)down / W1, ly0 = -1.f + (float)down / H1;
    const float lxs = wl > 1 ? (2.f - 2.f * down / W1) / (wl - 1) : 0.f, lys = hl > 1 ? (2.f - 2.f * down / H1) / (hl - 1) : 0.f;

    float best = -INFINITY; int bi = 0x7fffffff;
    float m = -INFINITY, ssum = 0.f, sx = 0.f, sy = 0.f;   // online softmax over the strided subset
    for (int j = lane; j < n1; j += 32) {
        const float4* src = reinterpret_cast<const float4*>(f1b + (long long)j * kC);
        float acc = 0.f;
#pragma unroll
        for (int c4 = 0; c4 < kC / 4; ++c4) {
            const float4 v = __ldg(src + c4);
            acc = fmaf(q_s[warp][4 * c4], v.x, acc); acc = fmaf(q_s[warp][4 * c4 + 1], v.y, acc);
            acc = fmaf(q_s[warp][4 * c4 + 2], v.z, acc); acc = fmaf(q_s[warp][4 * c4 + 3], v.w, acc);
        }
        const float sj = acc * inv_sqrt_c;
        if (sj > best) { best = sj; bi = j; }          // j increases per lane: first max kept
        const int y1 = j / W1, x1 = j - y1 * W1;
        if ((y1 % down) == 0 && (x1 % down) == 0 && (y1 / down) < hl && (x1 / down) < wl) {
            const float gx = lx0 + lxs * (x1 / down), gy = ly0 + lys * (y1 / down);
            if (sj > m) { const float sc = expf(m - sj); ssum *= sc; sx *= sc; sy *= sc; m = sj; }
            const float e = expf(sj - m);
            ssum += e; sx = fmaf(e, gx, sx); sy = fmaf(e, gy, sy);
        }
    }
    // warp merge: argmax (lowest index on ties) and the online-softmax triples
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, best, o);
        const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
        if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
        const float om = __shfl_xor_sync(0xffffffffu, m, o);
        const float os = __shfl_xor_sync(0xffffffffu, ssum, o);
        const float ox = __shfl_xor_sync(0xffffffffu, sx, o);
        const float oy = __shfl_xor_sync(0xffffffffu, sy, o);
        const float nm = fmaxf(m, om);
        const float a = (m == -INFINITY) ? 0.f : expf(m - nm), c = (om == -INFINITY) ? 0.f : expf(om - nm);
        ssum = ssum * a + os * c; sx = sx * a + ox * c; sy = sy * a + oy * c; m = nm;
    }
    if (lane == 0) {
        // extra logit = float(argmax index) (tiny.py:196), weighted with grid[best]
        const float extra = (float)bi;
        const float nm = fmaxf(m, extra);
        const float a = (m == -INFINITY) ? 0.f : expf(m - nm), e = expf(extra - nm);
        const int by = bi / W1, bx = bi - by * W1;
        const float tot = ssum * a + e;
        const float px = (sx * a + e * pixel_center(bx, W1)) / tot;
        const float py = (sy * a + e * pixel_center(by, H1)) / tot;
        out[((long long)b * 2 + 0) * n0 + qi] = px;
        out[((long long)b * 2 + 1) * n0 + qi] = py;
    }
}


// Merge of the per-(tile half) records written by the EPI_TINY epilogue of the tcgen05 GEMM (tc_gemm.cuh): slots are
// visited in ascending column order, so a strict comparison keeps the first maximum like torch.argmax; the online
// softmax triples merge as in the warp merge above; then the index-as-logit term of tiny.py:196.
__global__ void __launch_bounds__(256)
tiny_merge_kernel(const float* __restrict__ part, float* __restrict__ out, int B, int n0, int slots, int H1, int W1) {
    const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (long long)B * n0) return;
    const int b = (int)(idx / n0), qi = (int)(idx - (long long)b * n0);
    float best = -INFINITY; int bi = 0x7fffffff;
    float m = -INFINITY, ssum = 0.f, sx = 0.f, sy = 0.f;
    const float* rec = part + (long long)b * slots * 6 * n0 + qi;
    for (int sl = 0; sl < slots; ++sl, rec += 6ll * n0) {
        const float ov = rec[0];
        const int oi = __float_as_int(rec[n0]);
        if (ov > best) { best = ov; bi = oi; }
        const float om = rec[2ll * n0], os = rec[3ll * n0], ox = rec[4ll * n0], oy = rec[5ll * n0];
        const float nm = fmaxf(m, om);
        const float a = (m == -INFINITY) ? 0.f : expf(m - nm), c = (om == -INFINITY) ? 0.f : expf(om - nm);
        ssum = ssum * a + os * c; sx = sx * a + ox * c; sy = sy * a + oy * c; m = nm;
    }
    const float extra = (float)bi;
    const float nm = fmaxf(m, extra);
    const float a = (m == -INFINITY) ? 0.f : expf(m - nm), e = expf(extra - nm);
    const int by = bi / W1, bx = bi - by * W1;
    const float tot = ssum * a + e;
    out[((long long)b * 2 + 0) * n0 + qi] = (sx * a + e * pixel_center(bx, W1)) / tot;
    out[((long long)b * 2 + 1) * n0 + qi] = (sy * a + e * pixel_center(by, H1)) / tot;
}

}  // namespace roma

extern "C" int roma_cls_to_flow_refine(const void* cls, roma_strides4 cs, int dtype,
                                       float* flow_out, int B, int C, int H, int W, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(cls && flow_out, ROMA_ERR_BAD_ARG, "roma_cls_to_flow_refine: null pointer");
    ROMA_REQUIRE(B > 0 && C > 0 && H > 0 && W > 0, ROMA_ERR_BAD_ARG, "roma_cls_to_flow_refine: bad sizes");
    int res = 1;
    while (res * res < C) ++res;
    ROMA_REQUIRE(res * res == C, ROMA_ERR_BAD_ARG, "roma_cls_to_flow_refine: C=%d is not a square", C);
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(flow_out) & 7u) == 0, ROMA_ERR_ALIGNMENT, "roma_cls_to_flow_refine: flow_out must be 8-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int grid = cdiv((long long)B * H * W * 32, 256);
    switch (dtype) {
        case ROMA_F32:  cls_to_flow_kernel<float><<<grid, 256, 0, st>>>((const float*)cls, cs, flow_out, B, C, H, W, res); break;
        case ROMA_F16:  cls_to_flow_kernel<__half><<<grid, 256, 0, st>>>((const __half*)cls, cs, flow_out, B, C, H, W, res); break;
        case ROMA_BF16: cls_to_flow_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)cls, cs, flow_out, B, C, H, W, res); break;
        default: return fail(ROMA_ERR_BAD_ARG, "roma_cls_to_flow_refine: unknown dtype %d", dtype);
    }
    return check_launch("roma_cls_to_flow_refine");
}

extern "C" int roma_tiny_corr_posembed(const float* feat0, const float* feat1, float* out,
                                       int B, int C, int H0, int W0, int H1, int W1, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(feat0 && feat1 && out, ROMA_ERR_BAD_ARG, "roma_tiny_corr_posembed: null pointer");
    ROMA_REQUIRE(B > 0 && H0 > 0 && W0 > 0 && H1 >= 4 && W1 >= 4, ROMA_ERR_BAD_ARG, "roma_tiny_corr_posembed: bad sizes");
    ROMA_REQUIRE(H1 % 4 == 0 && W1 % 4 == 0, ROMA_ERR_UNSUPPORTED,
                 "roma_tiny_corr_posembed: H1, W1 must be multiples of 4 (the reference's strided subset and grid_lr disagree otherwise)");
    ROMA_REQUIRE(((reinterpret_cast<uintptr_t>(feat0) | reinterpret_cast<uintptr_t>(feat1)) & 15u) == 0, ROMA_ERR_ALIGNMENT,
                 "roma_tiny_corr_posembed: features must be 16-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int grid = cdiv((long long)B * H0 * W0, 8);
    switch (C) {
        case 16: tiny_corr_posembed_kernel<16><<<grid, 256, 0, st>>>(feat0, feat1, out, B, H0, W0, H1, W1); break;
        case 24: tiny_corr_posembed_kernel<24><<<grid, 256, 0, st>>>(feat0, feat1, out, B, H0, W0, H1, W1); break;
        case 32: tiny_corr_posembed_kernel<32><<<grid, 256, 0, st>>>(feat0, feat1, out, B, H0, W0, H1, W1); break;
        case 64: tiny_corr_posembed_kernel<64><<<grid, 256, 0, st>>>(feat0, feat1, out, B, H0, W0, H1, W1); break;
        default: return fail(ROMA_ERR_UNSUPPORTED, "roma_tiny_corr_posembed: C=%d (supported: 16, 24, 32, 64)", C);
    }
    return check_launch("roma_tiny_corr_posembed");
}

// Tensor-core form of the same op: scores = f0 f1^T / sqrt(C) as a tcgen05 GEMM with the fp32 features split on the
// fly into three bf16 planes (fp32-grade products, tc_gemm.cuh), the arg-max / online-softmax reduction fused on the
// TMEM drain, one small merge kernel over the column tiles.  The (n0 x n1) volume is never written.
extern "C" size_t roma_tiny_corr_posembed_workspace_bytes(int B, int H0, int W0, int H1, int W1) {
    if (B <= 0 || H0 <= 0 || W0 <= 0 || H1 <= 0 || W1 <= 0) return 0;
    const long long n0 = (long long)H0 * W0, n1 = (long long)H1 * W1;
    const long long slots = 2 * ((n1 + roma::tc::BN - 1) / roma::tc::BN);
    return (size_t)((long long)B * slots * 6 * n0) * sizeof(float);
}

extern "C" int roma_tiny_corr_posembed_tc(const float* feat0, const float* feat1, float* out,
                                          int B, int C, int H0, int W0, int H1, int W1,
                                          void* workspace, size_t workspace_bytes, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(feat0 && feat1 && out && workspace, ROMA_ERR_BAD_ARG, "roma_tiny_corr_posembed_tc: null pointer");
    ROMA_REQUIRE(B > 0 && H0 > 0 && W0 > 0 && H1 >= 4 && W1 >= 4 && C > 0, ROMA_ERR_BAD_ARG, "roma_tiny_corr_posembed_tc: bad sizes");
    ROMA_REQUIRE(H1 % 4 == 0 && W1 % 4 == 0, ROMA_ERR_UNSUPPORTED,
                 "roma_tiny_corr_posembed_tc: H1, W1 must be multiples of 4 (the reference's strided subset and grid_lr disagree otherwise)");
    ROMA_REQUIRE(C % 4 == 0 && ((reinterpret_cast<uintptr_t>(feat0) | reinterpret_cast<uintptr_t>(feat1)) & 15u) == 0, ROMA_ERR_ALIGNMENT,
                 "roma_tiny_corr_posembed_tc: features must be 16-byte aligned rows (C %% 4 == 0)");
    ROMA_REQUIRE(workspace_bytes >= roma_tiny_corr_posembed_workspace_bytes(B, H0, W0, H1, W1), ROMA_ERR_WORKSPACE,
                 "roma_tiny_corr_posembed_tc: workspace too small");
    ROMA_REQUIRE(B <= 65535, ROMA_ERR_UNSUPPORTED, "roma_tiny_corr_posembed_tc: batch too large");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    const int n0 = H0 * W0, n1 = H1 * W1;
    tc::Launch L{};
    tc::Problem& P = L.p[0];
    P = tc::make_problem(feat0, feat1, nullptr, n0, n1, C, C, C, 0, (long long)n0 * C, (long long)n1 * C, 0, tc::EPI_TINY);
    const int down = 4, wl = W1 / down, hl = H1 / down;
    P.part = reinterpret_cast<float*>(workspace);
    P.scale = 1.0f / sqrtf((float)C);
    P.w1 = W1;
    P.lx0 = -1.f + (float)down / W1; P.ly0 = -1.f + (float)down / H1;
    P.lxs = wl > 1 ? (2.f - 2.f * down / W1) / (wl - 1) : 0.f;
    P.lys = hl > 1 ? (2.f - 2.f * down / H1) / (hl - 1) : 0.f;
    L.nprob = 1; L.tiles0 = P.tiles_m * P.tiles_n;
    if (int rc = tc::launch_gemm_nt<float, 3>(L, B, st, "roma_tiny_corr_posembed_tc(gemm)")) return rc;
    tiny_merge_kernel<<<cdiv((long long)B * n0, 256), 256, 0, st>>>(P.part, out, B, n0, 2 * P.tiles_n, H1, W1);
    return check_launch("roma_tiny_corr_posembed_tc(merge)");
}
