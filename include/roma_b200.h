/*
 * roma_b200 — C ABI of the B200-native (sm_100a) RoMa matching hot path.
 *
 * The reference (techshoww/RoMa) is pure Python/PyTorch and has NO FFI or operator
 * registry: its "plugin API" is a set of Python call sites that all
 * resolve to ATen library calls.  Each entry point below replaces one such group of
 * call sites; the reference file:line it stands in for is cited on the declaration.
 * `INTEGRATION.md` shows the ctypes binding a maintainer of the reference would add.
 *
 * Conventions (all entry points)
 *  - plain pointers + sizes; device pointers unless stated; no torch types.
 *  - the CALLER owns every buffer (outputs and workspaces); the library never
 *    allocates or frees device memory and keeps no pointer across calls.
 *  - asynchronous on `stream` (a cudaStream_t passed as void*); no hidden syncs.
 *    One entry point (roma_gp_posterior, two-lane schedule) additionally uses a
 *    high-priority helper stream and events owned by the library (one set per device,
 *    created on first use): it forks from and joins back to `stream` inside the call,
 *    so for the caller the call is still ordered on `stream` alone.
 *  - return 0 on success, a negative ROMA_ERR_* for bad arguments, a positive
 *    cudaError_t for a failed launch.  `roma_last_error_string()` describes the
 *    last failure on the calling thread.  Nothing throws across the ABI.
 *  - there is NO CPU fallback and no non-sm_100 code path.
 *  - feature maps are channels-last: element (b, y, x, c) lives at
 *    ((b*h + y)*w + x)*C + c.  Flow / certainty / delta tensors carry explicit
 *    element strides so NCHW tensors and permuted views are consumed in place.
 */
#ifndef ROMA_B200_H
#define ROMA_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ROMA_B200_ABI_VERSION 1

/* feature / output element types */
#define ROMA_F32  0
#define ROMA_F16  1
#define ROMA_BF16 2

/* error codes (negative); positive return values are cudaError_t */
#define ROMA_OK               0
#define ROMA_ERR_BAD_ARG     -1   /* null pointer, non-positive size, unknown dtype   */
#define ROMA_ERR_UNSUPPORTED -2   /* shape/layout the kernels do not implement       */
#define ROMA_ERR_ALIGNMENT   -3   /* pointer or stride not aligned as documented     */
#define ROMA_ERR_WORKSPACE   -4   /* workspace too small (see *_workspace_bytes)     */
#define ROMA_ERR_ARCH        -5   /* device is not compute capability 10.x           */

int         roma_abi_version(void);
const char* roma_last_error_string(void);
/* 0 if `device` is an sm_100-class GPU usable by this library, else ROMA_ERR_ARCH / cudaError_t */
int         roma_check_device(int device);
/* number of kernels this library has launched in the calling process so far (monotonic; for bench accounting) */
unsigned long long roma_kernel_launches(void);

/* element strides of a (b, c, y, x)-indexed fp32 tensor */
typedef struct roma_strides4 {
    int64_t b, c, y, x;
} roma_strides4;

/* ---------------------------------------------------------------------------------------
 * K2  refiner gather: bilinear warp of f1 + local correlation + displacement embedding,
 *     written straight into the ConvRefiner input buffer `d`.
 *
 * Replaces (reference):
 *   romatch/models/matcher.py:109        x_hat = F.grid_sample(y, flow^T, bilinear, align_corners=False)
 *   romatch/models/matcher.py:111-120    in_displacement = flow - pixel_grid ; disp_emb(40/32*scale_factor*.)
 *   romatch/utils/local_correlation.py:4-48   local_correlation(x, y, r, flow)
 *   romatch/models/matcher.py:130,132    d = cat(x, x_hat, emb[, local_corr])
 *
 * f0 : (B, h, w, C) channels-last features of image A (query), dtype `dtype`
 * f1 : (B, h1, w1, C) channels-last features of image B (support), dtype `dtype`
 * flow : fp32, element (b, c, y, x) at flow[b*fs.b + c*fs.c + y*fs.y + x*fs.x]; c=0 is x, c=1 is y,
 *        normalised to [-1, 1] with align_corners=False semantics.
 * d  : (B, h, w, d_channels) channels-last, dtype `dtype`, dense.  Segments (channel offsets):
 *        [off_x,    off_x+C)      copy of f0           (skipped when off_x    < 0)
 *        [off_xhat, off_xhat+C)   warped f1            (skipped when off_xhat < 0)
 *        [off_emb,  off_emb+E)    disp_w @ (disp_scale*(flow - grid)) + disp_b  (skipped when off_emb < 0 or E == 0)
 *        [off_corr, off_corr+K)   local correlation, K=(2r+1)^2, k = ky*(2r+1)+kx   (skipped when r == 0 or off_corr < 0)
 * disp_w : fp32 (E, 2) row-major, disp_b : fp32 (E)   (host-visible DEVICE pointers)
 * Zero padding: taps outside f1 contribute 0 (grid_sample padding_mode="zeros").
 * The correlation is sum_c f0[c] * f1_sampled[c] / sqrt(C).
 */
/* Testing / profiling aid for the aligned 16-bit local-correlation case (C in {64, 128, 256, 512}).
 * 0 (default) = for r = 2, 3 and equally sized maps, 8 x 8 pixel tiles whose windows fit a shared patch of f1 (coherent
 *     flow) are computed as ONE tcgen05 GEMM of the tile's f0 rows against the patch (TMA-loaded once, TMEM accumulator;
 *     local_corr_tile.cuh); the remaining tiles, and every other case, take the per-pixel kernel of path 2;
 * 1 = SIMT FMA dot products with shuffle reductions;
 * 2 = per pixel, neighbour dot products as warp-level tensor-core MMAs (mma.sync m16n8k16, fragments filled straight
 *     from 16-byte global loads) for every pixel. */
int roma_local_corr_set_path(int path);
int roma_refiner_gather(const void* f0, const void* f1, const float* flow, roma_strides4 flow_strides,
                        const float* disp_w, const float* disp_b, float disp_scale,
                        void* d, int d_channels, int off_x, int off_xhat, int off_emb, int off_corr,
                        int B, int C, int h, int w, int h1, int w1, int E, int r, int dtype,
                        void* stream);
/* The same call with a workspace of roma_refiner_gather_workspace_bytes(B, h, w, r) bytes (16-byte aligned), which enables, for
 * 16-bit features on the aligned fast path:
 *   r = 2, 3: the shared-patch tile kernel of path 0 above — one byte per 8 x 8 pixel tile (the tile map: which tiles the tile
 *             kernel takes, so that the per-pixel kernel launched after it skips them);
 *   r = 7 on maps of <= 4096 positions (scale 16: 40 x 40, 48 x 48): the FULL correlation f0 . f1^T (B x hw x hw fp32 in the
 *             workspace: 164 MB at B = 16, 40 x 40) as one TMA-fed tcgen05 GEMM, whatever the flow; the per-pixel kernel then
 *             gathers its (2r+2)^2 = 256 products per pixel from its row instead of pulling 256 KB of features per pixel.
 * 0 for other radii.  Without a workspace (or through roma_refiner_gather) every pixel takes the per-pixel kernel.  Same
 * results either way up to the fp32 summation order of the correlations. */
size_t roma_refiner_gather_workspace_bytes(int B, int h, int w, int r);
int roma_refiner_gather_ws(const void* f0, const void* f1, const float* flow, roma_strides4 flow_strides,
                           const float* disp_w, const float* disp_b, float disp_scale,
                           void* d, int d_channels, int off_x, int off_xhat, int off_emb, int off_corr,
                           int B, int C, int h, int w, int h1, int w1, int E, int r, int dtype,
                           void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3a  flow / certainty update + bilinear resample to the next scale.
 *
 * Replaces (reference):
 *   romatch/models/matcher.py:397-402   flow += ins*(dx/(4*W_full), dy/(4*H_full)); certainty += dcert
 *   romatch/models/matcher.py:349-360, 407-417   F.interpolate(flow|certainty, size, bilinear, align_corners=False)
 *
 * flow_in (B,2,h,w), cert_in (B,1,h,w), delta (B,3,h,w: dx, dy, dcert) — fp32 with explicit strides.
 *   cert_in may be NULL (treated as 0: first scale of the coarse pass, matcher.py:347).
 *   delta may be NULL (pure resample, e.g. the 560^2 -> 108^2 entry of the upsample pass).
 * ins = scale (16, 8, ...), den_x = 4*W_full, den_y = 4*H_full: flow += ins*(delta/den)  (matcher.py:397-398,
 *   same operation order as the reference so the update is bit-identical)
 * flow_out (B,h,w,2) interleaved fp32, cert_out (B,h,w) fp32: updated values at this scale (may be NULL).
 * flow_next (B,h2,w2,2), cert_next (B,h2,w2): bilinear resample of the UPDATED values (may be NULL with h2=w2=0).
 */
int roma_flow_update_resample(const float* flow_in, roma_strides4 flow_strides,
                              const float* cert_in, roma_strides4 cert_strides,
                              const float* delta, roma_strides4 delta_strides,
                              float ins, float den_x, float den_y,
                              float* flow_out, float* cert_out,
                              float* flow_next, float* cert_next,
                              int B, int h, int w, int h2, int w2, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3a  match epilogue: certainty attenuation, sigmoid, out-of-range mask, clamp, grid concat
 *      and the symmetric side-by-side layout, in one pass and without a host sync.
 *
 * Replaces (reference): romatch/models/matcher.py:656-662 and 684-730.
 *
 * flow (Bd,2,H,W), cert (Bd,1,H,W) fp32 strided: finest-scale decoder outputs.  Bd = 2*B when
 *   `symmetric` (first B elements A->B, last B elements B->A, matcher.py:516-528) else B.
 * cert16 (Bd,1,h16,w16) fp32 strided or NULL (attenuate_cert False).
 * warp : fp32 (B, H, Wout, 4), Wout = 2*W when symmetric else W.
 * certainty : fp32 (B, H, Wout)   probabilities after sigmoid and masking.
 * logits_out : optional fp32 (B, H, Wout): cert - low (pre-sigmoid, pre-mask), for parity checks.
 */
int roma_match_epilogue(const float* flow, roma_strides4 flow_strides,
                        const float* cert, roma_strides4 cert_strides,
                        const float* cert16, roma_strides4 cert16_strides, int h16, int w16,
                        float* warp, float* certainty, float* logits_out,
                        int B, int H, int W, int symmetric, void* stream);

/* ---------------------------------------------------------------------------------------
 * K1  GP coarse matcher posterior mean.
 *
 * Replaces (reference):
 *   romatch/models/matcher.py:154-163   CosKernel: K = exp((x.y^T/(|x||y|+eps) - 1)/T)
 *   romatch/models/matcher.py:232-247,224   f = cos(8*pi*pos_conv(grid))
 *   romatch/models/matcher.py:249-264   mu = K_xy (K_yy + sigma I)^-1 f, returned as (b, D, h1, w1)
 *
 * x : (B, Nx, C) token-major features of image A (channels-last feature map flattened), dtype `dtype`
 * y : (B, Ny, C) same for image B; Ny = hy*wy
 * pos_w : fp32 (D, 2), pos_b : fp32 (D)   (GP.pos_conv weight/bias)
 * mu : fp32 (B, Nx, D) token-major == channels-last (B, D, h1, w1)
 * workspace: roma_gp_workspace_bytes(B, Nx, Ny, C, D) bytes, 256-byte aligned.
 */
size_t roma_gp_workspace_bytes(int B, int Nx, int Ny, int C, int D);
/* Testing / profiling aid: 0 = auto (default): tcgen05 tensor-core path when the shapes allow, fp32 operands of the solve
 * split into THREE bf16 planes (fp32-grade products; what parity with the reference needs on ill-conditioned K_yy);
 * 1 = force the fp32 SIMT path; 2 = require the tcgen05 path (3-plane solve; ROMA_ERR_UNSUPPORTED from roma_gp_posterior
 * when it cannot be used); 3 = tcgen05 path with a TWO-plane solve (3 instead of 6 MMAs per k-step, ~2^-16 relative
 * products: 20 % faster, measurably less accurate — opt-in). */
int roma_gp_set_path(int path);
/* Schedule of the tcgen05 factorisation: 1 (default) = lanes — the dependent chain of the blocked Cholesky (diag, which also
 * applies the rank-128 update of the previous block column to its own tile, and panel A) on a high-priority helper stream,
 * its feeders (update of the next block column, syrk of the remaining diagonal tiles) on a second one — both owned by the
 * library, one pair per device — and all right-hand-side work trailing on the caller's stream; ordered by events, forked
 * from and joined to the caller's stream inside the call; 0 = every launch on the caller's stream (left-looking). */
int roma_gp_set_lanes(int on);
/* 1 (default): the finished panels of the factorisation (L and the solved right-hand sides) are ALSO written as bf16 hi / lo planes
 * by the panel kernels' epilogue, and the update / syrk / posterior GEMMs read those planes through TMA with no conversion
 * in their main loop (gemm_planes_kernel: 3-stage ring, three / six plane pairs per k-block for the 2- / 3-plane solve);
 * 0: fp32 operands re-split on the fly at every visit (the round-1 form).  Needs the two-lane schedule; same products either way. */
int roma_gp_set_planes(int on);
/* The tensor-core building block of K1, exposed for unit tests and timing: batched C = A * B^T with
 * A (batch, M, K), B (batch, N, K) row-major in `dtype`, C (batch, M, N) fp32.  fp32 operands are split on the
 * fly into `split` (2 or 3) bf16 planes (3 / 6 tcgen05.mma per k-step, fp32 TMEM accumulate); bf16 / fp16
 * operands need split = 1 and are exact.  Rows must be 16-byte aligned. */
int roma_tc_gemm_nt(const void* A, const void* B, float* C, int batch, int M, int N, int K,
                    int dtype, int split, void* stream);
/* Testing aid: 1 (default) = 16-bit operands of the tcgen05 GEMMs (the GP Gram matrices, roma_tc_gemm_nt with
 * fp16 / bf16) are fed by TMA (cp.async.bulk.tensor, warp-specialised producer / MMA issuer); 0 = the
 * register-staged main loop.  Results are identical. */
int roma_tc_set_tma(int on);
int roma_gp_posterior(const void* x, const void* y, const float* pos_w, const float* pos_b,
                      float* mu, int B, int Nx, int Ny, int C, int D, int hy, int wy,
                      float T, float sigma_noise, float eps, int dtype,
                      void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3b  Gaussian kde over 4-D matches, streaming (no N x N temporary).
 *
 * Replaces (reference): romatch/utils/kde.py:4-12 (torch.cdist + exp + row-sum) and the
 *   p computation of romatch/models/matcher.py:490-491.
 *
 * x : fp32 (N, 4) query points; refs are x[::down] (down >= 1).
 * half != 0 reproduces the reference's fp16 rounding chain (kde.py:6-7), else fp32.
 * density : fp32 (N)  (holds the fp16-rounded value when half != 0)
 * p_out : optional fp32 (N): 1/(density+1), 1e-7 where density < 10   (matcher.py:490-491)
 * workspace: roma_kde_workspace_bytes(N) bytes (deterministic split-sum partials + the 128 KB table of the half
 *   path; a smaller buffer of at least N floats is accepted and only reduces parallelism / disables the table).
 */
size_t roma_kde_workspace_bytes(int N);
/* Half path selection.  0 = auto (default): the squared distance d2 = half(|a|^2 + |b|^2 - 2ab) of every pair comes
 * from a warp-level tensor-core product over the 6-wide augmented half vectors (what ATen's _euclidean_dist feeds
 * the GPU's HGEMM), the rest of the chain from a lookup table of its single half-valued argument.  Testing aids:
 * 1 = evaluate the whole op chain directly per pair with sequential fp32 FMAs; 2 = table + sequential FMAs
 * (bit-identical per pair to 1); 3 = as 0 but without the symmetric evaluation below.
 * Symmetric evaluation (mode 0, down == 1, N <= 65536): the kernel value of a pair is the same for (i, j) and (j, i),
 * so every pair is evaluated once and added to the row sum of i and to a per-warp-range column sum of j
 * (ceil(N / 128) x N fp32 partials in the workspace, fixed summation order, no atomics: deterministic). */
int roma_kde_set_direct(int direct);
int roma_kde(const float* x, int N, int down, double std, int half,
             float* density, float* p_out, void* workspace, size_t workspace_bytes, void* stream);

/* Testing aids for the exact half path (tests/test_kde_exact_gpu.py): d2(i, j) = half(x1_[i] . x2_[j]) of ATen's
 * _euclidean_dist for ALL pairs of a small point set (N <= 16384) by one candidate arithmetic form — 0 = sequential fp32
 * FMAs, 1 = mma.sync m16n8k8, 2 = mma.sync m16n8k16, 3 / 4 = 2 / 1 with the operand roles exchanged — written as half
 * bit patterns to out_half_bits (N x N uint16); and the 31745-entry table of the rest of the chain (clamp_min(0), sqrt,
 * **2, neg, / (2 std^2), exp, each rounded to half) as fp32 values indexed by the half bit pattern of d2 (lut: 32768 floats). */
int roma_kde_debug_d2(const float* x, int N, int variant, void* out_half_bits, void* stream);
int roma_kde_debug_lut(float* lut, double std, void* stream);

/* The k largest of n scores as int64 indices ordered by (value descending, index ascending); out_values (same dtype as
 * values) may be null.  Replaces (reference): the top-k inside both torch.multinomial(replacement=False) draws of
 * RegressionMatcher.sample (romatch/models/matcher.py:482-494; TinyRoMa.sample, tiny.py:314-325) — multinomial without
 * replacement is topk(p / q, k) with q ~ Exp(1), and ATen's multi-block radix select + stable descending sort of the
 * survivors yields exactly this order, ties included, so the sampled rows stay those of the reference (the exponential
 * draw itself stays on the torch generator: that is what consumes the Philox stream as the reference does).
 * values: n contiguous elements (dtype ROMA_F32 / ROMA_F16 / ROMA_BF16), 0 < k <= n < 2^31; no NaN.
 * workspace: roma_topk_workspace_bytes(n, k) bytes, 256-byte aligned (select state + 2 k 64-bit words + sort scratch;
 * the size query needs a CUDA device and returns 0 without one).  Allocation-free, stream-ordered. */
size_t roma_topk_workspace_bytes(long long n, long long k);
int roma_topk_desc(const void* values, int dtype, long long n, long long k, long long* out_indices, void* out_values,
                   void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * K3c  match_keypoints: warp keypoints of A into B, mutual nearest neighbours.
 *
 * Replaces (reference): romatch/models/matcher.py:576-591.
 *
 * x_A (Na,2), x_B (Nb,2) fp32 normalised coords; warp (H, Wp, 4) fp32; certainty (H, Wp) fp32.
 * The last two warp channels are sampled bilinearly at x_A (zeros padding, align_corners=False).
 * workspace: roma_match_keypoints_workspace_bytes(Na, Nb).
 * inds_A, inds_B: int64 arrays of `capacity` entries — mutual-NN pairs with cert > thresh, ordered by
 *   (a, b) like torch.nonzero; *count (device int32) receives the TOTAL number of pairs found (pairs
 *   beyond `capacity` are counted but not stored; without exact distance ties there are <= min(Na,Nb)).
 */
size_t roma_match_keypoints_workspace_bytes(int Na, int Nb);
int roma_match_keypoints(const float* x_A, int Na, const float* x_B, int Nb,
                         const float* warp, const float* certainty, int H, int Wp,
                         float thresh, int64_t* inds_A, int64_t* inds_B, int capacity, int* count,
                         void* workspace, size_t workspace_bytes, void* stream);

/* roma_match_keypoints_static: the same computation and outputs as roma_match_keypoints, but the call ALLOCATES
 * NOTHING and takes no workspace: its scratch is a file-scope `static __device__` array inside the library.
 * No counterpart in the reference: added for callers that cannot call
 * cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid allocation).
 * Limit: Na, Nb <= roma_match_keypoints_static_max_points() (16384 keypoints per image; 512 KiB of device memory per
 * GPU that loads the library, held until the process exits).  Larger sizes: ROMA_ERR_UNSUPPORTED with a message —
 * use roma_match_keypoints with a caller workspace.  Launches on the CURRENT device (the one the pointers live on).
 * Every call on a device shares the one array: calls that overlap on the GPU (two streams, two host threads) RACE;
 * serialise them on one stream.  Nothing in the array needs clearing between calls (every entry is written before
 * it is read within a call); it keeps the last call's contents. */
int roma_match_keypoints_static_max_points(void);
int roma_match_keypoints_static(const float* x_A, int Na, const float* x_B, int Nb,
                                const float* warp, const float* certainty, int H, int Wp,
                                float thresh, int64_t* inds_A, int64_t* inds_B, int capacity, int* count,
                                void* stream);

/* ---------------------------------------------------------------------------------------
 * (f-1) cls_to_flow_refine: softmax over res^2 bins, argmax, 5-tap weighted mean.
 *
 * Replaces (reference): romatch/utils/utils.py:301-323.
 * cls : (B, C=res*res, H, W) logits, dtype `dtype`, explicit strides; flow_out: fp32 (B, H, W, 2).
 */
int roma_cls_to_flow_refine(const void* cls, roma_strides4 cls_strides, int dtype,
                            float* flow_out, int B, int C, int H, int W, void* stream);

/* ---------------------------------------------------------------------------------------
 * (f-2) ConvRefiner body: depthwise 5x5 convolution (padding 2, stride 1) + bias + optional ReLU, channels-last.
 *
 * Replaces (reference): the conv1 -> BatchNorm2d -> ReLU triple of every ConvRefiner block as it runs at
 *   inference, romatch/models/matcher.py:77-103 (create_block) / :139-140; the eval-mode BatchNorm is an affine
 *   map that the caller folds into `weight` and `bias`.
 * x, out : (B, h, w, C) channels-last, dtype `dtype`, C even, out != x.
 * weight : fp32 (25, C), tap-major (tap = ky*5 + kx, cross-correlation like conv2d); bias : fp32 (C).
 */
int roma_dwconv5x5_bias_act(const void* x, const float* weight, const float* bias, void* out,
                            int B, int h, int w, int C, int relu, int dtype, void* stream);
/* One whole ConvRefiner block of the finest scale (hidden width 22 padded to 24): depthwise 5x5 + folded BatchNorm +
 * ReLU immediately followed by the block's 1x1 convolution 24 -> 24 (matcher.py:77-103: conv1, norm, relu, conv2),
 * the depthwise result staying in shared memory (two passes over the activation instead of four).
 * x, out: (B, h, w, 24) channels-last fp16 / bf16; dw_weight fp32 (25, 24) tap-major, dw_bias fp32 (24);
 * pw_weight fp32 (24, 24) [out][in], pw_bias fp32 (24).  Other widths: ROMA_ERR_UNSUPPORTED (use the two calls). */
int roma_dwconv5x5_pointwise(const void* x, const float* dw_weight, const float* dw_bias, const float* pw_weight,
                             const float* pw_bias, void* out, int B, int h, int w, int C, int relu, int dtype,
                             void* stream);
/* (f-2) ConvRefiner head: out[p, o] = bias[o] + sum_c weight[o, c] * x[p, c], o < OUT <= 3, fp32 accumulation.
 *
 * Replaces (reference): romatch/models/matcher.py:141 `self.out_conv(d.float())` (fp32 copy of the activation +
 *   1x1 cudnn convolution + bias add); the activation is read once in its storage dtype.
 * x      : (npix, C) channels-last activation (npix = B*h*w), dtype `dtype`, C a multiple of 8 (4 for fp32),
 *          16-byte aligned.
 * weight : fp32 (OUT, C); bias: fp32 (OUT).
 * out    : fp32 (npix, 4): (dx, dy, dcertainty, 0) per pixel (components >= OUT are 0), 16-byte aligned; pass it
 *          to roma_flow_update_resample as `delta` with strides (b: h*w*4, c: 1, y: w*4, x: 4).
 */
int roma_refiner_out_conv(const void* x, const float* weight, const float* bias, float* out,
                          long long npix, int C, int OUT, int dtype, void* stream);
/* Narrow 1x1 convolution, OUT <= 32, output in the storage dtype: out[p, o] = bias[o] + sum_c weight[o, c] * x[p, c].
 *
 * Replaces (reference): the finest-scale feature projection proj["1"] = Conv2d(64, 9, 1) + BatchNorm2d,
 *   romatch/models/model_zoo/roma_models.py:131-142 as applied at matcher.py:372-375 (eval-mode BatchNorm folded
 *   into weight / bias by the caller); the 24 -> 24 pointwise convolutions (conv2 of matcher.py:77-103) of the finest
 *   ConvRefiner, where a GEMM with N = K = 24 leaves cuBLAS at a third of HBM speed.
 * x : (npix, C) channels-last, dtype `dtype`, C a multiple of 8 (4 for fp32), x and out 16-byte aligned;
 * weight : fp32 (OUT, C); bias : fp32 (OUT); out : (npix, OUT) packed, dtype `dtype`.  fp32 accumulation.
 */
int roma_pointwise_narrow(const void* x, const float* weight, const float* bias, void* out,
                          long long npix, int C, int OUT, int dtype, void* stream);
/* Testing aid: 0 (default) = 16-bit inputs with C <= 64 use warp-level tensor-core MMAs (weights rounded to the
 * input dtype, as autocast does in the reference); 1 = always the fp32-weight FFMA2 kernel. */
int roma_pointwise_narrow_set_simt(int on);
/* Testing aid: 0 = automatic (TMA-tiled input when C * sizeof(dtype) is a multiple of 16 and x is 16-byte
 * aligned, per-thread cp.async staging otherwise), 1 = always the cp.async variant.  Results are identical. */
int roma_dwconv_set_path(int path);

/* ---------------------------------------------------------------------------------------
 * Encoder: LayerNorm over the last dimension, out = (x - mean) / sqrt(var + eps) * gamma + beta (biased variance,
 * fp32 statistics), one read and one write of the token matrix.
 *
 * Replaces (reference): nn.LayerNorm in romatch/models/transformer/layers/block.py:88-101 (norm1, norm2 of every
 *   DINOv2 / embedding-decoder block) and dinov2.py:236 (final norm).
 * x, out : (rows, D); gamma, beta : (D); all dtype `dtype`, D a multiple of 8 (4 for fp32) and <= 2048 (1024 for
 * fp32), 16-byte aligned.  out may alias x.
 */
int roma_layernorm(const void* x, const void* gamma, const void* beta, void* out,
                   long long rows, int D, float eps, int dtype, void* stream);
/* LayerNorm(x + shift) with shift fp32 (D), added in fp32 in front of the statistics.  Lets a pre-norm transformer keep
 * its residual stream without the bias vectors of its residual branches (each branch = one GEMM accumulating in place),
 * the accumulated bias re-entering at every norm.  Otherwise as roma_layernorm. */
int roma_layernorm_shift(const void* x, const float* shift, const void* gamma, const void* beta, void* out,
                         long long rows, int D, float eps, int dtype, void* stream);
/* LayerNorm of fp32 rows written directly in a 16-bit dtype (out_dtype ROMA_F16 / ROMA_BF16; gamma, beta fp32): the fp32
 * layer_norm + cast that autocast performs in front of the Linear consuming the normalised tokens
 * (romatch/models/transformer/layers/block.py:88-101 under transformer/__init__.py's autocast), same bits, one pass. */
int roma_layernorm_cast(const void* x, const void* gamma, const void* beta, void* out,
                        long long rows, int D, float eps, int out_dtype, void* stream);
/* sum = x + y (rounded to the storage dtype, as ATen's add) and out = LayerNorm(sum), one pass: the residual add of a
 * pre-norm block and the norm that follows it (block.py:88-101: x = x + ls(attn(norm1(x))); ... norm2(x); the second add
 * of a block pairs with the next block's norm1 / the final norm).  All tensors of one dtype; sum may alias x (in place);
 * out must be distinct.  Same bits as roma-less `x + y` followed by roma_layernorm. */
int roma_add_layernorm(const void* x, const void* y, const void* gamma, const void* beta, void* sum, void* out,
                       long long rows, int D, float eps, int dtype, void* stream);
/* out = x + y, x / out fp32, y 16-bit, n % 4 == 0: the residual add of the embedding decoder's fp32 token stream and an
 * autocast branch (block.py:88-101: x = x + ls(attn(norm(x)))), which ATen runs unvectorised. */
int roma_residual_add(const float* x, const void* y, float* out, long long n, int y_dtype, void* stream);

/* Exact (erf) GELU over a flat tensor of n elements (n % 8 == 0 for 16-bit, % 4 for fp32; 16-byte aligned; in place allowed).
 * Replaces (reference): nn.GELU() of every transformer MLP, romatch/models/transformer/layers/mlp.py:23, 31, 37.
 * x * Phi(x) with Phi(-|x|) = erfc(|x| / sqrt 2) / 2 from a relative-accuracy erfc fit (< 1.2e-7 fractional error), so the
 * 16-bit output is the correctly rounded exact GELU (ATen's 0.5 * x * (1 + erf) form loses relative accuracy in the negative
 * tail: the two differ by one unit in the last place in ~2-3 % of the elements, all with |y| < 0.05). */
int roma_gelu(const void* x, void* out, long long n, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * Encoder: 2x2 / stride-2 max pooling, channels-last (floor semantics: odd trailing row / column dropped).
 *
 * Replaces (reference): the nn.MaxPool2d(2, 2) layers of vgg19_bn.features walked by
 *   romatch/models/encoders.py:57-66 (VGG19.forward).
 * x : (B, h, w, C), out : (B, h/2, w/2, C), dtype `dtype`, C a multiple of 8 (4 for fp32), 16-byte aligned.
 */
int roma_maxpool2x2_nhwc(const void* x, void* out, int B, int h, int w, int C, int dtype, void* stream);

/* ---------------------------------------------------------------------------------------
 * KT  Tiny-RoMa fused correlation volume + approximate soft-argmax position embedding.
 *
 * Replaces (reference): romatch/models/tiny.py:241-254 (corr_volume) and :178-199 (pos_embed,
 *   inference branch) without materialising the (B, H1*W1, H0*W0) volume.
 * feat0 (B, H0*W0, C), feat1 (B, H1*W1, C) fp32 token-major; out : fp32 (B, 2, H0, W0).
 */
int roma_tiny_corr_posembed(const float* feat0, const float* feat1, float* out,
                            int B, int C, int H0, int W0, int H1, int W1, void* stream);
/* The same op on tensor cores (default in the host wrapper when C % 4 == 0): scores = feat0 feat1^T / sqrt(C) as a
 * tcgen05 / TMEM GEMM over fp32 features split on the fly into three bf16 planes (fp32-grade products), the arg-max
 * and online-softmax reduction fused on the TMEM drain (one record per 64-column tile half and query), one merge
 * kernel over the column tiles.  workspace: roma_tiny_corr_posembed_workspace_bytes(...) bytes, 16-byte aligned. */
size_t roma_tiny_corr_posembed_workspace_bytes(int B, int H0, int W0, int H1, int W1);
int roma_tiny_corr_posembed_tc(const float* feat0, const float* feat1, float* out,
                               int B, int C, int H0, int W0, int H1, int W1,
                               void* workspace, size_t workspace_bytes, void* stream);

/* ---------------------------------------------------------------------------------------
 * (f-3)  Input path: uint8 RGB batch -> ImageNet-normalised fp32 images at a pass resolution.
 *
 * Replaces (reference): romatch/models/matcher.py:606-637, 667-676 with romatch/utils/utils.py:165-261
 *   (resize, ToTensor, Normalize) in tensor form: (u8 / 255 - mean) / std, then — where (h, w) != (H, W) —
 *   F.interpolate(mode="bicubic", antialias=True, align_corners=False) with ATen's weights and summation order.
 * images : DEVICE uint8 (B, 3, H, W) planar, contiguous.   out : DEVICE fp32 (B, 3, h, w) planar.
 * mean3, std3 : HOST pointers to 3 floats each (passed to the kernel by value).
 * Downscale factors above 4 per dimension: ROMA_ERR_UNSUPPORTED.
 */
int roma_prepare_images(const unsigned char* images, float* out, int B, int H, int W, int h, int w,
                        const float* mean3, const float* std3, void* stream);

/* (f-3)  Input path with the reference's OWN resize: PIL `Image.resize((w, h), Image.BICUBIC)` of an 8-bit RGB image,
 * then ToTensor and Normalize — bit-exact integer arithmetic.
 *
 * Replaces (reference): romatch/utils/utils.py:223-232 (TupleResize -> PIL resize), :234-239 (ToTensorScaled, / 255),
 *   :254-261 (Normalize) as applied by romatch/models/matcher.py:628-634 and :667-676 to file / PIL inputs.  The resize
 *   arithmetic is Pillow's (third-party, libImaging/Resample.c, 8 bits per channel): double-precision bicubic
 *   coefficients quantised to 22 fractional bits, horizontal then vertical pass, each rounded and clipped to 8 bits;
 *   a dimension that does not change is passed through.
 * images : DEVICE uint8, B images of 3 channels, addressed images[b*stride_b + c*stride_c + y*stride_y + x*stride_x]
 *          (strides in bytes: (H, W, 3) interleaved as decoded by PIL = {H*W*3, 1, W*3, 3}; planar = {3*H*W, H*W, W, 1}).
 * out    : DEVICE fp32 (B, 3, h, w) planar = (resized / 255 - mean) / std, or NULL.
 * out_u8 : DEVICE uint8 (B, 3, h, w) planar = the resized 8-bit image, or NULL (at least one output is required).
 * mean3, std3 : HOST pointers to 3 floats each (required with `out`).
 * Downscale factors above 4 per dimension: ROMA_ERR_UNSUPPORTED.
 */
int roma_prepare_images_pil(const unsigned char* images, float* out, unsigned char* out_u8, int B, int H, int W,
                            int h, int w, long long stride_b, long long stride_c, long long stride_y,
                            long long stride_x, const float* mean3, const float* std3, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* ROMA_B200_H */
