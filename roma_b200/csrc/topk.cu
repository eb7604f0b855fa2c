// roma_topk_desc: the k largest of a 1-D score vector, returned as int64 indices ordered by (value descending, index
// ascending) — the selection step of torch.multinomial(replacement=False) = topk(p / Exp(1)) that both draws of
// RegressionMatcher.sample make (reference romatch/models/matcher.py:482-494).  ATen runs it as a 4-pass multi-block
// radix select over the 32 value bits (4 launches per pass), a gather, a radix sort of the k survivors with their
// indices and two more gathers (~30 launches, ~0.28 ms for k = 40000 of 1.49 M); its tie order is "gathered in index
// order, then a stable descending sort", i.e. exactly the order produced here.
//
// Pipeline (n > 131072): radix select of the k-th largest ordered key with 11-bit digits from the top (3 passes for fp32,
// 2 for 16-bit scores; histogram in shared memory, the last block to finish picks the digit) -> K and the number of
// ties at K to keep; an ORDERED compaction in two launches (per-block counts of "> K" and "== K" over contiguous
// chunks, scanned by the last block; then every block writes its survivors at their index-ordered position: the first
// `need` ties in index order are the ones kept) -> k (key, index) pairs in ascending index order; one STABLE descending
// radix sort on the key bits only (CUB DeviceRadixSort::SortPairsDescending, 32 or 16 bits, indices as int64 values,
// written straight into the caller's output).  Short inputs (the second draw: 10000 of 40000 half scores) skip the
// select and sort every element.  Streams over n values: 1.49 M floats = 6 MB per pass, L2-resident after the first.
#define THRUST_CUB_WRAPPED_NAMESPACE roma_b200_cub
#include "common.cuh"
#include <algorithm>
#include <cub/device/device_radix_sort.cuh>

namespace roma {
namespace topk {

constexpr int kDigitBits = 11;
constexpr int kBins = 1 << kDigitBits;
constexpr int kMaxPasses = 3;            // 32 key bits / 11
constexpr int kThreads = 256;
constexpr int kWarps = kThreads / 32;
constexpr int kBatch = 8;               // elements per thread whose loads are issued together
constexpr int kMaxBlocks = 2048;         // chunks of the ordered compaction (= blocks of its two kernels)
constexpr long long kDirectMax = 131072; // at most this many elements: sort them all

struct State {                           // zeroed by one memset per call
    unsigned int prefix;                 // digits chosen so far (right-aligned); after the last pass: K
    unsigned int remaining;              // results still to take inside the prefix bucket; after the last pass: ties to keep
    unsigned int ticket;
    unsigned int pad;
    unsigned int hist[kMaxPasses][kBins];
    unsigned int gt[kMaxBlocks];         // per chunk: elements with key > K, then (after the scan) those in earlier chunks
    unsigned int eq[kMaxBlocks];         // same for key == K
};

template <typename T> struct Key;
template <> struct Key<float> {
    static constexpr int kBits = 32;
    __device__ static __forceinline__ unsigned int of(float v) {
        const unsigned int u = __float_as_uint(v);
        return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
    }
};
template <> struct Key<__half> {
    static constexpr int kBits = 16;
    __device__ static __forceinline__ unsigned int of(__half v) {
        const unsigned int u = __half_as_ushort(v);
        return ((u & 0x8000u) ? ~u : (u | 0x8000u)) & 0xFFFFu;
    }
};
template <> struct Key<__nv_bfloat16> {
    static constexpr int kBits = 16;
    __device__ static __forceinline__ unsigned int of(__nv_bfloat16 v) {
        const unsigned int u = __bfloat16_as_ushort(v);
        return ((u & 0x8000u) ? ~u : (u | 0x8000u)) & 0xFFFFu;
    }
};

// One pass of the radix select: histogram of the digit at [shift, shift + width) over the keys whose higher bits equal
// the prefix chosen so far; the last block picks the digit d with count(> d) < remaining <= count(>= d).
template <typename T>
__global__ void __launch_bounds__(kThreads)
select_pass_kernel(const T* __restrict__ vals, unsigned int n, unsigned int k, int pass, int shift, int width, State* __restrict__ s) {
    __shared__ unsigned int h[kBins];
    __shared__ unsigned int scan[kThreads];
    __shared__ unsigned int is_last;
    const unsigned int prefix = s->prefix, need = pass == 0 ? k : s->remaining;   // read before the last block rewrites them
    const int hi = shift + width;                                   // bits above the digit: must equal the prefix
    for (int i = threadIdx.x; i < kBins; i += kThreads) h[i] = 0;
    __syncthreads();
    const unsigned int dmask = (1u << width) - 1u;
    const int lane = threadIdx.x & 31;
    for (unsigned int base = blockIdx.x * (kThreads * kBatch); base < n; base += gridDim.x * (kThreads * kBatch)) {   // uniform trips
        unsigned int key[kBatch];
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {                           // all loads of the batch in flight before the first use
            const unsigned int i = base + j * kThreads + threadIdx.x;
            key[j] = i < n ? Key<T>::of(vals[i]) : 0u;
        }
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
            const unsigned int i = base + j * kThreads + threadIdx.x;
            unsigned int d = 0xFFFFFFFFu;
            if (i < n && (pass == 0 || (key[j] >> hi) == prefix)) d = (key[j] >> shift) & dmask;
            // scores cluster in a few exponent bins (and zero-heavy inputs in one): a warp that agrees adds once
            const unsigned int d0 = __shfl_sync(0xFFFFFFFFu, d, 0);
            if (__all_sync(0xFFFFFFFFu, d == d0)) {
                if (lane == 0 && d0 != 0xFFFFFFFFu) atomicAdd(&h[d0], 32u);
            } else if (d != 0xFFFFFFFFu) {
                atomicAdd(&h[d], 1u);
            }
        }
    }
    __syncthreads();
    unsigned int* g = s->hist[pass];
    for (int i = threadIdx.x; i < kBins; i += kThreads)
        if (h[i]) atomicAdd(&g[i], h[i]);
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) is_last = (atomicAdd(&s->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    // ---- last block: counts from the top digit down (8 bins per thread, block scan in shared memory)
    constexpr int PER = kBins / kThreads;
    unsigned int loc[PER], sum = 0;
    const int top = kBins - 1 - threadIdx.x * PER;                  // this thread owns digits top, top-1, ..., top-PER+1
#pragma unroll
    for (int j = 0; j < PER; ++j) { loc[j] = __ldcg(&g[top - j]); sum += loc[j]; }
    scan[threadIdx.x] = sum;
    __syncthreads();
    for (int off = 1; off < kThreads; off <<= 1) {
        const unsigned int v = threadIdx.x >= off ? scan[threadIdx.x - off] : 0u;
        __syncthreads();
        scan[threadIdx.x] += v;
        __syncthreads();
    }
    unsigned int above = scan[threadIdx.x] - sum;                   // keys with a digit above this thread's range
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        if (above < need && need <= above + loc[j]) {               // exactly one (thread, j) satisfies this
            s->prefix = (prefix << width) | (unsigned int)(top - j);
            s->remaining = need - above;
        }
        above += loc[j];
    }
    if (threadIdx.x == 0) s->ticket = 0u;
}

// Ordered compaction, step 1: per contiguous chunk, how many keys are > K and how many == K; the last block turns the
// two count arrays into exclusive prefix sums over the chunks.
template <typename T>
__global__ void __launch_bounds__(kThreads)
count_kernel(const T* __restrict__ vals, unsigned int n, unsigned int chunk, State* __restrict__ s) {
    __shared__ unsigned int cg, ce, is_last;
    __shared__ unsigned int sg[kThreads], se[kThreads];
    const unsigned int K = s->prefix;
    if (threadIdx.x == 0) { cg = 0; ce = 0; }
    __syncthreads();
    const unsigned int start = blockIdx.x * chunk;
    const unsigned int stop = start + chunk < n ? start + chunk : n;
    unsigned int ng = 0, ne = 0;
    for (unsigned int base = start; base < stop; base += kThreads * kBatch) {
        unsigned int key[kBatch];
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
            const unsigned int i = base + j * kThreads + threadIdx.x;
            key[j] = i < stop ? Key<T>::of(vals[i]) : 0u;
        }
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
            const unsigned int i = base + j * kThreads + threadIdx.x;
            ng += i < stop && key[j] > K;
            ne += i < stop && key[j] == K;
        }
    }
    for (int off = 16; off; off >>= 1) {
        ng += __shfl_xor_sync(0xFFFFFFFFu, ng, off);
        ne += __shfl_xor_sync(0xFFFFFFFFu, ne, off);
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&cg, ng); atomicAdd(&ce, ne); }
    __syncthreads();
    if (threadIdx.x == 0) {
        s->gt[blockIdx.x] = cg;
        s->eq[blockIdx.x] = ce;
        __threadfence();
        is_last = (atomicAdd(&s->ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
    }
    __syncthreads();
    if (!is_last) return;
    __threadfence();
    constexpr int PER = kMaxBlocks / kThreads;
    unsigned int lg[PER], le[PER], tg = 0, te = 0;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const unsigned int b = threadIdx.x * PER + j;
        lg[j] = b < gridDim.x ? __ldcg(&s->gt[b]) : 0u;
        le[j] = b < gridDim.x ? __ldcg(&s->eq[b]) : 0u;
        tg += lg[j];
        te += le[j];
    }
    sg[threadIdx.x] = tg;
    se[threadIdx.x] = te;
    __syncthreads();
    for (int off = 1; off < kThreads; off <<= 1) {
        const unsigned int vg = threadIdx.x >= off ? sg[threadIdx.x - off] : 0u;
        const unsigned int ve = threadIdx.x >= off ? se[threadIdx.x - off] : 0u;
        __syncthreads();
        sg[threadIdx.x] += vg;
        se[threadIdx.x] += ve;
        __syncthreads();
    }
    unsigned int bg = sg[threadIdx.x] - tg, be = se[threadIdx.x] - te;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        const unsigned int b = threadIdx.x * PER + j;
        if (b < gridDim.x) { s->gt[b] = bg; s->eq[b] = be; }
        bg += lg[j];
        be += le[j];
    }
    if (threadIdx.x == 0) s->ticket = 0u;
}

// Step 2: every kept element goes to position (#"> K" before it) + min(#"== K" before it, need), "before" in index order.
template <typename T>
__global__ void __launch_bounds__(kThreads)
write_kernel(const T* __restrict__ vals, unsigned int n, unsigned int chunk, const State* __restrict__ s,
             unsigned int* __restrict__ keys, long long* __restrict__ idx, unsigned int k) {
    __shared__ unsigned int wg[2][kWarps], we[2][kWarps];
    const unsigned int K = s->prefix, need = s->remaining;
    unsigned int run_g = s->gt[blockIdx.x], run_e = s->eq[blockIdx.x];
    const unsigned int start = blockIdx.x * chunk;
    const unsigned int stop = start + chunk < n ? start + chunk : n;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const unsigned int lt = (1u << lane) - 1u;
    int buf = 0;
    for (unsigned int base = start; base < stop; base += kThreads * kBatch) {   // uniform trip count per block
        unsigned int key[kBatch];
#pragma unroll
        for (int j = 0; j < kBatch; ++j) {
            const unsigned int i = base + j * kThreads + threadIdx.x;
            key[j] = i < stop ? Key<T>::of(vals[i]) : 0u;
        }
#pragma unroll
        for (int j = 0; j < kBatch; ++j, buf ^= 1) {
            const unsigned int i = base + j * kThreads + threadIdx.x;
            const bool g = i < stop && key[j] > K, e = i < stop && key[j] == K;
            const unsigned int mg = __ballot_sync(0xFFFFFFFFu, g), me = __ballot_sync(0xFFFFFFFFu, e);
            if (lane == 0) { wg[buf][warp] = __popc(mg); we[buf][warp] = __popc(me); }
            __syncthreads();                                        // double-buffered warp totals: one barrier per round
            unsigned int pg = run_g, pe = run_e, tg = 0, te = 0;
#pragma unroll
            for (int w = 0; w < kWarps; ++w) {
                const unsigned int a = wg[buf][w], b = we[buf][w];
                if (w < warp) { pg += a; pe += b; }
                tg += a;
                te += b;
            }
            pg += __popc(mg & lt);
            pe += __popc(me & lt);
            if (g || (e && pe < need)) {
                const unsigned int pos = pg + (pe < need ? pe : need);
                if (pos < k) { keys[pos] = key[j]; idx[pos] = (long long)i; }
            }
            run_g += tg;
            run_e += te;
        }
    }
}

// Short inputs: every element is a candidate, already in index order.
template <typename T>
__global__ void __launch_bounds__(kThreads)
all_keys_kernel(const T* __restrict__ vals, unsigned int n, unsigned int* __restrict__ keys, long long* __restrict__ idx) {
    const unsigned int i = blockIdx.x * kThreads + threadIdx.x;
    if (i < n) { keys[i] = Key<T>::of(vals[i]); idx[i] = (long long)i; }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
gather_values_kernel(const T* __restrict__ vals, const long long* __restrict__ idx, unsigned int k, T* __restrict__ out) {
    const unsigned int i = blockIdx.x * kThreads + threadIdx.x;
    if (i < k) out[i] = vals[idx[i]];
}

static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }
static inline long long sorted_items(long long n, long long k) { return n <= kDirectMax ? n : k; }

static size_t sort_temp_bytes(long long m, int bits) {
    size_t bytes = 0;
    const cudaError_t e = ::roma_b200_cub::cub::DeviceRadixSort::SortPairsDescending(
        nullptr, bytes, (const unsigned int*)nullptr, (unsigned int*)nullptr, (const long long*)nullptr, (long long*)nullptr,
        (unsigned int)m, 0, bits, 0);
    if (e != cudaSuccess) { cudaGetLastError(); return 0; }
    return bytes;
}

struct Layout { size_t keys_in, keys_out, idx_in, idx_out, temp, total; };
static Layout layout(long long n, long long k, size_t temp_bytes) {
    const long long m = sorted_items(n, k);
    Layout L;
    L.keys_in = align256(sizeof(State));
    L.keys_out = L.keys_in + align256((size_t)m * 4);
    L.idx_in = L.keys_out + align256((size_t)m * 4);
    L.idx_out = L.idx_in + align256((size_t)m * 8);
    L.temp = L.idx_out + (m == k ? 0 : align256((size_t)m * 8));   // m == k: the sorted indices go straight to the caller's buffer
    L.total = L.temp + align256(temp_bytes);
    return L;
}

template <typename T>
static int run(const T* vals, long long n, long long k, long long* out_idx, T* out_vals, void* ws, size_t ws_bytes, cudaStream_t st) {
    const long long m = sorted_items(n, k);
    const size_t temp = sort_temp_bytes(m, Key<T>::kBits);
    ROMA_REQUIRE(temp > 0, ROMA_ERR_UNSUPPORTED, "roma_topk_desc: the radix-sort size query failed");
    const Layout L = layout(n, k, temp);
    ROMA_REQUIRE(ws_bytes >= L.total, ROMA_ERR_WORKSPACE, "roma_topk_desc: workspace too small (%zu < %zu bytes, see roma_topk_workspace_bytes)",
                 ws_bytes, L.total);
    char* base = reinterpret_cast<char*>(ws);
    State* s = reinterpret_cast<State*>(base);
    unsigned int* keys_in = reinterpret_cast<unsigned int*>(base + L.keys_in);
    unsigned int* keys_out = reinterpret_cast<unsigned int*>(base + L.keys_out);
    long long* idx_in = reinterpret_cast<long long*>(base + L.idx_in);
    long long* idx_out = m == k ? out_idx : reinterpret_cast<long long*>(base + L.idx_out);
    const unsigned int nn = (unsigned int)n, kk = (unsigned int)k;
    int rc;
    if (m == n) {
        all_keys_kernel<T><<<cdiv(n, kThreads), kThreads, 0, st>>>(vals, nn, keys_in, idx_in);
        if ((rc = check_launch("roma_topk_desc(keys)", 1))) return rc;
    } else {
        cudaMemsetAsync(s, 0, sizeof(State), st);
        const int blocks = (int)std::min<long long>((n + 8 * kThreads - 1) / (8 * kThreads), (long long)sm_count() * 8);
        int launches = 0;
        for (int pass = 0, hi = Key<T>::kBits; hi > 0; ++pass, ++launches) {
            const int width = hi >= kDigitBits ? kDigitBits : hi, shift = hi - width;
            select_pass_kernel<T><<<blocks, kThreads, 0, st>>>(vals, nn, kk, pass, shift, width, s);
            hi = shift;
        }
        const int cblocks = std::min<int>(blocks, kMaxBlocks);
        const unsigned int chunk = (unsigned int)(((n + cblocks - 1) / cblocks + kThreads - 1) / kThreads * kThreads);
        const int used = cdiv(n, chunk);
        count_kernel<T><<<used, kThreads, 0, st>>>(vals, nn, chunk, s);
        write_kernel<T><<<used, kThreads, 0, st>>>(vals, nn, chunk, s, keys_in, idx_in, kk);
        if ((rc = check_launch("roma_topk_desc(select, compact)", launches + 2))) return rc;
    }
    size_t tb = ws_bytes - L.temp;
    const cudaError_t e = ::roma_b200_cub::cub::DeviceRadixSort::SortPairsDescending(
        base + L.temp, tb, (const unsigned int*)keys_in, keys_out, (const long long*)idx_in, idx_out, (unsigned int)m, 0, Key<T>::kBits, st);
    ROMA_REQUIRE(e == cudaSuccess, ROMA_ERR_UNSUPPORTED, "roma_topk_desc: radix sort: %s", cudaGetErrorString(e));
    if (m != k) cudaMemcpyAsync(out_idx, idx_out, (size_t)k * 8, cudaMemcpyDeviceToDevice, st);
    if (out_vals) {
        gather_values_kernel<T><<<cdiv(k, kThreads), kThreads, 0, st>>>(vals, out_idx, kk, out_vals);
        return check_launch("roma_topk_desc(values)", 1);
    }
    return ROMA_OK;
}

}  // namespace topk
}  // namespace roma

extern "C" size_t roma_topk_workspace_bytes(long long n, long long k) {
    using namespace roma::topk;
    if (n <= 0 || k <= 0 || k > n || n > 0x7FFFFFFFll) return 0;
    const size_t temp = sort_temp_bytes(sorted_items(n, k), 32);
    if (!temp) return 0;
    return layout(n, k, temp).total + 256;
}

extern "C" int roma_topk_desc(const void* values, int dtype, long long n, long long k, long long* out_indices, void* out_values,
                              void* workspace, size_t workspace_bytes, void* stream) {
    using namespace roma;
    ROMA_REQUIRE(values && out_indices && workspace, ROMA_ERR_BAD_ARG, "roma_topk_desc: null pointer");
    ROMA_REQUIRE(n > 0 && n <= 0x7FFFFFFFll && k > 0 && k <= n, ROMA_ERR_BAD_ARG, "roma_topk_desc: need 0 < k <= n < 2^31 (n=%lld, k=%lld)", n, k);
    ROMA_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255u) == 0, ROMA_ERR_ALIGNMENT, "roma_topk_desc: workspace must be 256-byte aligned");
    cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
    switch (dtype) {
        case ROMA_F32:  return topk::run<float>(static_cast<const float*>(values), n, k, out_indices, static_cast<float*>(out_values), workspace, workspace_bytes, st);
        case ROMA_F16:  return topk::run<__half>(static_cast<const __half*>(values), n, k, out_indices, static_cast<__half*>(out_values), workspace, workspace_bytes, st);
        case ROMA_BF16: return topk::run<__nv_bfloat16>(static_cast<const __nv_bfloat16*>(values), n, k, out_indices, static_cast<__nv_bfloat16*>(out_values), workspace, workspace_bytes, st);
        default: return fail(ROMA_ERR_BAD_ARG, "roma_topk_desc: dtype %d", dtype);
    }
}
