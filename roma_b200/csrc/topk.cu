// roma_topk_desc: the k largest of a 1-D score vector, returned as int64 indices ordered by (value descending, index
// ascending) — the selection step of torch.multinomial(replacement=False) = topk(p / Exp(1)) that both draws of
// RegressionMatcher.sample make (reference romatch/models/matcher.py:482-494).  ATen runs it as a 4-pass multi-block
// radix select over the 32 value bits (4 launches per pass), a gather, a radix sort of the k survivors with their
// indices and two more gathers (~30 launches, ~0.35 ms for k = 40000 of 1.49 M); its tie order is "gathered in index
// order, then a stable descending sort", i.e. exactly the total order used here.
//
// Here every element becomes one 64-bit word  comp = ordered_key(value) << nb | (~index & (2^nb - 1))  (nb = bits of
// n - 1): all words are distinct, so "the k largest words" is unambiguous and carries the tie rule.  Radix select over
// comp with 11-bit digits from the top (histogram in shared memory, the last block to finish picks the digit; a pass
// exits at once when an earlier one already isolated exactly k words), one filter pass that appends the k words >= the
// threshold, one radix sort of k words (CUB DeviceRadixSort over the low KB + nb bits, double-buffered in the caller's
// workspace), one unpack.  Inputs of at most 131072 elements (the second draw: 10000 of 40000) skip the select and sort all words.  HBM/L2-bound streams over n values: 1.49 M floats = 6 MB per pass.
#define THRUST_CUB_WRAPPED_NAMESPACE roma_b200_cub
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>

namespace roma {
namespace topk {

constexpr int kDigitBits = 11;
constexpr int kBins = 1 << kDigitBits;
constexpr int kMaxPasses = 6;      // ceil(64 / 11)
constexpr int kThreads = 256;

struct State {                     // zeroed by one memset per call
    unsigned long long prefix;     // digits chosen so far (right-aligned)
    unsigned long long threshold;  // valid once done != 0: select comp >= threshold
    unsigned int remaining;        // words still to take inside the prefix bucket
    unsigned int done;
    unsigned int ticket;
    unsigned int appended;
    unsigned int hist[kMaxPasses][kBins];
};

template <typename T> struct Key;
template <> struct Key<float> {
    static constexpr int kBits = 32;
    __device__ static __forceinline__ unsigned long long of(float v) {
        const unsigned int u = __float_as_uint(v);
        return (u & 0x80000000u) ? (unsigned int)~u : (u | 0x80000000u);
    }
};
template <> struct Key<__half> {
    static constexpr int kBits = 16;
    __device__ static __forceinline__ unsigned long long of(__half v) {
        const unsigned int u = __half_as_ushort(v);
        return ((u & 0x8000u) ? ~u : (u | 0x8000u)) & 0xFFFFu;
    }
};
template <> struct Key<__nv_bfloat16> {
    static constexpr int kBits = 16;
    __device__ static __forceinline__ unsigned long long of(__nv_bfloat16 v) {
        const unsigned int u = __bfloat16_as_ushort(v);
        return ((u & 0x8000u) ? ~u : (u | 0x8000u)) & 0xFFFFu;
    }
};

template <typename T>
__device__ __forceinline__ unsigned long long compose(T v, unsigned int idx, int nb, unsigned int idx_mask) {
    return (Key<T>::of(v) << nb) | (unsigned long long)(~idx & idx_mask);
}

// One pass of the radix select: histogram of the digit at [shift, shift + width) over the words whose higher bits equal
// the prefix chosen so far; the last block picks the digit d with count(> d) < remaining <= count(>= d).
template <typename T>
__global__ void __launch_bounds__(kThreads)
select_pass_kernel(const T* __restrict__ vals, unsigned int n, unsigned int k, int nb, unsigned int idx_mask, int pass, int shift,
                   int width, State* __restrict__ s) {
    if (s->done) return;
    __shared__ unsigned int h[kBins];
    __shared__ unsigned int is_last;
    const unsigned long long prefix = s->prefix;
    const int hi = shift + width;                                   // bits above the digit: must equal the prefix
    for (int i = threadIdx.x; i < kBins; i += kThreads) h[i] = 0;
    __syncthreads();
    const unsigned int dmask = (1u << width) - 1u;
    for (unsigned int i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        const unsigned long long c = compose<T>(vals[i], i, nb, idx_mask);
        const bool in = pass == 0 || (c >> hi) == prefix;
        // warp-aggregated: scores cluster in a few exponent bins, 32 lanes on one shared address would serialise
        const unsigned int d = in ? (unsigned int)(c >> shift) & dmask : 0xFFFFFFFFu;
        const unsigned int act = __activemask();
        const unsigned int peers = __match_any_sync(act, d);
        if (in && (int)(__ffs(peers) - 1) == (int)(threadIdx.x & 31)) atomicAdd(&h[d], __popc(peers));
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
    // ---- last block: suffix counts from the top digit down (8 bins per thread, block scan in shared memory)
    constexpr int PER = kBins / kThreads;
    unsigned int loc[PER], sum = 0;
    const int top = kBins - 1 - threadIdx.x * PER;                  // this thread owns digits top, top-1, ..., top-PER+1
#pragma unroll
    for (int j = 0; j < PER; ++j) { loc[j] = __ldcg(&g[top - j]); sum += loc[j]; }
    __shared__ unsigned int scan[kThreads];
    scan[threadIdx.x] = sum;
    __syncthreads();
    for (int off = 1; off < kThreads; off <<= 1) {
        const unsigned int v = threadIdx.x >= off ? scan[threadIdx.x - off] : 0u;
        __syncthreads();
        scan[threadIdx.x] += v;
        __syncthreads();
    }
    const unsigned int need = pass == 0 ? k : s->remaining;
    unsigned int above = scan[threadIdx.x] - sum;                   // words with a digit above this thread's range
#pragma unroll
    for (int j = 0; j < PER; ++j) {
        if (above < need && need <= above + loc[j]) {               // exactly one (thread, j) satisfies this
            const unsigned long long p = (prefix << width) | (unsigned long long)(top - j);
            s->prefix = p;
            s->remaining = need - above;
            if (need - above == loc[j] || shift == 0) { s->threshold = p << shift; s->done = 1u; }
        }
        above += loc[j];
    }
    if (threadIdx.x == 0) s->ticket = 0u;
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
filter_kernel(const T* __restrict__ vals, unsigned int n, int nb, unsigned int idx_mask, State* __restrict__ s,
              unsigned long long* __restrict__ cand, unsigned int k) {
    const unsigned long long thr = s->threshold;
    for (unsigned int i = blockIdx.x * kThreads + threadIdx.x; i < n; i += gridDim.x * kThreads) {
        const unsigned long long c = compose<T>(vals[i], i, nb, idx_mask);
        const bool take = c >= thr;
        const unsigned int act = __activemask();
        const unsigned int m = __ballot_sync(act, take);
        if (m) {
            const int lane = threadIdx.x & 31;
            unsigned int base = 0;
            if (lane == __ffs(m) - 1) base = atomicAdd(&s->appended, __popc(m));
            base = __shfl_sync(act, base, __ffs(m) - 1);
            const unsigned int pos = base + __popc(m & ((1u << lane) - 1u));
            if (take && pos < k) cand[pos] = c;
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(kThreads)
unpack_kernel(const unsigned long long* __restrict__ sorted, unsigned int k, unsigned int idx_mask, const T* __restrict__ vals,
              long long* __restrict__ out_idx, T* __restrict__ out_vals) {
    const unsigned int i = blockIdx.x * kThreads + threadIdx.x;
    if (i >= k) return;
    const unsigned int idx = ~(unsigned int)sorted[i] & idx_mask;
    out_idx[i] = (long long)idx;
    if (out_vals) out_vals[i] = vals[idx];
}

// Short inputs (n <= kDirectMax): no select — every word goes through the sort, the first k come out.
template <typename T>
__global__ void __launch_bounds__(kThreads)
compose_kernel(const T* __restrict__ vals, unsigned int n, int nb, unsigned int idx_mask, unsigned long long* __restrict__ words) {
    const unsigned int i = blockIdx.x * kThreads + threadIdx.x;
    if (i < n) words[i] = compose<T>(vals[i], i, nb, idx_mask);
}

constexpr long long kDirectMax = 131072;
static inline long long sorted_items(long long n, long long k) { return n <= kDirectMax ? n : k; }

static inline size_t align256(size_t v) { return (v + 255) & ~(size_t)255; }

static size_t sort_temp_bytes(long long k, int bits) {
    size_t bytes = 0;
    ::roma_b200_cub::cub::DoubleBuffer<unsigned long long> keys(nullptr, nullptr);
    if (::roma_b200_cub::cub::DeviceRadixSort::SortKeysDescending(nullptr, bytes, keys, (unsigned int)k, 0, bits, 0) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return bytes;
}

static int index_bits(long long n) { int nb = 1; while (nb < 32 && (1ll << nb) < n) ++nb; return nb; }

template <typename T>
static int run(const T* vals, long long n, long long k, long long* out_idx, T* out_vals, void* ws, size_t ws_bytes, cudaStream_t st) {
    const int nb = index_bits(n), total = Key<T>::kBits + nb;
    const unsigned int idx_mask = nb >= 32 ? 0xFFFFFFFFu : ((1u << nb) - 1u);
    char* base = reinterpret_cast<char*>(ws);
    State* s = reinterpret_cast<State*>(base);
    const long long m = sorted_items(n, k);
    const size_t off_a = align256(sizeof(State)), off_b = off_a + align256((size_t)m * 8), off_t = off_b + align256((size_t)m * 8);
    const size_t temp = sort_temp_bytes(m, total);
    ROMA_REQUIRE(temp > 0, ROMA_ERR_UNSUPPORTED, "roma_topk_desc: the radix-sort size query failed");
    ROMA_REQUIRE(ws_bytes >= off_t + temp, ROMA_ERR_WORKSPACE, "roma_topk_desc: workspace too small (%zu < %zu bytes, see roma_topk_workspace_bytes)",
                 ws_bytes, off_t + temp);
    unsigned long long* cand_a = reinterpret_cast<unsigned long long*>(base + off_a);
    unsigned long long* cand_b = reinterpret_cast<unsigned long long*>(base + off_b);
    const unsigned int kk = (unsigned int)k;
    int rc;
    if (m == n) {
        compose_kernel<T><<<cdiv(n, kThreads), kThreads, 0, st>>>(vals, (unsigned int)n, nb, idx_mask, cand_a);
        if ((rc = check_launch("roma_topk_desc(compose)", 1))) return rc;
    } else {
        cudaMemsetAsync(s, 0, sizeof(State), st);
        const int blocks = (int)std::min<long long>((n + 8 * kThreads - 1) / (8 * kThreads), (long long)sm_count() * 8);
        int launches = 0;
        for (int pass = 0, hi = total; hi > 0; ++pass) {
            const int width = hi >= kDigitBits ? kDigitBits : hi, shift = hi - width;
            select_pass_kernel<T><<<blocks, kThreads, 0, st>>>(vals, (unsigned int)n, kk, nb, idx_mask, pass, shift, width, s);
            hi = shift;
            ++launches;
        }
        filter_kernel<T><<<blocks, kThreads, 0, st>>>(vals, (unsigned int)n, nb, idx_mask, s, cand_a, kk);
        if ((rc = check_launch("roma_topk_desc(select, filter)", launches + 1))) return rc;
    }
    ::roma_b200_cub::cub::DoubleBuffer<unsigned long long> keys(cand_a, cand_b);
    size_t tb = ws_bytes - off_t;
    const cudaError_t e = ::roma_b200_cub::cub::DeviceRadixSort::SortKeysDescending(base + off_t, tb, keys, (unsigned int)m, 0, total, st);
    ROMA_REQUIRE(e == cudaSuccess, ROMA_ERR_UNSUPPORTED, "roma_topk_desc: radix sort: %s", cudaGetErrorString(e));
    unpack_kernel<T><<<cdiv(k, kThreads), kThreads, 0, st>>>(keys.Current(), kk, idx_mask, vals, out_idx, out_vals);
    return check_launch("roma_topk_desc(unpack)", 1);
}

}  // namespace topk
}  // namespace roma

extern "C" size_t roma_topk_workspace_bytes(long long n, long long k) {
    using namespace roma::topk;
    if (n <= 0 || k <= 0 || k > n || n > 0x7FFFFFFFll) return 0;
    const long long m = sorted_items(n, k);
    const size_t temp = sort_temp_bytes(m, 64);
    if (!temp) return 0;
    return align256(sizeof(State)) + 2 * align256((size_t)m * 8) + align256(temp) + 256;
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
