// TMA (cp.async.bulk.tensor) plumbing shared by the kernels: host-side tensor-map encoding without linking
// libcuda (the driver entry point is fetched through the runtime), and the device-side PTX wrappers.
#pragma once
#include <cuda.h>
#include "common.cuh"

namespace roma {
namespace tma {

using EncodeTiledFn = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                   const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                   CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline EncodeTiledFn encode_fn() {
    static EncodeTiledFn fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q = cudaDriverEntryPointSymbolNotFound;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            p = nullptr;
        return reinterpret_cast<EncodeTiledFn>(p);
    }();
    return fn;
}

inline CUtensorMapDataType data_type(int roma_dtype) {
    return roma_dtype == ROMA_F32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
         : roma_dtype == ROMA_F16 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT16 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
}

// Tiled map of a dense row-major tensor of `rank` dims; dims[0] is the innermost (contiguous) one.
// strides_bytes[i] is the byte stride of dims[i + 1] (rank - 1 entries, multiples of 16).  Out-of-bounds
// elements of a box (including negative coordinates) read as zero.  Returns 0 or a ROMA_ERR_* code.
inline int encode_tiled(CUtensorMap* map, int roma_dtype, int rank, const void* base, const uint64_t* dims,
                        const uint64_t* strides_bytes, const uint32_t* box, CUtensorMapSwizzle swizzle) {
    EncodeTiledFn fn = encode_fn();
    if (!fn) return fail(ROMA_ERR_UNSUPPORTED, "cuTensorMapEncodeTiled is not available from this driver");
    cuuint64_t gd[5], gs[4];
    cuuint32_t bx[5], es[5];
    for (int i = 0; i < rank; ++i) { gd[i] = dims[i]; bx[i] = box[i]; es[i] = 1; }
    for (int i = 0; i + 1 < rank; ++i) gs[i] = strides_bytes[i];
    const CUresult rc = fn(map, data_type(roma_dtype), (cuuint32_t)rank, const_cast<void*>(base), gd, gs, bx, es,
                           CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (rc != CUDA_SUCCESS) return fail(ROMA_ERR_BAD_ARG, "cuTensorMapEncodeTiled failed (CUresult %d)", (int)rc);
    return 0;
}

// ---------------------------------------------------------------------------------------------- device side
__device__ __forceinline__ uint32_t saddr(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(saddr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(saddr(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = saddr(bar);
    uint32_t ok;
    do {
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void fence_barrier_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void prefetch_map(const CUtensorMap* map) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(map) : "memory");
}
// Box load, 4-D tiled map: coordinates innermost first.  Completion is signalled on `bar` (complete_tx::bytes).
__device__ __forceinline__ void load_4d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(saddr(smem_dst)), "l"(map), "r"(saddr(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void load_2d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(saddr(smem_dst)), "l"(map), "r"(saddr(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void load_3d(void* smem_dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(saddr(smem_dst)), "l"(map), "r"(saddr(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}

}  // namespace tma
}  // namespace roma
