// tma.cuh — mbarrier + TMA (cp.async.bulk.tensor) PTX wrappers for sm_100a and the host-side tensor-map helper.
// Shared by poisson_tb.cu (temporally blocked Poisson tiles) and cleandiv_tma.cu (row ring of the lexicographic sweep).
#pragma once
#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200cfd {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t phase) {
    asm volatile(
        "{\n .reg .pred p;\n"
        "WAIT_%=:\n"
        " mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        " @p bra DONE_%=;\n"
        " bra WAIT_%=;\n"
        "DONE_%=:\n}\n" ::"r"(smem_u32(bar)), "r"(phase) : "memory");
}
// global -> shared, 2-D box at element coordinates (x, y); out-of-bounds elements are zero-filled; completion on `bar`
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, int x, int y, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(x), "r"(y), "r"(smem_u32(bar)) : "memory");
}
// the same for a 3-D map (tma_make_3d_swz32: coordinates (0, x / 4, y))
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, int c0, int c1, int c2, uint64_t* bar) {
    asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
                 ::"r"(smem_u32(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}
// shared -> global, 2-D box; elements outside the tensor are not written.  Bulk-group completion.
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int x, int y, const void* src) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                 ::"l"(map), "r"(x), "r"(y), "r"(smem_u32(src)) : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* map, int c0, int c1, int c2, const void* src) {
    asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%1, %2, %3}], [%4];"
                 ::"l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(src)) : "memory");
}
__device__ __forceinline__ void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N> __device__ __forceinline__ void tma_store_wait() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }
// generic-proxy writes to shared memory -> visible to the async proxy (before a TMA store reads them)
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ---- host side ----------------------------------------------------------------------------------------------------
inline PFN_cuTensorMapEncodeTiled_v12000 tma_encoder() {
    static PFN_cuTensorMapEncodeTiled_v12000 fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) { cudaGetLastError(); p = nullptr; }
        return (PFN_cuTensorMapEncodeTiled_v12000)p;
    }();
    return fn;
}
// 2-D map of a row-major (ny, nx) field with row pitch `pitch` elements; box = box_w x box_h elements.  dtype 0 = fp64.
inline bool tma_make_2d(CUtensorMap* m, int dtype, const void* base, int nx, int ny, long long pitch, int box_w, int box_h) {
    PFN_cuTensorMapEncodeTiled_v12000 enc = tma_encoder();
    if (!enc) return false;
    const int es = dtype == 0 ? 8 : 4;
    cuuint64_t dims[2] = {(cuuint64_t)nx, (cuuint64_t)ny};
    cuuint64_t strides[1] = {(cuuint64_t)(pitch * es)};
    cuuint32_t box[2] = {(cuuint32_t)box_w, (cuuint32_t)box_h};
    cuuint32_t estr[2] = {1, 1};
    return enc(m, dtype == 0 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT64 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<void*>(base), dims,
               strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
               CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// The same fp64 field seen as (ny, nx / 4, 4): 32-byte units of 4 doubles, so that the box rows can be written to shared
// memory with the 32-byte swizzle (the swizzle modes need an inner box extent no larger than their span).  The box holds
// the same bytes in the same order as tma_make_2d's, except that inside every odd 128-byte line the two 16-byte halves of each
// 32-byte unit are exchanged (byte-address bit 4 ^= bit 7; the destination must be 256-byte aligned).  A warp whose lane l
// reads the 32-byte unit l of a row with two 128-bit loads then hits every bank once per quarter-warp; unswizzled, lanes l
// and l + 4 collide and each load takes twice the shared-memory cycles.  nx and box_w must be multiples of 4.
inline bool tma_make_3d_swz32(CUtensorMap* m, const void* base, int nx, int ny, long long pitch, int box_w, int box_h) {
    PFN_cuTensorMapEncodeTiled_v12000 enc = tma_encoder();
    if (!enc || nx % 4 || box_w % 4) return false;
    cuuint64_t dims[3] = {4, (cuuint64_t)(nx / 4), (cuuint64_t)ny};
    cuuint64_t strides[2] = {32, (cuuint64_t)(pitch * 8)};
    cuuint32_t box[3] = {4, (cuuint32_t)(box_w / 4), (cuuint32_t)box_h};
    cuuint32_t estr[3] = {1, 1, 1};
    return enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
               CU_TENSOR_MAP_SWIZZLE_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace b200cfd
