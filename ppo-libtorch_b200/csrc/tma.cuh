// tma.cuh — the bulk-tensor copy engine (TMA) for the hidden = 256 pipeline: 2-D tensor maps over 16-bit
// row-major activation scratch in HBM ([rows][256], 512 B per row) moved as [box_rows][64] boxes with the 128-byte
// swizzle, i.e. straight into / out of the shared-memory "activation panel" layout that umma.cuh's descriptors read
// (one 128-byte line per row, 16-byte chunk c of row r at ((c ^ (r & 7)) << 4)).
//
// Host side: cuTensorMapEncodeTiled is resolved through cudaGetDriverEntryPoint (no link against libcuda).
// Device side: cp.async.bulk.tensor.2d loads complete on an mbarrier (complete_tx::bytes); stores are tracked with
// bulk async-groups (commit_group / wait_group.read) — SASS: UTMALDG / UTMASTG.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "umma.cuh"

namespace ppo {
namespace tma {

// [rows][cols] 16-bit row-major tensor, boxes of box_rows x 64 elements (128 B wide), 128-byte swizzle.
// Returns cudaSuccess or an error; *ok_out = false when the driver entry point is missing.
inline int encode_2d_sw128(CUtensorMap* out, const void* base, uint64_t rows, uint64_t cols, uint32_t box_rows) {
    typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                 const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                 CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    static EncodeFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || !p) return -1;
        fn = reinterpret_cast<EncodeFn>(p);
    }
    const cuuint64_t gdim[2] = {cols, rows};
    const cuuint64_t gstride[1] = {cols * 2};
    const cuuint32_t box[2] = {64, box_rows};
    const cuuint32_t estr[2] = {1, 1};
    const CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void*>(base), gdim, gstride, box, estr,
                          CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                          CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)r;
}

#ifdef __CUDACC__
__device__ __forceinline__ void prefetch_map(const CUtensorMap* m) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(m)) : "memory");
}
// one thread: expect `bytes` on the barrier (counts as this thread's arrival)
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(umma::smem_u32(bar)), "r"(bytes) : "memory");
}
// global -> shared box at (col0, row0); completes `box bytes` on the barrier
__device__ __forceinline__ void load_2d(const CUtensorMap* m, uint32_t dst_smem, uint64_t* bar, int col0, int row0) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
        ::"r"(dst_smem), "l"(reinterpret_cast<uint64_t>(m)), "r"(col0), "r"(row0), "r"(umma::smem_u32(bar))
        : "memory");
}
// shared -> global box at (col0, row0); the caller fenced the generic-proxy writes (fence.proxy.async) beforehand
__device__ __forceinline__ void store_2d(const CUtensorMap* m, uint32_t src_smem, int col0, int row0) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];"
                 ::"l"(reinterpret_cast<uint64_t>(m)), "r"(col0), "r"(row0), "r"(src_smem)
                 : "memory");
}
__device__ __forceinline__ void store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
// every committed store of this thread has finished READING shared memory (the source may be overwritten)
__device__ __forceinline__ void store_wait_read() { asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory"); }
// ... and has been written out completely
__device__ __forceinline__ void store_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
#endif

}  // namespace tma
}  // namespace ppo
