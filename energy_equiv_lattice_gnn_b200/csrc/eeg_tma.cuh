// Tensor-map (TMA) plumbing shared by the tcgen05 kernels that load 2-D fp32 operands with cp.async.bulk.tensor:
// the host-side encoder (cuTensorMapEncodeTiled fetched through the runtime, so libcuda is not a link-time dependency) and
// the device-side 2-D tile load.  Boxes land dense ([box rows][box cols]) in shared memory, SWIZZLE_NONE; rows / columns
// outside the tensor are zero-filled by the hardware (and still count towards the mbarrier's transaction bytes).
#pragma once
#include <cuda.h>   // CUtensorMap (types only)
#include <cuda_runtime.h>
#include <stdint.h>
#include "eeg_umma.cuh"

namespace eeg_tma {

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                             const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// nullptr when the driver does not offer the encoder
static inline EncodeFn encoder() {
  static EncodeFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      p = nullptr;
    (void)cudaGetLastError();
    return reinterpret_cast<EncodeFn>(p);
  }();
  return fn;
}

// Row-major fp32 matrix [rows][cols] with `pitch` floats between rows, loaded in [box_rows x box_cols] boxes.
// base must be 16-byte aligned, pitch a multiple of 4 floats, box_cols * 4 a multiple of 16, both box extents <= 256.
static inline bool make_map_2d(CUtensorMap* map, const float* base, int64_t rows, int64_t cols, int64_t pitch, int box_cols, int box_rows) {
  EncodeFn enc = encoder();
  if (enc == nullptr) return false;
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows}, strides[1] = {(cuuint64_t)pitch * 4};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows}, estr[2] = {1, 1};
  return enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// box at (column c0, row r0) of the tensor map -> shared memory (128-byte aligned), completion on the mbarrier
__device__ __forceinline__ void load_2d(void* smem_dst, const CUtensorMap* map, int c0, int r0, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                   eeg_umma::smem_u32(smem_dst)),
               "l"(map), "r"(c0), "r"(r0), "r"(eeg_umma::smem_u32(bar))
               : "memory");
}

}  // namespace eeg_tma
