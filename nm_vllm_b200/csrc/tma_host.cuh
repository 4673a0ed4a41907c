// Host-side CUtensorMap construction without linking libcuda: the driver entry point is fetched through the runtime.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

inline PFN_encodeTiled get_encode_tiled() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

// 2-D row-major tensor [rows, cols] of `elem_bytes`-sized elements, row pitch `pitch_bytes`; box = [box_rows, box_cols];
// 128-byte swizzle (box_cols * elem_bytes must be <= 128).  Out-of-bounds elements read as zero.
inline bool make_tmap_2d(CUtensorMap* map, const void* base, CUtensorMapDataType dt, int elem_bytes, uint64_t rows,
                         uint64_t cols, uint64_t pitch_bytes, uint32_t box_rows, uint32_t box_cols,
                         CUtensorMapSwizzle swz = CU_TENSOR_MAP_SWIZZLE_128B,
                         CUtensorMapL2promotion l2 = CU_TENSOR_MAP_L2_PROMOTION_L2_256B) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) return false;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {pitch_bytes};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  (void)elem_bytes;
  CUresult r = enc(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                   l2, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

// activations [M, K] viewed as 3-D {64 (k in chunk), M (rows), K/64 (chunks)}: one box = {64, box_rows, box_chunks} lands
// as [box_chunks][box_rows][128 B] with the 128-byte swizzle a K-major UMMA operand expects
inline bool make_tmap_act3d(CUtensorMap* map, const void* base, CUtensorMapDataType dt, uint64_t rows, uint64_t k,
                            uint64_t pitch_bytes, uint32_t box_rows, uint32_t box_chunks) {
  PFN_encodeTiled enc = get_encode_tiled();
  if (!enc) return false;
  cuuint64_t dims[3] = {64, rows, k / 64};
  cuuint64_t strides[2] = {pitch_bytes, 128};
  cuuint32_t box[3] = {64, box_rows, box_chunks};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, dt, 3, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  return r == CUDA_SUCCESS;
}

}  // namespace b200
