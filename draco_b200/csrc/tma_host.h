// Host side of the TMA operand ring (tma_pipe.cuh): the tensor map of a set of factorisation workspaces.  Header-only so
// that libdgibbs.so and libdgnrml.so both carry it; the driver entry point is looked up at run time (no -lcuda).
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

namespace dg {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static inline EncodeTiledFn tma_encode_fn() {
  static EncodeTiledFn fn = []() -> EncodeTiledFn {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qr;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qr) != cudaSuccess ||
        qr != cudaDriverEntryPointSuccess) {
      cudaGetLastError();
      return nullptr;
    }
    return (EncodeTiledFn)p;
  }();
  return fn;
}

// `work` holds `nslots` row-major workspaces of `wrows` rows x `npad` doubles, `slot_stride` doubles apart.  The map
// views them as (k mod 4, row, k div 4, slot); one box of (4, tile, kc / 4, 1) lands in shared memory as
// [k div 4][row][k mod 4].  `ring_off_bytes`: offset of the ring in the kernel's dynamic shared memory (the bulk copies
// want 128-byte alignment).  false: keep the cp.async loaders (DGIBBS_NO_TMA set, alignment, or no driver entry point).
static inline bool tma_work_map(CUtensorMap* tm, double* work, size_t wrows, int npad, size_t slot_stride, int nslots,
                                size_t ring_off_bytes, int tile, int kc) {
  memset(tm, 0, sizeof *tm);
  static const bool off = getenv("DGIBBS_NO_TMA") != nullptr;
  if (off || work == nullptr || nslots <= 0) return false;
  if (ring_off_bytes % 128 != 0 || (slot_stride * sizeof(double)) % 16 != 0 || ((uintptr_t)work) % 16 != 0 || npad % 4 != 0)
    return false;
  EncodeTiledFn enc = tma_encode_fn();
  if (!enc) return false;
  const cuuint64_t dims[4] = {4, (cuuint64_t)wrows, (cuuint64_t)npad / 4, (cuuint64_t)nslots};
  const cuuint64_t strides[3] = {(cuuint64_t)npad * sizeof(double), 4 * sizeof(double), (cuuint64_t)slot_stride * sizeof(double)};
  const cuuint32_t box[4] = {4, (cuuint32_t)tile, (cuuint32_t)kc / 4, 1};
  const cuuint32_t estr[4] = {1, 1, 1, 1};
  return enc(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 4, work, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
             CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace dg
