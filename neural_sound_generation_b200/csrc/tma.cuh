// Tensor-map (TMA) plumbing: host-side encoding of 4-D NHWC activation views through the driver entry point
// (no link-time dependency on libcuda) and the device-side cp.async.bulk.tensor wrappers.
#pragma once
#include <cuda.h>

#include "common.cuh"
#include "tc_common.cuh"

namespace nsg {
namespace tma {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static inline EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// bf16 NHWC tensor with 64 channels seen as a (possibly strided) pixel grid:
//   element (c, w, h, n) at base + 2*(c + w*w_stride + h*h_stride + n*n_stride)   [strides in elements]
// box = 64 channels x box_w x box_h x 1 image, 128-byte swizzle (one pixel = one 128-byte row), zero fill outside.
static inline int encode_nhwc64(CUtensorMap* tm, const void* base, long long dim_w, long long dim_h, long long dim_n,
                                long long w_stride, long long h_stride, long long n_stride, int box_w, int box_h) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return NSG_E_CUDA;
  }
  if (dim_w < 1) dim_w = 1;
  if (dim_h < 1) dim_h = 1;
  cuuint64_t dims[4] = {64, (cuuint64_t)dim_w, (cuuint64_t)dim_h, (cuuint64_t)dim_n};
  cuuint64_t strides[3] = {(cuuint64_t)w_stride * 2, (cuuint64_t)h_stride * 2, (cuuint64_t)n_stride * 2};
  cuuint32_t box[4] = {64, (cuuint32_t)box_w, (cuuint32_t)box_h, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 4, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for dims (64,%lld,%lld,%lld) box (%d,%d)", (int)r, dim_w, dim_h,
              dim_n, box_w, box_h);
    return NSG_E_CUDA;
  }
  return NSG_OK;
}

#ifdef __CUDACC__
__device__ __forceinline__ void prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(tm)) : "memory");
}
// 4-D tile load: coordinates (channel, w, h, n); completes `bytes of the box` on the mbarrier
__device__ __forceinline__ void load_4d(void* dst_smem, const CUtensorMap* tm, int c0, int c1, int c2, int c3,
                                        uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], "
      "[%6];" ::"r"(tc::smem_u32(dst_smem)),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(tc::smem_u32(bar))
      : "memory");
}
__device__ __forceinline__ void load_4d_u32(uint32_t dst_smem, const CUtensorMap* tm, int c0, int c1, int c2, int c3,
                                            uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], "
      "[%6];" ::"r"(dst_smem),
      "l"(reinterpret_cast<uint64_t>(tm)), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(tc::smem_u32(bar))
      : "memory");
}
#endif

}  // namespace tma
}  // namespace nsg
