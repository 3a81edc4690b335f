// Shared helpers for the nsg_b200 kernels (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/nsg_b200.h"

namespace nsg {

// ---- error plumbing (thread local text, integer codes across the ABI)
void set_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define NSG_REQUIRE(cond, ...)      \
  do {                              \
    if (!(cond)) {                  \
      nsg::set_error(__VA_ARGS__);  \
      return NSG_E_INVALID;         \
    }                               \
  } while (0)

#define NSG_CHECK_LAUNCH(what)                                   \
  do {                                                           \
    cudaError_t e__ = cudaPeekAtLastError();                     \
    if (e__ != cudaSuccess) return nsg::cuda_fail(e__, what);    \
  } while (0)

int num_sms();

// ---- programmatic dependent launch (PDL).  A kernel launched through launch_pdl may become resident while its
// predecessor in the stream is still running (once every CTA of the predecessor has executed pdl_trigger() or
// exited).  It must execute pdl_wait() before its first access to memory an earlier kernel wrote or still reads:
// pdl_wait() returns when ALL earlier work in the stream has completed and is visible, so the dependency stays exact
// and only the launch latency and the prologue (barrier / TMEM set-up, issue tables, layer-constant loads) move under
// the predecessor's tail.  Discipline used by every kernel here: pdl_trigger() only AFTER pdl_wait().  Then at most
// ONE earlier kernel (the direct predecessor) can still be running while a prologue executes, and a prologue may read
// anything written two or more kernels back in the stream (the packed weight images of a layer: written once at the
// start of the step).  Both instructions are no-ops in a kernel launched without the attribute.  g_pdl = 0
// (NSG_PDL=0 / nsg_pdl(0)) turns the attribute off.
extern int g_pdl;
// Host-side bookkeeping behind "two or more kernels back": launch_pdl notes every launch per stream, conv_pack_many
// notes the packs; pdl_weights_settled(s) is true when at least one kernel of this library was launched on s since
// the last pack on s (launches by anybody else are not seen, which errs on the safe side; work on other streams
// reaches s only through events, i.e. complete).  Consumers preload their weight images above pdl_wait() only then.
void pdl_note_launch(cudaStream_t s);
void pdl_note_pack(cudaStream_t s);
bool pdl_weights_settled(cudaStream_t s);
// `urgent`: the launch carries the device's greatest priority.  Used by the grid-barrier kernels (conv_tma.cu, BnFuse):
// their resident CTAs spin until the LAST CTA of the grid has arrived, so when an SM frees up next to a kernel of
// another stream (the weight gradients run beside the data-gradient chain) the pending CTA of the barrier kernel
// must be the one that gets it.
int greatest_priority();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_ex(bool urgent, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  int n = 0;
  if (g_pdl) {
    at[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (urgent) {
    at[n].id = cudaLaunchAttributePriority;
    at[n].val.priority = greatest_priority();
    ++n;
  }
  cfg.attrs = at;
  cfg.numAttrs = n;
  pdl_note_launch(stream);
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              Args&&... args) {
  return launch_pdl_ex(false, kernel, grid, block, smem, stream, static_cast<Args&&>(args)...);
}
#if defined(__CUDA_ARCH__)
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#else
__device__ __forceinline__ void pdl_wait() {}
__device__ __forceinline__ void pdl_trigger() {}
#endif

// ---- activation storage types: float or bf16, always computed in fp32
template <typename T>
struct Act;
template <>
struct Act<float> {
  static __device__ __forceinline__ float ld(const float* p) { return __ldg(p); }
  static __device__ __forceinline__ float4 ld4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
  static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
  static __device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
};
template <>
struct Act<__nv_bfloat16> {
  static __device__ __forceinline__ float ld(const __nv_bfloat16* p) {
    return __bfloat162float(__ldg(p));
  }
  static __device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
    uint2 r = __ldg(reinterpret_cast<const uint2*>(p));
    float4 v;
    v.x = __uint_as_float(r.x << 16);
    v.y = __uint_as_float(r.x & 0xffff0000u);
    v.z = __uint_as_float(r.y << 16);
    v.w = __uint_as_float(r.y & 0xffff0000u);
    return v;
  }
  static __device__ __forceinline__ void st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
  static __device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
    uint2 r;
    r.x = *reinterpret_cast<uint32_t*>(&a);
    r.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = r;
  }
};

// prologue transform (NSG_PRO_*) on one value of channel c
struct Pro {
  int mode;
  const float* scale;
  const float* shift;
};
__device__ __forceinline__ float pro_apply(int mode, float x, float sc, float sh) {
  if (mode == NSG_PRO_RELU) return fmaxf(x, 0.f);
  if (mode == NSG_PRO_AFFINE_RELU) return fmaxf(fmaf(sc, x, sh), 0.f);
  if (mode == NSG_PRO_AFFINE) return fmaf(sc, x, sh);
  return x;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

#define NSG_DISPATCH_DTYPE(dtype, ...)                 \
  [&]() -> int {                                       \
    if ((dtype) == NSG_F32) {                          \
      using T = float;                                 \
      return __VA_ARGS__();                            \
    } else if ((dtype) == NSG_BF16) {                  \
      using T = __nv_bfloat16;                         \
      return __VA_ARGS__();                            \
    }                                                  \
    nsg::set_error("unknown dtype %d", (int)(dtype));  \
    return NSG_E_INVALID;                              \
  }()

}  // namespace nsg
