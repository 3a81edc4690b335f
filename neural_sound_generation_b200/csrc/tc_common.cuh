// Inline-PTX wrappers for the Blackwell (sm_100a) tensor-core path: mbarrier, bulk async copy,
// TMEM allocation, tcgen05.mma / commit / ld, and the K-major SWIZZLE_128B operand layout.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

namespace nsg {
namespace tc {

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

// ---------------------------------------------------------------- mbarrier
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// ---------------------------------------------------------------- bounded waits
// Every wait in these kernels is bounded: a pipeline that can never advance (a lost TMA box, a miscounted arrival)
// must not hang the device.  After ~2 s without progress the waiting thread writes WHERE it was stuck -- translation
// unit, source line, block, thread, barrier parity -- into a few words of pinned HOST memory (they survive the dead
// context; nsg_hang_report() reads them) and traps.  The words are per translation unit (no relocatable device code
// in this build); nsg_device_ok() hands every unit the same host buffer before the first launch.
#ifndef NSG_TU_ID
#define NSG_TU_ID 0
#endif
static __device__ unsigned int* g_hang_words = nullptr;
static __device__ unsigned int g_hang_once = 0;     // first reporter of this translation unit wins
static __device__ __noinline__ void hang_trap(uint32_t line, uint32_t aux) {
  unsigned int* w = g_hang_words;
  if (w != nullptr && atomicCAS(&g_hang_once, 0u, 1u) == 0u) {
    volatile unsigned int* v = w;
    v[0] = 1u;
    v[1] = (uint32_t)NSG_TU_ID;
    v[2] = line;
    v[3] = blockIdx.x | (blockIdx.y << 20);
    v[4] = threadIdx.x;
    v[5] = aux;
    v[6] = gridDim.x;
    __threadfence_system();
  }
  __trap();
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
constexpr unsigned long long kHangNs = 2000000000ull;
__device__ __forceinline__ void mbar_wait_tagged(uint64_t* bar, uint32_t parity, uint32_t line) {
  if (mbar_try_wait(bar, parity)) return;
  const unsigned long long t0 = global_ns();
  for (uint32_t spins = 1;; ++spins) {
    if (mbar_try_wait(bar, parity)) return;
    if ((spins & 1023u) == 0 && global_ns() - t0 > kHangNs) hang_trap(line, parity);
  }
}
// for whole warps that wait on a tile-granular event: a polling warp is always eligible and takes issue slots from
// the working warps of its scheduler (measured: 25 % of all issued instructions in the VQ search were the poll loop)
__device__ __forceinline__ void mbar_wait_backoff_tagged(uint64_t* bar, uint32_t parity, uint32_t line) {
  if (mbar_try_wait(bar, parity)) return;
  const unsigned long long t0 = global_ns();
  for (uint32_t spins = 1;; ++spins) {
    __nanosleep(128);
    if (mbar_try_wait(bar, parity)) return;
    if ((spins & 255u) == 0 && global_ns() - t0 > kHangNs) hang_trap(line, parity | 0x100u);
  }
}
#define mbar_wait(bar, parity) mbar_wait_tagged(bar, parity, __LINE__)
#define mbar_wait_backoff(bar, parity) mbar_wait_backoff_tagged(bar, parity, __LINE__)

// ---------------------------------------------------------------- bulk async copy global -> shared (TMA engine, 1-D)
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst_smem)),
      "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// generic-proxy writes to shared memory -> visible to the async proxy (tensor core / TMA)
__device__ __forceinline__ void fence_proxy_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

// ---------------------------------------------------------------- TMEM
template <uint32_t COLS>
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem) {   // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)),
               "n"(COLS)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
template <uint32_t COLS>
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr) {     // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "n"(COLS) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// D[tmem] (+)= A[smem] * B[smem]^T, bf16 operands, fp32 accumulate; issued by ONE thread
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on an mbarrier when every previously issued tcgen05.mma of this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
// The same two, for an issue loop that the WHOLE warp runs convergently: elect.sync picks the issuing lane (the same
// one every time for a full mask -- tcgen05.commit tracks the operations of the thread that executes it) and only the
// tcgen05 instruction is predicated.  `valid` = 0 turns the instruction off without a branch.
__device__ __forceinline__ void umma_bf16_elect(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc,
                                                uint32_t accumulate, uint32_t valid = 1u) {
  asm volatile(
      "{\n\t.reg .pred p, q, v;\n\t"
      "setp.ne.b32 q, %4, 0;\n\t"
      "setp.ne.b32 v, %5, 0;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "and.pred p, p, v;\n\t"
      "@p tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, q;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate), "r"(valid)
      : "memory");
}
__device__ __forceinline__ void umma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "@p tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(bar))
      : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread i of the warp receives TMEM lane (base_lane + i)
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// pins the values of r to a point after the preceding tmem_ld_wait (volatile asm statements keep their order, and
// every later use of r depends on these): needed when other work is placed between a tmem_ld32 and its wait
__device__ __forceinline__ void tmem_ld_fence(uint32_t (&r)[32]) {
#pragma unroll
  for (int i = 0; i < 32; ++i) asm volatile("" : "+r"(r[i])::"memory");
}

// ---------------------------------------------------------------- operand layout
// K-major, SWIZZLE_128B canonical layout: rows of 64 bf16 (128 B); 8-row groups of 1024 B; inside a group
// the 16-byte chunk c of row r is stored at chunk (c ^ (r & 7)).  Buffers must be 1024-byte aligned.
__device__ __host__ __forceinline__ uint32_t sw128_offset(uint32_t row, uint32_t chunk16) {
  return row * 128u + ((chunk16 ^ (row & 7u)) << 4);
}
// shared-memory matrix descriptor (sm_100 format): start>>4 | LBO=1 | SBO=1024>>4 | version=1 | SWIZZLE_128B
__device__ __forceinline__ uint64_t sw128_desc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | (1ull << 16) | (64ull << 32) | (1ull << 46) | (2ull << 61);
}
// instruction descriptor, kind::f16: D=f32, A=B=bf16, both K-major, dense
__device__ __host__ constexpr uint32_t idesc_bf16(uint32_t m, uint32_t n) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
// x = hi + lo with hi = bf16(x), lo = bf16(x - hi)
__device__ __forceinline__ void split_bf16(float x, float& hi, float& lo) {
  hi = __bfloat162float(__float2bfloat16_rn(x));
  lo = x - hi;
}

}  // namespace tc
}  // namespace nsg
