// Cross-GPU statistic exchange over peer-mapped ("symmetric") memory, shared by the kernels that need the statistics
// of the GLOBAL batch (conv_tma.cu: grid-barrier launches and the one-CTA stand-alone form; fused_elem.cu: bn_act / bn_bwd).
#pragma once
#include "common.cuh"
#include "tc_common.cuh"

namespace nsg {

using namespace tc;

constexpr int PEER_C = 64;
// Synchronised BatchNorm (one process per GPU; every rank must normalise with the statistics of the GLOBAL batch to
// equal the single-process reference): the 128 fp64 partial sums of a layer are exchanged through peer-mapped
// ("symmetric") memory over NVLink inside the kernel that needs them, instead of a 1 KB ncclAllReduce launch per
// layer and direction (~26 us each, 20 per step).  Every rank owns an area
//     slots [NSG_PEER_SLOTS][8 ranks][128] fp64   |   flags [NSG_PEER_SLOTS][8 ranks] u32
// mapped on all ranks.  One exchange (slot s) on rank r:  write my sums into slot (s, r) of EVERY rank; fence; raise
// flag (s, r) on every rank; wait until my flags (s, 0..world-1) are up; sum my slots (s, 0..world-1) in rank order
// (every rank adds the same numbers in the same order: bit-identical results everywhere); lower my flags.  A slot is
// used once per step, and a rank can only reach it again after the step's gradient all-reduce, which every rank enters
// after its own last exchange -- so lowering the flags after the read cannot race with the next step's writers.
struct PeerSync {
  int world, rank, slot;
  double* slots[NSG_PEER_MAX];
  unsigned int* flags[NSG_PEER_MAX];
};
constexpr size_t kPeerSlotDoubles = (size_t)NSG_PEER_MAX * 2 * PEER_C;

// executed by 128+ threads of ONE CTA (thread index e = 0.. within the group, `bar` = their named barrier id / count);
// sums[0..127] in global memory is replaced by the sum over ranks
__device__ __forceinline__ void peer_exchange(const PeerSync& ps, double* sums, int e, int bar_id, int bar_threads) {
  const size_t base = (size_t)ps.slot * kPeerSlotDoubles;
  if (e < 2 * PEER_C) {
    const double v = __ldcg(sums + e);
    for (int p = 0; p < ps.world; ++p)
      asm volatile("st.relaxed.sys.global.f64 [%0], %1;" ::"l"(ps.slots[p] + base + (size_t)ps.rank * 2 * PEER_C + e), "d"(v) : "memory");
  }
  __threadfence_system();
  asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(bar_threads) : "memory");
  if (e < ps.world) {
    unsigned int* mine_on_peer = ps.flags[e] + (size_t)ps.slot * NSG_PEER_MAX + ps.rank;
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(mine_on_peer), "r"(1u) : "memory");
    const unsigned int* theirs = ps.flags[ps.rank] + (size_t)ps.slot * NSG_PEER_MAX + e;
    unsigned int v = 0;
    const unsigned long long t0 = global_ns();
    for (;;) {                                         // ranks may be seconds apart (first steps): ~40 s
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(theirs) : "memory");
      if (v != 0 || global_ns() - t0 > 20 * kHangNs) break;
      __nanosleep(64);
    }
    if (v == 0) hang_trap(__LINE__, (uint32_t)e);   // a rank that never arrives must not hang the device
  }
  asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(bar_threads) : "memory");
  if (e < 2 * PEER_C) {
    double acc = 0.0;
    for (int r = 0; r < ps.world; ++r) {
      double v;
      asm volatile("ld.relaxed.sys.global.f64 %0, [%1];" : "=d"(v) : "l"(ps.slots[ps.rank] + base + (size_t)r * 2 * PEER_C + e) : "memory");
      acc += v;
    }
    __stcg(sums + e, acc);
  }
  __threadfence();
  asm volatile("bar.sync %0, %1;" ::"r"(bar_id), "r"(bar_threads) : "memory");
  if (e < ps.world) ps.flags[ps.rank][(size_t)ps.slot * NSG_PEER_MAX + e] = 0u;
}


// validated device-side copy of the public nsg_peer_sync (conv_tma.cu)
int make_peer(const nsg_peer_sync* p, PeerSync* out);

}  // namespace nsg
