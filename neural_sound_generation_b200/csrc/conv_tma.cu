// TMA-fed, persistent tcgen05 convolution kernels for the 64 -> 64 channel layers over OPERAND-READY bf16 NHWC
// activations (the BatchNorm/ReLU that precedes a convolution in models.py:148-183 has already been applied by the
// producer of the tensor, see fused_elem.cu), so the tensor core reads exactly what the TMA engine wrote:
//
//   conv_tma  (Conv2d / ConvTranspose2d forward and both data gradients):
//       one CTA per SM, looping over 128-pixel output tiles (box_w x box_h pixels of one image);
//       warp 0   : TMA producer -- all taps' packed weights once (bulk copies, resident for the CTA's lifetime),
//                  then per tile and tap one 4-D tensor-map load of the shifted 128 x 64 input box into a ring stage;
//                  zero padding, image borders and ragged tiles are the TMA unit's out-of-bounds zero fill;
//                  stride-2 gathers use four parity views of the tensor (strided tensor maps);
//       warp 1   : MMA issuer -- 4 tcgen05.mma (M128 N64 K16) per tap into one of two 64-column TMEM accumulators;
//       warps 2-9: two epilogue groups (one per accumulator): tcgen05.ld, bias / residual add / ReLU mask /
//                  BatchNorm statistics (forward moments or backward reductions), bf16 store.  The epilogue of tile i
//                  overlaps the loads and MMAs of tile i+1.
//   wgrad_tma (weight gradient): pixels are the GEMM K axis; the 64-pixel TMA boxes of the gathered operand P (one
//       per tap) and of Q are MN-major SWIZZLE_128B operands as they land; two taps share one M = 128 instruction;
//       an all-ones slot yields the bias gradient; partial sums leave through 16-byte vector atomics into a packed
//       [tap][a][b] fp32 image that wgrad_unpack folds into the reference's parameter layout.
#include <cstdlib>
#include "common.cuh"
#include "conv_geom.cuh"
#define NSG_TU_ID 2
#include "tc_common.cuh"
#include "tma.cuh"
#include "peer.cuh"

namespace nsg {

using namespace tc;

constexpr int TM_M = 128;                       // pixels per tile
constexpr int WG_THREADS = 224;                 // wgrad_tma_kernel: producer warp, 2 MMA-issuing warps, 4 epilogue warps
constexpr int TM_THREADS = 352;                 // conv_tma_kernel: producer warp, 2 MMA-issuing warps, 8 epilogue warps
constexpr int TM_C = 64;                        // channels
constexpr uint32_t TM_A_BYTES = TM_M * 128;     // one activation box
constexpr uint32_t TM_W_BYTES = TM_C * 128;     // one tap's weight image (bf16, K-major SWIZZLE_128B)
constexpr int TM_MAX_STAGES = 8;
constexpr int TM_MAX_TAPS = 16;
constexpr int TM_TILE_TABLE = 32;
// Accumulators per tile the K = 16 steps rotate over (the epilogue adds them).  Measured (tools/microbench/mma_rate.cu):
// back-to-back tcgen05.mma into ONE accumulator already run at the shared-memory operand rate (48 cycles for
// M128 N64 K16), so independent chains buy nothing: 1.
constexpr int TM_CHAINS = 1;
constexpr int TM_ACC_COLS = TM_CHAINS * TM_C;       // TMEM columns per tile accumulator set                 // local tiles with precomputed coordinates (more: decoded on the fly)
constexpr uint32_t TM_SMEM_LIMIT = 223 * 1024;      // dynamic part; the kernels also hold <= 4 KB of static shared memory

// Taps that differ only by a vertical shift share ONE activation box: a "group" is (tensor map, horizontal shift,
// first row) and its box holds box_h + halo rows of box_w pixels; tap t of the group reads the 128-pixel sub-block
// that starts roff[t] rows into the box (box_w is a multiple of 8 pixels, so every sub-block starts on a 1024-byte
// swizzle atom).  A 3x3 tap set costs 3 boxes per tile instead of 9 -- the L2 -> SM path is what bounds these kernels.
struct TmaPhase {
  int sub_h, sub_w;              // output sub-grid of this stride phase
  int tiles_h, tiles_w, tile0;   // tile grid and index of its first tile
  int ntaps, ngroups;
  int out_stride, off_h, off_w;  // oh = ohs * out_stride + off_h
  signed char gdh[TM_MAX_TAPS], gdw[TM_MAX_TAPS];          // per group: box origin relative to the tile origin
  unsigned char gmap[TM_MAX_TAPS], gfirst[TM_MAX_TAPS + 1];  // per group: tensor map, first tap (taps sorted by group)
  unsigned short aoff[TM_MAX_TAPS];                        // per tap: offset of its sub-block inside the box, 16-byte units
  unsigned char wslot[TM_MAX_TAPS];                        // per tap: weight image
};

struct TmaConvParams {
  int batch, full_h, full_w;
  int in_stride;                 // 1: one tensor map; 2: four parity maps
  int box_w, box_h, log2_bw;
  int halo;                      // extra box rows shared by vertically shifted taps
  uint32_t slot_bytes, row_bytes;  // one ring slot; one box row (pixels per box line x 128 bytes)
  uint32_t box_bytes;            // bytes one TMA box delivers (<= slot_bytes: slots are 1024-byte multiples)
  uint32_t sbo16;                // byte distance between the 8-pixel groups of the A operand, 16-byte units
  int swap;                      // 1: the box's fast pixel axis is H ("w" quantities of the tile grid run along H)
  int issuers;                   // 1 / 2 MMA-issuing threads (2: local tiles alternate between warps 1 and 10)
  int elect;                     // 1: the issuing warps run converged and elect.sync issues (NSG_CONV_ELECT)
  int nphases, total_tiles, nw, stages;
  unsigned long long* dbg;       // optional clock64 stamps of CTA 0 (nsg_debug_timeline), null in production
  int w_early;                   // 1: the weight images were written >= 2 kernels back: load them above pdl_wait()
  TmaPhase ph[4];
};

__device__ __forceinline__ void tm_stamp(unsigned long long* dbg, int slot) {
  if (dbg && blockIdx.x == 0) dbg[slot] = (unsigned long long)clock64();
}

// ---------------------------------------------------------------------------------------------- packing
struct PackItem {
  const float* w;
  uint8_t* dst;
  int ntaps;
  int kmax, rmax;                // K indices >= kmax and rows >= rmax are written as zeros (small operands)
  long long in_stride, out_stride;
};
struct PackList {
  int n;
  int first_block[25];           // prefix of blocks (2 per tap)
  PackItem it[24];
};

// bf16 image per tap: rows = GEMM output channel, 64 GEMM input channels along K, K-major SWIZZLE_128B
__global__ void __launch_bounds__(256) conv_pack_many_kernel(const __grid_constant__ PackList L) {
  pdl_wait();
  pdl_trigger();
  int li = 0;
  while (li + 1 < L.n && (int)blockIdx.x >= L.first_block[li + 1]) ++li;
  const PackItem& p = L.it[li];
  const int gid = ((int)blockIdx.x - L.first_block[li]) * 256 + threadIdx.x;
  const int c = gid & 7, r = (gid >> 3) & 63, t = gid >> 9;
  if (t >= p.ntaps) return;
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i)
    v[i] = (c * 8 + i < p.kmax && r < p.rmax)
               ? __ldg(p.w + t + (long long)(c * 8 + i) * p.in_stride + (long long)r * p.out_stride) : 0.f;
  *reinterpret_cast<uint4*>(p.dst + (size_t)t * TM_W_BYTES + sw128_offset(r, c)) =
      make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}

// ---------------------------------------------------------------------------------------------- helpers
__device__ __forceinline__ void ld8bf(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void unpack8bf(const uint4& r, float (&v)[8]) {
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void st8bf(__nv_bfloat16* p, const float* v) {
  *reinterpret_cast<uint4*>(p) =
      make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}
// 16 bf16 = one whole 32-byte L2 sector per store instruction.  With 16-byte stores every sector of the output is
// written in two partial pieces, and the L2 fetched it from DRAM to merge them: the 1 -> 64 forward read 42 MB of DRAM
// against 2.6 MB of algorithmic input (ncu, profiles/r1_final_ncu_full_summary.txt).  p must be 32-byte aligned.
__device__ __forceinline__ void st16bf(__nv_bfloat16* p, const float* v) {
  asm volatile("st.global.v8.b32 [%0], {%1, %2, %3, %4, %5, %6, %7, %8};" ::"l"(p), "r"(pack_bf16(v[0], v[1])),
               "r"(pack_bf16(v[2], v[3])), "r"(pack_bf16(v[4], v[5])), "r"(pack_bf16(v[6], v[7])),
               "r"(pack_bf16(v[8], v[9])), "r"(pack_bf16(v[10], v[11])), "r"(pack_bf16(v[12], v[13])),
               "r"(pack_bf16(v[14], v[15]))
               : "memory");
}
// v[i] of lane l -> sum over lanes of v[lane] (32x32 transpose-reduce); v is destroyed
__device__ __forceinline__ float transpose_sum32(float (&v)[32], int lane) {
#pragma unroll
  for (int half = 16; half >= 1; half >>= 1) {
    const bool up = (lane & half) != 0;
#pragma unroll
    for (int i = 0; i < half; ++i) {
      const float keep = up ? v[i + half] : v[i];
      const float send = up ? v[i] : v[i + half];
      v[i] = keep + __shfl_xor_sync(0xffffffffu, send, half);
    }
  }
  return v[0];
}
__device__ __forceinline__ void named_bar(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

// ---------------------------------------------------------------------------------------------- conv_tma kernel
struct TmaTile {
  int p, n, th, tw;
};
__device__ __forceinline__ TmaTile decode_tile(const TmaConvParams& P, int t) {
  TmaTile r;
  r.p = 0;
#pragma unroll
  for (int i = 1; i < 4; ++i)
    if (i < P.nphases && t >= P.ph[i].tile0) r.p = i;
  const TmaPhase& ph = P.ph[r.p];
  int local = t - ph.tile0;
  r.tw = local % ph.tiles_w;
  local /= ph.tiles_w;
  r.th = local % ph.tiles_h;
  r.n = local / ph.tiles_h;
  return r;
}

__device__ __forceinline__ TmaTile fetch_tile(const TmaConvParams& P, const int4* table, int i) {
  if (i < TM_TILE_TABLE) {
    const int4 v = table[i];
    TmaTile r;
    r.p = v.x; r.n = v.y; r.th = v.z; r.tw = v.w;
    return r;
  }
  return decode_tile(P, (int)blockIdx.x + i * (int)gridDim.x);
}

// (PeerSync / peer_exchange: the cross-GPU statistic exchange over peer-mapped memory lives in peer.cuh)
// stand-alone form for the statistics that are not produced by a grid-barrier launch (encoder.1 forward, the two
// backward reductions fed by the single-channel / bottleneck kernels): one CTA, replaces ncclAllReduce of 1 KB
__global__ void __launch_bounds__(128) peer_allreduce_kernel(double* sums, PeerSync ps) {
  pdl_wait();
  pdl_trigger();
  peer_exchange(ps, sums, threadIdx.x, 3, 128);
}

// BatchNorm fused across a grid barrier (NACC > 2, "hold" launches).  A CTA of the persistent grid owns at most NACC
// tiles; their fp32 accumulators all stay in TMEM (NACC x 64 columns).  Pass 1 (overlapping the remaining MMAs) reads
// each accumulator once for the per-channel sums -- BatchNorm moments of the convolution output (forward) or the
// BatchNorm-backward reductions of the masked data gradient (backward) -- the CTAs meet at a counter in global memory,
// and pass 2 reads the accumulators again and stores the tensor the NEXT convolution reads (normalised + ReLU
// [+ residual] forward; dh = gamma*invstd*(g - mean(g) - xhat*mean(g*xhat)) backward).  The separate bn_act / bn_bwd
// launch of fused_elem.cu and its read + write of the whole activation disappear (models.py:151-158).
struct BnFuse {
  int mode;                        // NSG_BNF_*
  int relu_out;
  const float* gamma;
  const float* beta;
  float eps, momentum;
  double count;                    // elements per channel behind the statistics
  float* running_mean;
  float* running_var;
  long long* nbt;
  float* coef;                     // TRAIN: [4][64] mean | invstd | scale | shift published by CTA 0 (nullable)
  const __nv_bfloat16* res;        // TRAIN: residual added after the normalisation (nullable)
  unsigned int* barrier;           // zero-initialised counter of this launch
  float* dgamma;
  float* dbeta;
  __nv_bfloat16* out2;
  PeerSync peer;                   // world > 1: the statistics are summed over the GPUs of the node inside the barrier
};

// All CTAs of the grid are resident (grid <= SM count, 1 CTA per SM; kernels running beside this one on other streams
// never wait for it), so a counter + spin is a legal grid barrier.  A lost CTA must not hang the device: the spin is
// bounded (2 s) and traps with a report (tc_common.cuh).
__device__ __forceinline__ void grid_barrier_wait(unsigned int* ctr, unsigned int expected,
                                                  unsigned long long limit_ns = kHangNs) {
  unsigned int v = 0;
  const unsigned long long t0 = global_ns();
  for (;;) {
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
    if (v >= expected) break;
    if (global_ns() - t0 > limit_ns) hang_trap(__LINE__, v);
    __nanosleep(64);
  }
  __threadfence();
}
__device__ __forceinline__ void grid_barrier(unsigned int* ctr, unsigned int expected) {
  __threadfence();
  atomicAdd(ctr, 1u);
  grid_barrier_wait(ctr, expected);
}

// operands of one (row, 32-channel half): 64 bytes each
__device__ __forceinline__ void ld_half(const void* src, long long o, uint4 (&r)[4]) {
  const uint4* p = reinterpret_cast<const uint4*>(static_cast<const __nv_bfloat16*>(src) + o);
#pragma unroll
  for (int g8 = 0; g8 < 4; ++g8) r[g8] = __ldg(p + g8);
}
// x[j] (op)= operand channel j
__device__ __forceinline__ void add_half(float (&x)[32], const uint4 (&r)[4]) {
#pragma unroll
  for (int g8 = 0; g8 < 4; ++g8) {
    float a[8];
    unpack8bf(r[g8], a);
#pragma unroll
    for (int j = 0; j < 8; ++j) x[g8 * 8 + j] += a[j];
  }
}
__device__ __forceinline__ void mask_half(float (&x)[32], const uint4 (&r)[4], int mask_mode, const float* sPar, int ch0) {
#pragma unroll
  for (int g8 = 0; g8 < 4; ++g8) {
    float mk[8];
    unpack8bf(r[g8], mk);
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int ch = ch0 + g8 * 8 + j;
      const float mv = mask_mode == NSG_MASK_AFFINE ? fmaf(sPar[TM_C + ch], mk[j], sPar[2 * TM_C + ch]) : mk[j];
      x[g8 * 8 + j] = mv > 0.f ? x[g8 * 8 + j] : 0.f;
    }
  }
}
// the streaming part of the epilogue for one (row, half): x = osc*acc + osh [+ add] [masked] [ReLU]
// (per-channel coefficients come from shared memory 4 at a time: 8 LDS.128 per 32 channels instead of 32 LDS.32; the
// output scale differs from 1 only when eval-mode BatchNorm is folded in -- `scaled` -- and is not even loaded otherwise)
__device__ __forceinline__ void ld32s(const float* p, float (&v)[32]) {
  const float4* q = reinterpret_cast<const float4*>(p);
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float4 t = q[i];
    v[4 * i] = t.x; v[4 * i + 1] = t.y; v[4 * i + 2] = t.z; v[4 * i + 3] = t.w;
  }
}
__device__ __forceinline__ void epi_half(float (&x)[32], const uint32_t (&acc)[32], const EpiArgs& epi, const float* sPar,
                                         long long oh, bool valid, int ch0, bool use_add, const uint4 (&radd)[4],
                                         bool scaled) {
  {
    float sh[32];
    ld32s(sPar + ch0, sh);
    if (scaled) {
      float sc[32];
      ld32s(sPar + 5 * TM_C + ch0, sc);
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = valid ? fmaf(__uint_as_float(acc[j]), sc[j], sh[j]) : 0.f;
    } else {
#pragma unroll
      for (int j = 0; j < 32; ++j) x[j] = valid ? __uint_as_float(acc[j]) + sh[j] : 0.f;
    }
  }
  if (use_add) add_half(x, radd);
  if (valid && epi.mask_mode != NSG_MASK_NONE) {
    uint4 rm[4];
    ld_half(epi.mask_src, oh, rm);
    mask_half(x, rm, epi.mask_mode, sPar, ch0);
  }
  if (epi.tanh_out == 2) {
#pragma unroll
    for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
  }
}

// The MMA-issuing loop of conv_tma_kernel, issuer `me` (0: warp 1, 1: warp 10).  ELECT = false: lane 0 runs it alone.
// ELECT = true: the WHOLE warp runs it, converged -- every value is warp-uniform, all lanes poll the barriers, and only
// the tcgen05 instructions are predicated on elect.sync (tc_common.cuh); without the divergent `if (lane == 0)` the
// compiler needs no ELECT / BRA.U.ANY loop around every UTCHMMA.
template <int NACC, bool ELECT>
__device__ __forceinline__ void conv_issue_tiles(const TmaConvParams& P, int me, int lane, int ntiles_mine, const int4* sTile,
                                                 const int* sGrp, const uint32_t* sTapA, const uint32_t* sTapB, uint8_t* sA,
                                                 uint64_t* full, uint64_t* empty, uint64_t* wfull, uint64_t* acc_full,
                                                 uint64_t* acc_empty, uint32_t tmem_base) {
  {
    const int stages = P.stages;
    const bool split = P.issuers == 2;                   // tile i belongs to issuer i & 1
    if ((ELECT || lane == 0) && ntiles_mine > me && (me == 0 || split)) {
      const uint32_t idesc = idesc_bf16(TM_M, TM_C);
      mbar_wait(wfull, 0);
      if (me == 0) tm_stamp(P.dbg, 6);                   // weights resident
      // this issuer's part of the slot ring (the producer above)
      const int ring0 = split && me ? stages >> 1 : 0;
      const int ring_size = split ? (me ? stages - (stages >> 1) : stages >> 1) : stages;
      int it = 0, s = 0;
      uint32_t fparity = 0;
      const uint32_t slot_units = P.slot_bytes >> 4;
      const uint32_t a_base = (((smem_u32(sA) >> 4) + (uint32_t)ring0 * slot_units) & 0x3FFFu) | (1u << 16);     // low word of sw128_desc(first slot)
      uint64_t* const rfull = full + ring0;
      uint64_t* const rempty = empty + ring0;
      // SWIZZLE_128B K-major descriptors.  A: 8-pixel groups P.sbo16 * 16 bytes apart (1024 when a tile line is a whole
      // number of groups; the box line pitch in the one-box form, where a group is ONE box line of 8 pixels and the
      // +-1-pixel taps start 128 bytes off the swizzle atom -- the hardware applies the swizzle to the absolute
      // shared-memory address, exactly as the TMA unit wrote it: tools/microbench/desc_probe.cu).  B: the weight images.
      const uint64_t DESC_A_HI = ((uint64_t)P.sbo16 << 32) | (1ull << 46) | (2ull << 61);
      constexpr uint64_t DESC_B_HI = (64ull << 32) | (1ull << 46) | (2ull << 61);
      for (int i = me; i < ntiles_mine; i += split ? 2 : 1) {
        const TmaTile tl = fetch_tile(P, sTile, i);
        const TmaPhase& ph = P.ph[tl.p];
        const int ngroups = ph.ngroups;
        // streaming: the accumulators in turn.  hold: tile i lives in accumulator i % NACC; with more tiles than
        // accumulators the FIRST ntiles - NACC tiles give theirs up after pass 1 (pass 2 re-reads them from memory)
        const int as = i % NACC, use = i / NACC;
        if (use > 0) mbar_wait(&acc_empty[as], (uint32_t)((use - 1) & 1));
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)as * TM_ACC_COLS;
        for (int g = 0; g < ngroups; ++g, ++it) {
          // the group's tap descriptors are fetched four at a time before they are needed: the issuing thread runs
          // dependent scalar code, so every load left between two tcgen05.mma shows up as an idle tensor pipe
          const uint32_t gi = (uint32_t)sGrp[tl.p * TM_MAX_TAPS + g];
          const int first = (gi >> 24) & 15, cnt = (int)(gi >> 28) + 1;
          uint32_t ao[4], bo[4];
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int e = tl.p * TM_MAX_TAPS + ((first + j) & 15);
            ao[j] = sTapA[e];
            bo[j] = sTapB[e];
          }
          mbar_wait(&rfull[s], fparity);
          if (it == 0 && me == 0) tm_stamp(P.dbg, 7);               // first activation box landed
          if (i == 1 && g < 3) tm_stamp(P.dbg, 21 + 3 * g);      // 21,24,27: tile 1, box of group g landed
          tc_fence_after();
          const uint32_t box = a_base + (uint32_t)s * slot_units;
          for (int j0 = 0; j0 < cnt; j0 += 4) {
            uint32_t an[4], bn_[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {                // (the next four, for groups of more than four taps)
              const int e = tl.p * TM_MAX_TAPS + ((first + j0 + 4 + j) & 15);
              an[j] = sTapA[e];
              bn_[j] = sTapB[e];
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const uint64_t a = DESC_A_HI | (uint64_t)(box + ao[j]);
              const uint64_t b = DESC_B_HI | (uint64_t)bo[j];
              const uint32_t accum = (first + j0 + j) != 0;
              if (ELECT) {                               // (taps past the group's count are predicated off: no branch)
                const uint32_t valid = (j0 + j < cnt) ? 1u : 0u;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                  umma_bf16_elect(d + (ks % TM_CHAINS) * TM_C, a + 2 * ks, b + 2 * ks, idesc,
                                  (ks >= TM_CHAINS) ? 1u : accum, valid);
              } else if (j0 + j < cnt) {
#pragma unroll
                for (int ks = 0; ks < 4; ++ks)
                  umma_bf16(d + (ks % TM_CHAINS) * TM_C, a + 2 * ks, b + 2 * ks, idesc,
                            (ks >= TM_CHAINS) ? 1u : accum);
              }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              ao[j] = an[j];
              bo[j] = bn_[j];
            }
          }
          if (i == 1 && g < 3) tm_stamp(P.dbg, 22 + 3 * g);      // 22,25,28: its instructions issued
          if (ELECT) umma_commit_elect(&rempty[s]); else umma_commit(&rempty[s]);
          if (i == 1 && g < 3) tm_stamp(P.dbg, 23 + 3 * g);      // 23,26,29: commit issued
          if (++s == ring_size) {
            s = 0;
            fparity ^= 1;
          }
        }
        if (ELECT) umma_commit_elect(&acc_full[as]); else umma_commit(&acc_full[as]);
        tm_stamp(P.dbg, 8 + (i < 3 ? i : 3));           // 8..11: MMAs of tile i issued
      }
    }
  }
}

// HOLD = false: streaming (NACC accumulators in turn, the epilogue of tile i overlaps the MMAs of the following tiles).
// HOLD = true: every tile of the CTA keeps its own accumulator for the second pass (BnFuse above).
// Warps: 0 = TMA producer, 1 and 10 = MMA issuers (even / odd local tiles when P.issuers == 2), 2-9 = epilogue.
// Two issuers because ONE thread cannot keep the tensor pipe busy at N = 64: the uniform-datapath sequence around each
// tcgen05.mma (elect, descriptor adds, branch) costs 75-85 cycles in this kernel against the instruction's 48, and the
// hand-over between two tiles 700-1400 more.
template <int NACC, bool HOLD>
__global__ void __launch_bounds__(TM_THREADS, 1)
conv_tma_kernel(const __grid_constant__ CUtensorMap tm0, const __grid_constant__ CUtensorMap tm1,
                const __grid_constant__ CUtensorMap tm2, const __grid_constant__ CUtensorMap tm3,
                const uint8_t* __restrict__ wimg, __nv_bfloat16* __restrict__ out, EpiArgs epi, BnFuse bn,
                const __grid_constant__ TmaConvParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  uint8_t* sW = sm;                                         // nw weight images
  uint8_t* sA = sm + (size_t)P.nw * TM_W_BYTES;             // ring of activation boxes
  uint8_t* tail = sA + (size_t)P.stages * P.slot_bytes;
  // [9][64]: 0 output shift (bias), 1 mscale, 2 mshift, 3 mean, 4 invstd, 5 output scale, 6-8 pass-2 coefficients
  __shared__ __align__(16) float sPar[9 * TM_C];
  // [lane quarter][2][64] partial statistic sums: written once, after the CTA's last tile, when every TMA load has been
  // consumed -- the first 4 KB of the (then idle) activation ring serve (static shared memory is full: 223 KB dynamic
  // + 4 KB static is the 227 KB limit)
  float* sStat = reinterpret_cast<float*>(sA);
  // issue tables: the producer and MMA threads are single threads running dependent scalar code, so everything they
  // would otherwise derive per tile / tap (integer divisions, byte loads from the parameter bank, descriptor
  // packing) is computed once, in parallel, here
  __shared__ uint32_t sTapA[4 * TM_MAX_TAPS];               // per (phase, tap): A sub-block offset, descriptor units
  __shared__ uint32_t sTapB[4 * TM_MAX_TAPS];               // per (phase, tap): low word of the weight descriptor
  __shared__ int sGrp[4 * TM_MAX_TAPS];                     // per (phase, group): map | dw+128 << 8 | dh+128 << 16
  __shared__ int4 sTile[TM_TILE_TABLE];                     // per local tile: phase, image, th, tw
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail);
  uint64_t* full = bars;                                    // [8]
  uint64_t* empty = bars + TM_MAX_STAGES;                   // [8]
  uint64_t* wfull = bars + 2 * TM_MAX_STAGES;               // [1]
  uint64_t* acc_full = wfull + 1;                           // [8]
  uint64_t* acc_empty = acc_full + 8;                       // [8]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 8);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stages = P.stages;
  if (tid == 0) tm_stamp(P.dbg, 0);

  // Everything up to the first pdl_wait() below is independent of the preceding kernel (common.cuh: PDL): shared-memory
  // tables, barriers, TMEM, and the layer's weight images, which were packed at the start of the step.
  const int ntiles_mine = ((int)blockIdx.x < P.total_tiles)
                              ? (P.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  if (tid >= 64 && tid < 128) {
    const int e = tid - 64, p = e >> 4, t = e & 15;
    if (p < P.nphases) {
      const TmaPhase& ph = P.ph[p];
      sTapA[e] = (uint32_t)ph.aoff[t];
      sTapB[e] = (((smem_u32(sW) + (uint32_t)ph.wslot[t] * TM_W_BYTES) >> 4) & 0x3FFFu) | (1u << 16);
      const int gcount = t < ph.ngroups ? (int)ph.gfirst[t + 1] - (int)ph.gfirst[t] : 0;
      sGrp[e] = (int)ph.gmap[t] | (((int)ph.gdw[t] + 128) << 8) | (((int)ph.gdh[t] + 128) << 16) |
                ((int)ph.gfirst[t] << 24) | ((gcount > 0 ? gcount - 1 : 0) << 28);       // (count - 1: up to 16 taps)
    }
  }
  if (tid >= 128 && tid < 128 + TM_TILE_TABLE) {
    const int i = tid - 128;
    if (i < ntiles_mine) {
      const TmaTile tl = decode_tile(P, (int)blockIdx.x + i * (int)gridDim.x);
      sTile[i] = make_int4(tl.p, tl.n, tl.th, tl.tw);
    }
  }
  if (tid == 0) {
    for (int s = 0; s < TM_MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(wfull, 1);
    for (int a = 0; a < 8; ++a) mbar_init(&acc_full[a], 1);
    for (int a = 0; a < 8; ++a) mbar_init(&acc_empty[a], 256);
    fence_mbar_init();
    tma::prefetch_desc(&tm0);
    if (P.in_stride == 2) {
      tma::prefetch_desc(&tm1);
      tma::prefetch_desc(&tm2);
      tma::prefetch_desc(&tm3);
    }
  }
  if (warp == 1) tmem_alloc<NACC * TM_ACC_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) tm_stamp(P.dbg, 1);

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && ntiles_mine > 0) {
      if (!P.w_early) pdl_wait();
      mbar_arrive_expect_tx(wfull, (uint32_t)P.nw * TM_W_BYTES);
      for (int t = 0; t < P.nw; ++t)
        bulk_g2s(sW + (size_t)t * TM_W_BYTES, wimg + (size_t)t * TM_W_BYTES, TM_W_BYTES, wfull);
      pdl_wait();                              // the activations are the preceding kernel's output
      pdl_trigger();
      // the ring of activation slots: ONE ring for a single MMA issuer; with two, each issuer owns a part of it (its
      // full / empty barriers then have exactly one waiter on either side -- a thread that skipped the other issuer's
      // phases of a shared barrier could not tell phase u from phase u - 2 by parity)
      const bool split = P.issuers == 2;
      const int size0 = split ? stages >> 1 : stages, size1 = stages - (stages >> 1);
      int s0 = 0, s1 = 0;
      uint32_t ep0 = 1, ep1 = 1;               // first pass over a ring: slots are free (wait on the preceding phase)
      const uint32_t slot_bytes = P.slot_bytes;
      const uint32_t sA_u32 = smem_u32(sA);
      for (int i = 0; i < ntiles_mine; ++i) {
        const TmaTile tl = fetch_tile(P, sTile, i);
        const int ngroups = P.ph[tl.p].ngroups;
        const int w0 = tl.tw * P.box_w, h0 = tl.th * P.box_h;
        const bool second = split && (i & 1);
        for (int g = 0; g < ngroups; ++g) {
          const int s = second ? size0 + s1 : s0;
          mbar_wait(&empty[s], second ? ep1 : ep0);
          mbar_arrive_expect_tx(&full[s], P.box_bytes);
          const int gi = sGrp[tl.p * TM_MAX_TAPS + g];
          const int par = gi & 3;
          const CUtensorMap* tm = par == 0 ? &tm0 : (par == 1 ? &tm1 : (par == 2 ? &tm2 : &tm3));
          tma::load_4d_u32(sA_u32 + (uint32_t)s * slot_bytes, tm, 0, w0 + ((gi >> 8) & 255) - 128,
                           h0 + ((gi >> 16) & 255) - 128, tl.n, &full[s]);
          if (second) {
            if (++s1 == size1) {
              s1 = 0;
              ep1 ^= 1;
            }
          } else if (++s0 == size0) {
            s0 = 0;
            ep0 ^= 1;
          }
        }
        tm_stamp(P.dbg, 2 + (i < 3 ? i : 3));          // 2..5: producer finished issuing tile i
      }
    }
  } else if (warp == 1 || warp == 10) {
    // ------------------------------------------------------------------ MMA issuers (conv_issue_tiles above)
    if (P.elect)
      conv_issue_tiles<NACC, true>(P, warp == 1 ? 0 : 1, lane, ntiles_mine, sTile, sGrp, sTapA, sTapB, sA, full, empty, wfull,
                                   acc_full, acc_empty, tmem_base);
    else
      conv_issue_tiles<NACC, false>(P, warp == 1 ? 0 : 1, lane, ntiles_mine, sTile, sGrp, sTapA, sTapB, sA, full, empty, wfull,
                                    acc_full, acc_empty, tmem_base);
  } else {
    // ------------------------------------------------------------------ epilogue: all 8 warps work on every tile;
    // thread = (output row, 32-channel half).  The per-channel statistics are carried in registers across the CTA's
    // tiles and leave through ONE 32 x 32 shuffle transpose per statistic at the end (they used to cost two
    // transposes per tile and half, 4000-4700 cycles against 1500 for the plain epilogue).
    const int half = (warp - 2) >> 2;                // channel half
    const int q = warp & 3;                          // TMEM lane quarter this warp may read
    const int r = q * 32 + lane;
    const int ch0 = half * 32;
    const int bw = r & (P.box_w - 1), bh = r >> P.log2_bw;
    const int e = tid - 64;
    pdl_wait();                                      // epilogue operands / coefficients come from preceding kernels
    pdl_trigger();
    if (e < TM_C) {
      const int c = e;
      const float bias = epi.bias ? __ldg(epi.bias + c) : 0.f;
      float osc = 1.f, osh = bias;
      if (bn.mode == NSG_BNF_EVAL) {                 // BatchNorm in eval mode folds into the output affine
        const float is = 1.0f / sqrtf(__ldg(bn.running_var + c) + bn.eps);
        osc = __ldg(bn.gamma + c) * is;
        osh = fmaf(bias - __ldg(bn.running_mean + c), osc, __ldg(bn.beta + c));
      }
      sPar[0 * TM_C + c] = osh;
      sPar[1 * TM_C + c] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mscale + c) : 1.f;
      sPar[2 * TM_C + c] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mshift + c) : 0.f;
      sPar[3 * TM_C + c] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.mean + c) : 0.f;
      sPar[4 * TM_C + c] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.invstd + c) : 0.f;
      sPar[5 * TM_C + c] = osc;
    }
    named_bar(2, 256);
    const bool do_stats = epi.stats_mode != NSG_STATS_NONE;
    const bool bwd_stats = epi.stats_mode == NSG_STATS_BNBWD;
    float s1[32], s2[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) s1[j] = s2[j] = 0.f;
    auto row_offset = [&](int i, bool& valid) -> long long {
      const TmaTile tl = fetch_tile(P, sTile, i);
      const TmaPhase& ph = P.ph[tl.p];
      // (tile grid "w" / "h" = the box's fast / slow pixel axis: W / H, or H / W when P.swap)
      const int fi = tl.tw * P.box_w + bw, si = tl.th * P.box_h + bh;
      const int ohs = P.swap ? fi : si, ows = P.swap ? si : fi;
      valid = ohs < ph.sub_h && ows < ph.sub_w;
      return valid ? ((((long long)tl.n * P.full_h + (ohs * ph.out_stride + ph.off_h)) * P.full_w +
                       (ows * ph.out_stride + ph.off_w)) * TM_C + ch0) : 0;
    };
    for (int i = 0; i < ntiles_mine; ++i) {
      bool valid;
      const long long oh = row_offset(i, valid);
      const bool use_add = valid && epi.add_src != nullptr;
      uint4 radd[4];
      if (use_add) ld_half(epi.add_src, oh, radd);       // requested before the accumulator wait
      const int as = i % NACC;
      mbar_wait_backoff(&acc_full[as], (uint32_t)((i / NACC) & 1));
      if (tid == 64) tm_stamp(P.dbg, 12 + (i < 3 ? i : 3));     // 12..15: accumulator of tile i ready
      tc_fence_after();
      uint32_t acc[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)as * TM_ACC_COLS + (uint32_t)ch0, acc);
      tmem_ld_wait();
      if (i + NACC < ntiles_mine) {                    // a later tile of this CTA takes over the accumulator
        tc_fence_before();
        mbar_arrive(&acc_empty[as]);
      }
      float x[32];
      epi_half(x, acc, epi, sPar, oh, valid, ch0, use_add, radd, bn.mode == NSG_BNF_EVAL);
      if (valid && out != nullptr) {
        st16bf(out + oh, &x[0]);
        st16bf(out + oh + 16, &x[16]);
      }
      if (do_stats) {
        if (bwd_stats) {
          if (valid) {
            uint4 rx[4];
            ld_half(epi.xhat_src, oh, rx);
            float mu[32], is[32];
            ld32s(sPar + 3 * TM_C + ch0, mu);
            ld32s(sPar + 4 * TM_C + ch0, is);
#pragma unroll
            for (int g8 = 0; g8 < 4; ++g8) {
              float hx[8];
              unpack8bf(rx[g8], hx);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                s1[g8 * 8 + j] += x[g8 * 8 + j];
                s2[g8 * 8 + j] = fmaf(x[g8 * 8 + j], (hx[j] - mu[g8 * 8 + j]) * is[g8 * 8 + j], s2[g8 * 8 + j]);
              }
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            s1[j] += x[j];
            s2[j] = fmaf(x[j], x[j], s2[j]);
          }
        }
      }
      if (tid == 64) tm_stamp(P.dbg, 16 + (i < 3 ? i : 3));     // 16..19: epilogue of tile i done
    }
    if (do_stats) {
      const float t1 = transpose_sum32(s1, lane);
      const float t2 = transpose_sum32(s2, lane);
      // one slot per (TMEM lane quarter, statistic, channel), summed in a FIXED order: the CTA's contribution does not
      // depend on which warp finishes first, so the forward pass is reproducible run to run (the fp64 atomics across
      // CTAs differ at 1e-16, far below the fp32 coefficients derived from them)
      sStat[q * 2 * TM_C + ch0 + lane] = t1;
      sStat[q * 2 * TM_C + TM_C + ch0 + lane] = t2;
      named_bar(1, 256);
      if (e < 2 * TM_C)
        atomicAdd(epi.stats + e, (double)((sStat[e] + sStat[2 * TM_C + e]) + (sStat[4 * TM_C + e] + sStat[6 * TM_C + e])));
    }
    if (HOLD) {
      // ---------------------------------------------------------------- grid barrier, coefficients, pass 2
      named_bar(1, 256);                               // every statistic atomic of this CTA has been issued
      if (bn.peer.world > 1) {
        // synchronised BatchNorm: CTA 0 waits for the grid, swaps the sums with the other GPUs, then opens a second gate
        if (tid == 64) grid_barrier(bn.barrier, gridDim.x);
        if (blockIdx.x == 0) {
          named_bar(1, 256);
          peer_exchange(bn.peer, epi.stats, e, 1, 256);
          named_bar(1, 256);
          if (tid == 64) {
            __threadfence();
            atomicExch(bn.barrier + 1, 1u);
          }
        } else if (tid == 64) {
          grid_barrier_wait(bn.barrier + 1, 1u, 30 * kHangNs);    // the other GPUs may be seconds away
        }
      } else if (tid == 64) {
        grid_barrier(bn.barrier, gridDim.x);
      }
      named_bar(1, 256);
      if (e < TM_C) {
        const int c = e;
        const double a0 = __ldcg(epi.stats + c), a1 = __ldcg(epi.stats + TM_C + c);
        if (bn.mode == NSG_BNF_TRAIN) {
          const double m = a0 / bn.count;
          double var = a1 / bn.count - m * m;
          if (var < 0.0) var = 0.0;
          const float mu = (float)m;
          const float is = (float)(1.0 / sqrt(var + (double)bn.eps));
          const float sc = __ldg(bn.gamma + c) * is;
          const float sh = __ldg(bn.beta + c) - mu * sc;
          sPar[6 * TM_C + c] = sc;
          sPar[7 * TM_C + c] = sh;
          if (blockIdx.x == 0) {
            if (bn.running_mean) {
              const double unbiased = bn.count > 1.0 ? var * bn.count / (bn.count - 1.0) : var;
              bn.running_mean[c] = (1.f - bn.momentum) * bn.running_mean[c] + bn.momentum * mu;
              bn.running_var[c] = (1.f - bn.momentum) * bn.running_var[c] + bn.momentum * (float)unbiased;
              if (c == 0 && bn.nbt) *bn.nbt += 1;
            }
            if (bn.coef) {
              bn.coef[0 * TM_C + c] = mu;
              bn.coef[1 * TM_C + c] = is;
              bn.coef[2 * TM_C + c] = sc;
              bn.coef[3 * TM_C + c] = sh;
            }
          }
        } else {                                       // NSG_BNF_BWD: dh = A*g + B*h + C (fused_elem.cu: bn_bwd_kernel)
          const double sg = a0 / bn.count, sgx = a1 / bn.count;
          const float mu = sPar[3 * TM_C + c], is = sPar[4 * TM_C + c];
          const float k0 = __ldg(bn.gamma + c) * is;
          const float kb = k0 * is * (float)sgx;
          sPar[6 * TM_C + c] = k0;
          sPar[7 * TM_C + c] = -kb;
          sPar[8 * TM_C + c] = kb * mu - k0 * (float)sg;
          if (blockIdx.x == 0) {
            if (bn.dbeta) bn.dbeta[c] += (float)a0;
            if (bn.dgamma) bn.dgamma[c] += (float)a1;
          }
        }
      }
      named_bar(1, 256);
      for (int i = 0; i < ntiles_mine; ++i) {
        bool valid;
        const long long oh = row_offset(i, valid);
        // tiles whose accumulator was handed on read what pass 1 stored (h forward, g backward: written by this thread)
        const bool in_tmem = i + NACC >= ntiles_mine;
        // tcgen05.ld / tcgen05.wait::ld are .sync.aligned: the WHOLE warp executes them, converged -- rows outside a
        // ragged tile (box wider than the tensor) take part and only skip the arithmetic and the store.  The row's
        // global operands are requested in between so that they travel under the TMEM read.
        uint32_t acc[32];
        if (in_tmem) tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(i % NACC) * TM_ACC_COLS + (uint32_t)ch0, acc);
        const bool train = bn.mode == NSG_BNF_TRAIN;
        const bool use_res = valid && train && bn.res != nullptr;
        const bool use_add = valid && !train && in_tmem && epi.add_src != nullptr;
        uint4 rs[4], rr[4], radd[4], rx[4];
        if (valid && !in_tmem) {
          const uint4* p = reinterpret_cast<const uint4*>(out + oh);
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) rs[g8] = p[g8];
        }
        if (use_res) ld_half(bn.res, oh, rr);
        if (use_add) ld_half(epi.add_src, oh, radd);
        if (valid && !train) ld_half(epi.xhat_src, oh, rx);
        if (in_tmem) tmem_ld_wait();
        if (!valid) continue;
        float x[32];
        if (!in_tmem) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = 0.f;
          add_half(x, rs);
        }
        if (train) {
          if (in_tmem) {
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = __uint_as_float(acc[j]) + sPar[ch0 + j];       // (TRAIN: output scale 1)
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = fmaf(x[j], sPar[6 * TM_C + ch0 + j], sPar[7 * TM_C + ch0 + j]);
          if (use_res) add_half(x, rr);
          if (bn.relu_out) {
#pragma unroll
            for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
          }
        } else {
          if (in_tmem) epi_half(x, acc, epi, sPar, oh, true, ch0, use_add, radd, false);     // g, exactly as pass 1 formed it
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {
            float hx[8];
            unpack8bf(rx[g8], hx);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int ch = ch0 + g8 * 8 + j;
              x[g8 * 8 + j] = fmaf(sPar[6 * TM_C + ch], x[g8 * 8 + j], fmaf(sPar[7 * TM_C + ch], hx[j], sPar[8 * TM_C + ch]));
            }
          }
        }
        st16bf(bn.out2 + oh, &x[0]);
        st16bf(bn.out2 + oh + 16, &x[16]);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<NACC * TM_ACC_COLS>(tmem_base);
  if (tid == 0) tm_stamp(P.dbg, 20);
}

// ---------------------------------------------------------------------------------------------- conv_tma host side
static unsigned long long* g_timeline = nullptr;       // device buffer of 32 stamps, set by nsg_debug_timeline(1)
int debug_timeline(int enable, unsigned long long* host_out) {
  if (enable && !g_timeline) {
    cudaError_t e = cudaMalloc(&g_timeline, 32 * sizeof(unsigned long long));
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc(timeline)");
    cudaMemset(g_timeline, 0, 32 * sizeof(unsigned long long));
  }
  if (host_out && g_timeline) {
    cudaError_t e = cudaMemcpy(host_out, g_timeline, 32 * sizeof(unsigned long long), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return cuda_fail(e, "cudaMemcpy(timeline)");
  }
  if (!enable && g_timeline) {
    cudaFree(g_timeline);
    g_timeline = nullptr;
  }
  return NSG_OK;
}
static int tma_shape_ok(const nsg_conv_geom* cg, const char* what) {
  if (cg->cin != TM_C || cg->cout != TM_C || cg->dtype != NSG_BF16) {
    set_error("%s: the TMA kernels serve bf16 tensors with cin == cout == 64 (got %d -> %d, dtype %d)", what, cg->cin,
              cg->cout, cg->dtype);
    return NSG_E_UNSUPPORTED;
  }
  if (cg->ksize < 1 || cg->ksize > 4 || cg->stride < 1 || cg->stride > 2) {
    set_error("%s: ksize %d / stride %d unsupported (1..4 / 1..2)", what, cg->ksize, cg->stride);
    return NSG_E_UNSUPPORTED;
  }
  return NSG_OK;
}

static void pick_box_w(int sub_w, int pixels, int* box_w, int* box_h, int* log2_bw) {
  int bw = 8, l2 = 3;
  while (bw < pixels && bw < sub_w) {
    bw <<= 1;
    ++l2;
  }
  *box_w = bw;
  *box_h = pixels / bw;
  *log2_bw = l2;
}

// group the taps of one phase by (tensor map, horizontal shift); returns the number of groups and the largest
// vertical span - 1 (= halo rows).  in_stride 2 addresses the four parity views of the tensor.
struct TapGroups {
  int ngroups, halo;
  int gmap[TM_MAX_TAPS], gdw[TM_MAX_TAPS], gdh[TM_MAX_TAPS], gcount[TM_MAX_TAPS];
  int tap[TM_MAX_TAPS][TM_MAX_TAPS];       // tap indices (into TapGeom) per group
  int edh[TM_MAX_TAPS];                    // effective vertical shift per tap
};
static void group_taps(const TapGeom& g, TapGroups* G) {
  G->ngroups = 0;
  G->halo = 0;
  for (int t = 0; t < g.ntaps; ++t) {
    int map = 0, dw = g.dw[t], dh = g.dh[t];
    if (g.in_stride == 2) {
      map = (dh & 1) * 2 + (dw & 1);
      dw >>= 1;
      dh >>= 1;
    }
    G->edh[t] = dh;
    int gi = -1;
    for (int i = 0; i < G->ngroups; ++i)
      if (G->gmap[i] == map && G->gdw[i] == dw) gi = i;
    if (gi < 0) {
      gi = G->ngroups++;
      G->gmap[gi] = map; G->gdw[gi] = dw; G->gdh[gi] = dh; G->gcount[gi] = 0;
    }
    if (dh < G->gdh[gi]) G->gdh[gi] = dh;
    G->tap[gi][G->gcount[gi]++] = t;
  }
  for (int i = 0; i < G->ngroups; ++i)
    for (int j = 0; j < G->gcount[i]; ++j) {
      const int span = G->edh[G->tap[i][j]] - G->gdh[i];
      if (span > G->halo) G->halo = span;
    }
}

// 128-pixel tile shape (box_w x box_h, box_w a power of two >= 8).  Measured on B200 with every SM loading
// (profiles/r1_tma_v9_conv_timeline.txt): boxes of one CTA land one after the other at 22-30 bytes per cycle whatever
// their shape (16 KB 64x2, 24 KB 32x6 and 32 KB 64x4 boxes all fit that line), so the cost to minimise is the bytes
// the TMA unit moves, tiles * groups * (box_h + halo) * box_w -- ragged tiles and halo rows together; ties go to the
// shape with fewer box rows.
static void pick_box(const TapGeom* phases, const TapGroups* groups, int np, int batch, int pixels, int halo,
                     int* box_w, int* box_h, int* log2_bw) {
  long long best_rows = -1, best_bytes = -1;
  for (int l2 = 3; (1 << l2) <= pixels; ++l2) {
    const int bw = 1 << l2, bh = pixels / bw;
    if (bh + halo > 256) continue;                       // TMA box dimension limit
    long long rows = 0, bytes = 0;
    for (int i = 0; i < np; ++i) {
      if (phases[i].sub_h <= 0 || phases[i].sub_w <= 0) continue;
      const long long tiles = (long long)((phases[i].sub_w + bw - 1) / bw) * ((phases[i].sub_h + bh - 1) / bh) * batch;
      rows += tiles * groups[i].ngroups * (bh + halo);
      bytes += tiles * groups[i].ngroups * (bh + halo) * bw;
    }
    if (best_rows < 0 || bytes < best_bytes || (bytes == best_bytes && rows <= best_rows)) {
      best_rows = rows;
      best_bytes = bytes;
      *box_w = bw; *box_h = bh; *log2_bw = l2;
    }
  }
}

// tensor maps over `in` (N, in_h, in_w, 64) bf16: plain view (in_stride 1) or the four parity views (in_stride 2)
static int encode_input_maps(CUtensorMap* tm, const void* in, int batch, int in_h, int in_w, int in_stride, int box_w,
                             int box_h, int swap = 0) {
  // swap: the box's fast pixel axis (tensor-map dimension 1) is H instead of W
  const long long C = TM_C;
  auto enc = [&](CUtensorMap* t, const void* base, long long dw_, long long dh_, long long ws, long long hs) -> int {
    if (swap)
      return tma::encode_nhwc64(t, base, dh_, dw_, batch, hs, ws, (long long)in_h * in_w * C, box_w, box_h);
    return tma::encode_nhwc64(t, base, dw_, dh_, batch, ws, hs, (long long)in_h * in_w * C, box_w, box_h);
  };
  if (in_stride == 1) {
    if (int rc = enc(&tm[0], in, in_w, in_h, C, (long long)in_w * C)) return rc;
    tm[1] = tm[2] = tm[3] = tm[0];
    return NSG_OK;
  }
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(in) + ((long long)ph * in_w + pw) * C;
      const long long dw_ = (in_w - pw + 1) / 2, dh_ = (in_h - ph + 1) / 2;
      if (dw_ < 1 || dh_ < 1) {       // degenerate parity view (1-pixel-wide tensor): never addressed in range
        if (int rc = enc(&tm[ph * 2 + pw], in, 1, 1, 2 * C, 2LL * in_w * C)) return rc;
        continue;
      }
      if (int rc = enc(&tm[ph * 2 + pw], base, dw_, dh_, 2 * C, 2LL * in_w * C)) return rc;
    }
  return NSG_OK;
}

int make_peer(const nsg_peer_sync* p, PeerSync* out) {
  NSG_REQUIRE(p->world >= 2 && p->world <= NSG_PEER_MAX && p->rank >= 0 && p->rank < p->world,
              "peer sync: world %d / rank %d (at most %d GPUs of one node)", p->world, p->rank, NSG_PEER_MAX);
  NSG_REQUIRE(p->slot >= 0 && p->slot < NSG_PEER_SLOTS, "peer sync: slot %d out of range", p->slot);
  PeerSync s{};
  s.world = p->world; s.rank = p->rank; s.slot = p->slot;
  for (int i = 0; i < p->world; ++i) {
    NSG_REQUIRE(p->areas[i], "peer sync: area of rank %d missing", i);
    s.slots[i] = static_cast<double*>(p->areas[i]);
    s.flags[i] = reinterpret_cast<unsigned int*>(static_cast<char*>(p->areas[i]) +
                                                 (size_t)NSG_PEER_SLOTS * kPeerSlotDoubles * sizeof(double));
  }
  *out = s;
  return NSG_OK;
}

int peer_allreduce_stats(double* sums, const nsg_peer_sync* p, cudaStream_t stream) {
  NSG_REQUIRE(sums && p, "peer_allreduce_stats: null pointer");
  PeerSync s{};
  if (int rc = make_peer(p, &s)) return rc;
  launch_pdl(peer_allreduce_kernel, dim3(1), dim3(128), 0, stream, sums, s);
  NSG_CHECK_LAUNCH("peer_allreduce_kernel");
  return NSG_OK;
}

// validated device-side copy of the public nsg_bn_fuse
static int make_bnfuse(const nsg_bn_fuse* b, const EpiArgs& e, BnFuse* out) {
  BnFuse f{};
  if (b && b->mode != NSG_BNF_NONE) {
    NSG_REQUIRE(b->mode == NSG_BNF_TRAIN || b->mode == NSG_BNF_EVAL || b->mode == NSG_BNF_BWD, "conv_bn_tma: mode %d", b->mode);
    f.mode = b->mode; f.relu_out = b->relu_out; f.gamma = b->gamma; f.beta = b->beta; f.eps = b->eps;
    f.momentum = b->momentum; f.count = b->count; f.running_mean = b->running_mean; f.running_var = b->running_var;
    f.nbt = b->num_batches_tracked; f.coef = b->coef; f.res = static_cast<const __nv_bfloat16*>(b->res);
    f.barrier = b->barrier; f.dgamma = b->dgamma; f.dbeta = b->dbeta; f.out2 = static_cast<__nv_bfloat16*>(b->out2);
    if (b->peer && b->peer->world > 1) {
      if (int rc = make_peer(b->peer, &f.peer)) return rc;
      NSG_REQUIRE(f.mode != NSG_BNF_EVAL, "conv_bn_tma: eval mode exchanges nothing");
    } else {
      f.peer.world = 1;
    }
    NSG_REQUIRE(f.gamma, "conv_bn_tma: gamma missing");
    if (f.mode == NSG_BNF_EVAL) {
      NSG_REQUIRE(f.beta && f.running_mean && f.running_var, "conv_bn_tma: eval mode needs beta and the running statistics");
      NSG_REQUIRE(e.stats_mode == NSG_STATS_NONE, "conv_bn_tma: eval mode accumulates no statistics");
    } else {
      NSG_REQUIRE(f.barrier && f.out2 && f.count > 0, "conv_bn_tma: barrier counter, second output and count are required");
      NSG_REQUIRE((reinterpret_cast<uintptr_t>(f.out2) & 31) == 0 && (reinterpret_cast<uintptr_t>(f.res) & 15) == 0,
                  "conv_bn_tma: out2 must be 32-byte and res 16-byte aligned");
      if (f.mode == NSG_BNF_TRAIN)
        NSG_REQUIRE(f.beta && e.stats_mode == NSG_STATS_MOMENTS, "conv_bn_tma: train mode needs beta and NSG_STATS_MOMENTS");
      else
        NSG_REQUIRE(e.stats_mode == NSG_STATS_BNBWD, "conv_bn_tma: backward mode needs NSG_STATS_BNBWD operands");
    }
  }
  *out = f;
  return NSG_OK;
}

template <int NACC, bool HOLD>
static int launch_conv_tma(int grid, size_t smem, cudaStream_t stream, const CUtensorMap* tm, const void* packed, void* out,
                           const EpiArgs& e, const BnFuse& f, const TmaConvParams& P) {
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(conv_tma_kernel<NACC, HOLD>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)TM_SMEM_LIMIT);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(conv_tma_kernel)");
    configured = true;
  }
  launch_pdl_ex(HOLD, conv_tma_kernel<NACC, HOLD>, dim3(grid), dim3(TM_THREADS), smem, stream, tm[0], tm[1], tm[2], tm[3],
                static_cast<const uint8_t*>(packed), static_cast<__nv_bfloat16*>(out), e, f, P);
  NSG_CHECK_LAUNCH("conv_tma_kernel");
  return NSG_OK;
}

// tiles per CTA of the persistent grid for this geometry (what decides whether a hold launch fits in TMEM)
static int conv_tma_tiles_per_cta(const nsg_conv_geom* cg) {
  TapGeom phases[4];
  const int np = build_phases(cg, phases);
  static TapGroups groups[4];
  int halo = 0;
  for (int i = 0; i < np; ++i) {
    group_taps(phases[i], &groups[i]);
    if (phases[i].sub_h > 0 && phases[i].sub_w > 0 && groups[i].halo > halo) halo = groups[i].halo;
  }
  int bw, bh, l2;
  pick_box(phases, groups, np, cg->batch, TM_M, halo, &bw, &bh, &l2);
  long long tiles = 0;
  for (int i = 0; i < np; ++i) {
    if (phases[i].sub_h <= 0 || phases[i].sub_w <= 0) continue;
    tiles += (long long)((phases[i].sub_w + bw - 1) / bw) * ((phases[i].sub_h + bh - 1) / bh) * cg->batch;
  }
  if (tiles == 0) return 0;
  const long long grid = tiles < num_sms() ? tiles : num_sms();
  return (int)((tiles + grid - 1) / grid);
}
int conv_bn_tma_fits(const nsg_conv_geom* cg) {
  if (cg->cin != TM_C || cg->cout != TM_C || cg->dtype != NSG_BF16 || cg->ksize < 1 || cg->ksize > 4 || cg->stride < 1 ||
      cg->stride > 2)
    return 0;
  return conv_tma_tiles_per_cta(cg) <= 64 ? 1 : 0;
}

int conv_gemm_tma(const nsg_conv_geom* cg, const void* in, const void* packed, const nsg_epilogue* epi,
                  const nsg_bn_fuse* bnf, void* out, cudaStream_t stream) {
  if (int rc = tma_shape_ok(cg, "conv_gemm_tma")) return rc;
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 127) == 0, "conv_gemm_tma: packed weights must be 128-byte aligned");
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 31) == 0,
              "conv_gemm_tma: the input must be 16-byte and the output 32-byte aligned");
  EpiArgs e{};
  if (int rc = make_epi(epi, &e)) return rc;
  NSG_REQUIRE(e.tanh_out == 0 || e.tanh_out == 2,
              "conv_gemm_tma: output activation must be none (0) or ReLU (2); tanh belongs to the cout == 1 layer");
  BnFuse f{};
  if (int rc = make_bnfuse(bnf, e, &f)) return rc;
  const bool hold = f.mode == NSG_BNF_TRAIN || f.mode == NSG_BNF_BWD;
  NSG_REQUIRE(out || hold, "conv_gemm_tma: null output");
  TapGeom phases[4];
  const int np = build_phases(cg, phases);
  TmaConvParams P{};
  P.batch = cg->batch; P.full_h = cg->out_h; P.full_w = cg->out_w;
  P.in_stride = phases[0].in_stride;
  NSG_REQUIRE(P.in_stride == 1 || P.in_stride == 2, "conv_gemm_tma: input stride %d", P.in_stride);
  static TapGroups groups[4];
  int halo = 0;
  for (int i = 0; i < np; ++i) {
    group_taps(phases[i], &groups[i]);
    if (phases[i].sub_h > 0 && phases[i].sub_w > 0 && groups[i].halo > halo) halo = groups[i].halo;
  }
  // ONE box per (tile, tensor map) for kernels wider than one pixel.  The tile is 8 pixels along the box's fast pixel
  // axis x 16 lines; the box carries the halo pixels of BOTH axes, and tap (dh, dw) reads the sub-block that starts
  // (its line shift x line pitch + its pixel shift) x 128 bytes into the box: a group of 8 A rows is one box line, the
  // descriptor's group stride (SBO) is the line pitch, and a start 128 bytes off the 1024-byte swizzle atom is legal
  // because the tensor core swizzles on the absolute shared-memory address, as the TMA unit does when it writes
  // (tools/microbench/desc_probe.cu: every tap of a 3 x 3 kernel from one 10 x 18-pixel box).  A 3 x 3 tile then loads
  // 23 KB instead of three 24 KB boxes -- the L2 -> shared-memory path is what bounds these kernels.  The fast axis is
  // whichever of W / H leaves fewer ragged tiles (20 x 64 latent grid: H, 3 x 4 tiles per image).
  static const bool one_box_enabled = [] { const char* v = getenv("NSG_CONV_ONE_BOX"); return !(v && v[0] == '0'); }();
  bool one_box = one_box_enabled && cg->ksize > 1;
  int hf[2] = {0, 0}, hs[2] = {0, 0};                   // halo along the fast / slow axis, per orientation
  long long ntile[2] = {0, 0};
  if (one_box) {
    for (int sw = 0; sw < 2; ++sw)
      for (int i = 0; i < np; ++i) {
        const TapGeom& g = phases[i];
        if (g.sub_h <= 0 || g.sub_w <= 0) continue;
        const int ef = sw ? g.sub_h : g.sub_w, es = sw ? g.sub_w : g.sub_h;
        ntile[sw] += (long long)((ef + 7) / 8) * ((es + 15) / 16) * cg->batch;
        for (int m = 0; m < 4; ++m) {                   // per tensor map: the span of its taps on either axis
          int lo_f = 1 << 20, hi_f = -(1 << 20), lo_s = 1 << 20, hi_s = -(1 << 20);
          for (int t = 0; t < g.ntaps; ++t) {
            int map = 0, dw = g.dw[t], dh = g.dh[t];
            if (g.in_stride == 2) {
              map = (dh & 1) * 2 + (dw & 1);
              dw >>= 1;
              dh >>= 1;
            }
            if (map != m) continue;
            const int df = sw ? dh : dw, ds = sw ? dw : dh;
            if (df < lo_f) lo_f = df;
            if (df > hi_f) hi_f = df;
            if (ds < lo_s) lo_s = ds;
            if (ds > hi_s) hi_s = ds;
          }
          if (hi_f >= lo_f && hi_f - lo_f > hf[sw]) hf[sw] = hi_f - lo_f;
          if (hi_s >= lo_s && hi_s - lo_s > hs[sw]) hs[sw] = hi_s - lo_s;
        }
      }
  }
  P.swap = 0;
  if (one_box) {
    P.swap = ntile[1] < ntile[0] ? 1 : 0;
    P.box_w = 8; P.box_h = 16; P.log2_bw = 3;
    P.halo = hs[P.swap];
    P.row_bytes = (uint32_t)(8 + hf[P.swap]) * 128;                       // one box line
    P.box_bytes = (uint32_t)(16 + hs[P.swap]) * P.row_bytes;
    P.slot_bytes = (P.box_bytes + 1023u) & ~1023u;
    P.sbo16 = P.row_bytes >> 4;
  } else {
    pick_box(phases, groups, np, cg->batch, TM_M, halo, &P.box_w, &P.box_h, &P.log2_bw);
    P.halo = halo;
    P.row_bytes = (uint32_t)P.box_w * 128;
    P.slot_bytes = (uint32_t)(P.box_h + halo) * P.row_bytes;
    P.box_bytes = P.slot_bytes;
    P.sbo16 = 1024 >> 4;
  }
  P.nw = cg->ksize * cg->ksize;
  int tiles = 0, nph = 0;
  for (int i = 0; i < np; ++i) {
    const TapGeom& g = phases[i];
    if (g.sub_h <= 0 || g.sub_w <= 0 || cg->batch == 0) continue;
    NSG_REQUIRE(g.ntaps > 0, "conv_gemm_tma: a stride phase without taps (ksize < stride) is not supported");
    TmaPhase& ph = P.ph[nph++];
    ph.sub_h = g.sub_h; ph.sub_w = g.sub_w;
    const int ef = P.swap ? g.sub_h : g.sub_w, es = P.swap ? g.sub_w : g.sub_h;
    ph.tiles_w = (ef + P.box_w - 1) / P.box_w;
    ph.tiles_h = (es + P.box_h - 1) / P.box_h;
    ph.tile0 = tiles;
    tiles += ph.tiles_w * ph.tiles_h * cg->batch;
    ph.ntaps = g.ntaps;
    ph.out_stride = g.out_stride; ph.off_h = g.off_h; ph.off_w = g.off_w;
    int nt = 0;
    if (one_box) {
      // groups = the tensor maps that have taps; box origin = the least shift on either axis
      int ng = 0;
      for (int m = 0; m < 4; ++m) {
        int lo_f = 1 << 20, lo_s = 1 << 20, cnt = 0;
        int df[TM_MAX_TAPS], ds[TM_MAX_TAPS], ti[TM_MAX_TAPS];
        for (int t = 0; t < g.ntaps; ++t) {
          int map = 0, dw = g.dw[t], dh = g.dh[t];
          if (g.in_stride == 2) {
            map = (dh & 1) * 2 + (dw & 1);
            dw >>= 1;
            dh >>= 1;
          }
          if (map != m) continue;
          df[cnt] = P.swap ? dh : dw;
          ds[cnt] = P.swap ? dw : dh;
          ti[cnt] = t;
          if (df[cnt] < lo_f) lo_f = df[cnt];
          if (ds[cnt] < lo_s) lo_s = ds[cnt];
          ++cnt;
        }
        if (cnt == 0) continue;
        ph.gmap[ng] = (unsigned char)m;
        ph.gdw[ng] = (signed char)lo_f;
        ph.gdh[ng] = (signed char)lo_s;
        ph.gfirst[ng] = (unsigned char)nt;
        for (int j = 0; j < cnt; ++j, ++nt) {
          ph.aoff[nt] = (unsigned short)((((uint32_t)(ds[j] - lo_s) * P.row_bytes) + (uint32_t)(df[j] - lo_f) * 128u) >> 4);
          ph.wslot[nt] = (unsigned char)g.woff[ti[j]];
        }
        ++ng;
      }
      ph.ngroups = ng;
      ph.gfirst[ng] = (unsigned char)nt;
    } else {
      const TapGroups& tg = groups[i];
      ph.ngroups = tg.ngroups;
      for (int gi = 0; gi < tg.ngroups; ++gi) {
        ph.gmap[gi] = (unsigned char)tg.gmap[gi];
        ph.gdw[gi] = (signed char)tg.gdw[gi];
        ph.gdh[gi] = (signed char)tg.gdh[gi];
        ph.gfirst[gi] = (unsigned char)nt;
        for (int j = 0; j < tg.gcount[gi]; ++j, ++nt) {
          const int t = tg.tap[gi][j];
          ph.aoff[nt] = (unsigned short)(((uint32_t)(tg.edh[t] - tg.gdh[gi]) * P.row_bytes) >> 4);
          ph.wslot[nt] = (unsigned char)g.woff[t];
        }
      }
      ph.gfirst[tg.ngroups] = (unsigned char)nt;
    }
  }
  if (tiles == 0) return NSG_OK;
  P.nphases = nph;
  P.total_tiles = tiles;
  const uint32_t tail = (2 * TM_MAX_STAGES + 17) * 8 + 64;
  int stages = (int)((TM_SMEM_LIMIT - 1024 - tail - (uint32_t)P.nw * TM_W_BYTES) / P.slot_bytes);
  if (stages > TM_MAX_STAGES) stages = TM_MAX_STAGES;
  NSG_REQUIRE(stages >= 2, "conv_gemm_tma: weights do not leave room for the activation ring");
  P.stages = stages;
  // two MMA-issuing threads when each can own at least two slots of the ring (NSG_CONV_ISSUERS=1: one, for A/B runs)
  static const int issuers = [] { const char* v = getenv("NSG_CONV_ISSUERS"); return v && v[0] == '1' ? 1 : 2; }();
  P.issuers = stages >= 4 ? issuers : 1;
  // (default on: 0.617 against 0.621 ms per step, the 4 x 4 stride-2 forward 12.5 against 14.3 us warm; NSG_CONV_ELECT=0
  // keeps the single-lane loop for A/B runs -- profiles/r2_elect_issue_ab.txt)
  static const int elect = [] { const char* v = getenv("NSG_CONV_ELECT"); return v && v[0] == '0' ? 0 : 1; }();
  P.elect = elect;
  P.dbg = g_timeline;
  P.w_early = pdl_weights_settled(stream) ? 1 : 0;
  const size_t smem = 1024 + (size_t)P.nw * TM_W_BYTES + (size_t)stages * P.slot_bytes + tail;
  CUtensorMap tm[4];
  if (int rc = encode_input_maps(tm, in, cg->batch, cg->in_h, cg->in_w, P.in_stride,
                                 one_box ? 8 + hf[P.swap] : P.box_w, one_box ? 16 + hs[P.swap] : P.box_h + halo, P.swap))
    return rc;
  // (one CTA per SM even when that leaves 124 CTAs with 2 tiles and 24 with 3: giving every CTA 3 tiles on 107 CTAs
  // was measured and is no faster -- 19.0 against 17.5 us for the 3x3 forward, profiles/r2_toggles_ab.txt)
  const int grid = tiles < num_sms() ? tiles : num_sms();
  // streaming: 4 accumulators, so that neither issuer waits for an epilogue before the CTA's fifth tile
  if (!hold) return launch_conv_tma<4, false>(grid, smem, stream, tm, packed, out, e, f, P);
  const int per_cta = (tiles + grid - 1) / grid;
  if (per_cta > 64) {
    set_error("conv_bn_tma: %d tiles per CTA do not fit (at most 64); run the convolution and the BatchNorm pass "
              "separately (nsg_conv_bn_tma_fits)", per_cta);
    return NSG_E_UNSUPPORTED;
  }
  // more tiles than the 8 TMEM accumulators: the surplus is read back in pass 2 from what pass 1 stored
  NSG_REQUIRE(per_cta <= 8 || out, "conv_bn_tma: with more than 8 tiles per CTA the first output (h / g) is required");
  // a grid barrier needs every CTA resident: one CTA per SM of THIS device
  NSG_REQUIRE(grid <= num_sms(), "conv_bn_tma: grid %d exceeds the SM count", grid);
  if (per_cta <= 4) return launch_conv_tma<4, true>(grid, smem, stream, tm, packed, out, e, f, P);
  return launch_conv_tma<8, true>(grid, smem, stream, tm, packed, out, e, f, P);
}

// ntaps[i] > 0: a 64 -> 64 layer with ntaps images.  ntaps[i] == -1 / -2: the single-channel layers' small operands,
// ONE image: -1 = rows are the 64 channels, K = the 16 taps (w[k + r*out_stride]); -2 = rows are the 16 taps,
// K = the 64 channels (w[r + k*in_stride]).
int conv_pack_many(int n, const float* const* w, void* const* dst, const int* ntaps, const long long* in_stride,
                   const long long* out_stride, cudaStream_t stream) {
  NSG_REQUIRE(n >= 0 && n <= 24, "conv_pack_many: at most 24 images per call (got %d)", n);
  if (n == 0) return NSG_OK;
  PackList L{};
  L.n = n;
  int blocks = 0;
  for (int i = 0; i < n; ++i) {
    NSG_REQUIRE(w[i] && dst[i] && ((ntaps[i] >= 1 && ntaps[i] <= TM_MAX_TAPS) || ntaps[i] == -1 || ntaps[i] == -2),
                "conv_pack_many: bad item %d", i);
    NSG_REQUIRE((reinterpret_cast<uintptr_t>(dst[i]) & 127) == 0, "conv_pack_many: image %d must be 128-byte aligned", i);
    L.first_block[i] = blocks;
    if (ntaps[i] > 0)
      L.it[i] = PackItem{w[i], static_cast<uint8_t*>(dst[i]), ntaps[i], 64, 64, in_stride[i], out_stride[i]};
    else if (ntaps[i] == -1)
      L.it[i] = PackItem{w[i], static_cast<uint8_t*>(dst[i]), 1, 16, 64, in_stride[i], out_stride[i]};
    else
      L.it[i] = PackItem{w[i], static_cast<uint8_t*>(dst[i]), 1, 64, 16, in_stride[i], out_stride[i]};
    blocks += (ntaps[i] > 0 ? ntaps[i] : 1) * 2;
  }
  L.first_block[n] = blocks;
  launch_pdl(conv_pack_many_kernel, dim3(blocks), dim3(256), 0, stream, L);
  pdl_note_pack(stream);
  NSG_CHECK_LAUNCH("conv_pack_many_kernel");
  return NSG_OK;
}

// ---------------------------------------------------------------------------------------------- weight gradient
constexpr int WG_KP = 128;                         // pixels per box (GEMM K per stage)
constexpr uint32_t WG_SLOT = WG_KP * 128;          // one WG_KP-pixel x 64-channel bf16 box
constexpr int C1W_KP = 64;                         // the single-channel layers' weight-gradient kernel (below)
constexpr uint32_t C1W_SLOT = C1W_KP * 128;
constexpr int WG_MAX_STAGES = 8;

// One blockIdx.y group = the taps of one kernel COLUMN (kw fixed).  Its taps differ only by a vertical shift (per
// row-parity view when in_stride = 2), so they are served by ONE halo box per view, (box_h + halo) rows tall: tap kh's
// operand is the box advanced by its shift * box_w pixels -- a multiple of the 1024-byte swizzle atom because
// box_w % 8 == 0.  Operands (taps, then an all-ones block for the column sums of Q, then a zero block if the count is
// odd) are paired in address order into M = 128 MMAs; the LBO field of the MN-major descriptor is the byte distance
// between the two 64-channel blocks of a pair, whatever it is.
constexpr int WG_MAX_OPS = 6;                      // 4 taps + ones + zeros
constexpr uint32_t WG_CONST_FLAG = 0x40000000u;    // operand offset is relative to the constant region, not the stage
struct WgGroup {
  int nboxes, nops;
  int bmap[2], bdw[2], bdh[2];   // per halo box: tensor map, coordinate offsets from the tile origin
  uint32_t off[WG_MAX_OPS];      // byte offset of each operand (stage-relative, or WG_CONST_FLAG | constant-relative)
  int tap[WG_MAX_OPS];           // tap index kh * k + kw; -2 = ones (column sums); -1 = zeros
};
struct TmaWgradParams {
  int batch, q_h, q_w;
  int box_w, box_h;              // Q box; P boxes are box_h + halo rows
  int tiles_h, tiles_w, total_tiles;
  int stages;
  uint32_t box_bytes, stage_bytes;   // one P halo box; boxes + Q
  int slice_elems;               // 0: add into ONE packed image with vector atomics; else CTA x stores its partial
                                 // image into slice x of [gridDim.x][ntaps][64][64] (no atomics, summed by the unpack)
  int dbg;
  int issuers;                   // 1 / 2 MMA-issuing threads (2: the operand pairs of a tile alternate between warps 1 and 6)
  unsigned long long* stamps;    // optional clock64 stamps of CTA (0, 0) (nsg_debug_timeline), null in production
  WgGroup g[4];
};

// instruction descriptor, kind::f16: D = f32, A = B = bf16, both MN-major
__device__ __host__ constexpr uint32_t idesc_bf16_mnmajor(uint32_t m, uint32_t n) {
  return idesc_bf16(m, n) | (1u << 15) | (1u << 16);
}
// MN-major SWIZZLE_128B descriptor: LBO = bytes between 64-element blocks along M/N, SBO = 1024 (8-pixel groups)
__device__ __forceinline__ uint64_t sw128_desc_mnmajor(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | (64ull << 32) |
         (1ull << 46) | (2ull << 61);
}
__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}

// Warps: 0 = TMA producer, 1 and 6 = MMA issuers, 2-5 = epilogue.  With two issuers each takes every other operand
// pair (= accumulator) of every tile: both wait on the stage's full barrier and both commit to its empty barrier.
template <int TMEM_COLS>
__global__ void __launch_bounds__(WG_THREADS, 1)
wgrad_tma_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tp0,
                 const __grid_constant__ CUtensorMap tp1, const __grid_constant__ CUtensorMap tp2,
                 const __grid_constant__ CUtensorMap tp3, float* __restrict__ dwp, float* __restrict__ colsum,
                 const __grid_constant__ TmaWgradParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  const WgGroup& G = P.g[blockIdx.y];
  uint8_t* sConst = sm + (size_t)P.stages * P.stage_bytes;               // [ones | zeros], WG_SLOT each
  uint64_t* bars = reinterpret_cast<uint64_t*>(sConst + 2 * WG_SLOT);
  uint64_t* full = bars;                         // [8]
  uint64_t* empty = bars + WG_MAX_STAGES;        // [8]
  uint64_t* acc_full = bars + 2 * WG_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int npairs = G.nops / 2;
  const int nissuers = (P.issuers == 2 && npairs >= 2) ? 2 : 1;
  unsigned long long* stamps = (blockIdx.x == 0 && blockIdx.y == 0) ? P.stamps : nullptr;
  if (tid == 0) tm_stamp(stamps, 0);
  const int ntiles_mine = ((int)blockIdx.x < P.total_tiles)
                              ? (P.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  for (int i = tid; i < (int)(2 * WG_SLOT / 16); i += WG_THREADS) {
    const uint32_t fill = i < (int)(WG_SLOT / 16) ? 0x3F803F80u : 0u;
    reinterpret_cast<uint4*>(sConst)[i] = make_uint4(fill, fill, fill, fill);
  }
  fence_proxy_async_smem();
  // Both single-thread loops below are bound by their own instruction latency (clock64 timeline: 1130 cycles per
  // 3x3 tile in the MMA loop, 700 in the producer, before these tables), so everything that does not change from
  // tile to tile is computed here, once: the A descriptor of every (stage, pair) and the B descriptor of every stage.
  __shared__ uint64_t sDA[WG_MAX_STAGES][WG_MAX_OPS / 2];
  __shared__ uint64_t sDB[WG_MAX_STAGES];
  if (tid >= 64 && tid < 64 + WG_MAX_STAGES * (WG_MAX_OPS / 2)) {
    const int s = (tid - 64) / (WG_MAX_OPS / 2), j = (tid - 64) % (WG_MAX_OPS / 2);
    if (s < P.stages && j < npairs) {
      const uint32_t base = smem_u32(sm + (size_t)s * P.stage_bytes), cbase = smem_u32(sConst);
      const uint32_t o0 = G.off[2 * j], o1 = G.off[2 * j + 1];
      const uint32_t a0 = (o0 & WG_CONST_FLAG) ? cbase + (o0 & ~WG_CONST_FLAG) : base + o0;
      const uint32_t a1 = (o1 & WG_CONST_FLAG) ? cbase + (o1 & ~WG_CONST_FLAG) : base + o1;
      sDA[s][j] = sw128_desc_mnmajor(a0, a1 - a0);
      if (j == 0) sDB[s] = sw128_desc_mnmajor(base + (uint32_t)G.nboxes * P.box_bytes, WG_SLOT);
    }
  }
  if (tid == 0) {
    for (int s = 0; s < WG_MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], (uint32_t)nissuers);
    }
    mbar_init(acc_full, (uint32_t)nissuers);
    fence_mbar_init();
    tma::prefetch_desc(&tq);
    tma::prefetch_desc(&tp0);
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) tm_stamp(stamps, 1);

  if (ntiles_mine > 0) {
    if (warp == 0) {
      // ---------------------------------------------------------------- TMA producer: the lanes decode 32 tiles'
      // coordinates at once (the divisions), lane 0 issues them
      const uint32_t q_off = (uint32_t)G.nboxes * P.box_bytes;
      const CUtensorMap* m0 = G.bmap[0] == 0 ? &tp0 : (G.bmap[0] == 1 ? &tp1 : (G.bmap[0] == 2 ? &tp2 : &tp3));
      const CUtensorMap* m1 = G.bmap[1] == 0 ? &tp0 : (G.bmap[1] == 1 ? &tp1 : (G.bmap[1] == 2 ? &tp2 : &tp3));
      const int bdw0 = G.bdw[0], bdh0 = G.bdh[0], bdw1 = G.bdw[1], bdh1 = G.bdh[1];
      const bool two = G.nboxes > 1, skip_p = (P.dbg & 2) != 0;
      const uint32_t tx = skip_p ? WG_SLOT : q_off + WG_SLOT;
      const uint32_t sm0 = smem_u32(sm);
      int s = 0;
      uint32_t wait_parity = 1;            // parity of the phase BEFORE the first completion: returns at once
      pdl_wait();                          // common.cuh (PDL): everything above is independent of earlier kernels
      pdl_trigger();
      for (int i0 = 0; i0 < ntiles_mine; i0 += 32) {
        int local = (int)blockIdx.x + (i0 + lane) * (int)gridDim.x;
        const int tw = local % P.tiles_w;
        local /= P.tiles_w;
        const int my_w0 = tw * P.box_w, my_h0 = (local % P.tiles_h) * P.box_h, my_n = local / P.tiles_h;
        const int cnt = min(32, ntiles_mine - i0);
        for (int t = 0; t < cnt; ++t) {
          const int w0 = __shfl_sync(0xffffffffu, my_w0, t), h0 = __shfl_sync(0xffffffffu, my_h0, t);
          const int n = __shfl_sync(0xffffffffu, my_n, t);
          if (lane == 0) {
            mbar_wait(&empty[s], wait_parity);
            const uint32_t stage = sm0 + (uint32_t)s * P.stage_bytes;
            mbar_arrive_expect_tx(&full[s], tx);
            tma::load_4d_u32(stage + q_off, &tq, 0, w0, h0, n, &full[s]);
            if (!skip_p) {
              tma::load_4d_u32(stage, m0, 0, w0 + bdw0, h0 + bdh0, n, &full[s]);
              if (two) tma::load_4d_u32(stage + P.box_bytes, m1, 0, w0 + bdw1, h0 + bdh1, n, &full[s]);
            }
            if (i0 + t < 10) tm_stamp(stamps, 2 + i0 + t);     // 2..11: producer issued tile i
          }
          if (++s == P.stages) {
            s = 0;
            wait_parity ^= 1;
          }
        }
      }
      if (lane == 0) tm_stamp(stamps, 12);
    } else if (warp == 1 || warp == 6) {
      // ---------------------------------------------------------------- MMA issuers
      const int me = warp == 1 ? 0 : 1;
      if (lane == 0 && me < nissuers) {
        const uint32_t idesc = idesc_bf16_mnmajor(128, TM_C);
        const int np = (P.dbg & 1) ? 0 : npairs;
        int s = 0;
        uint32_t parity = 0;
        for (int i = 0; i < ntiles_mine; ++i) {
          mbar_wait(&full[s], parity);
          tc_fence_after();
          if (i < 10 && me == 0) tm_stamp(stamps, 13 + i);   // 13..22: tile i landed
          const uint64_t db = sDB[s];
          for (int j = me; j < np; j += nissuers) {
            const uint64_t da = sDA[s][j];
            const uint32_t d = tmem_base + (uint32_t)j * TM_C;
            // K steps of 16 pixels = 2048 bytes = 128 in the descriptors' 16-byte address field (no carry: < 256 KB)
            umma_bf16(d, da, db, idesc, i != 0);
#pragma unroll
            for (int ks = 1; ks < WG_KP / 16; ++ks) umma_bf16(d, da + ks * 128, db + ks * 128, idesc, true);
          }
          umma_commit(&empty[s]);
          if (++s == P.stages) {
            s = 0;
            parity ^= 1;
          }
        }
        umma_commit(acc_full);
        if (me == 0) tm_stamp(stamps, 23);                   // all MMAs issued
      }
    } else if (warp < 6) {
      // ---------------------------------------------------------------- epilogue: thread = (block of the pair, channel a)
      pdl_wait();
      mbar_wait(acc_full, 0);
      tc_fence_after();
      if (tid == 64) tm_stamp(stamps, 24);                   // accumulators complete
      const int q = warp & 3;
      const int row = q * 32 + lane;
      const int a = row & 63, sp = row >> 6;
      for (int j = 0; j < npairs; ++j) {
        const int tap = G.tap[2 * j + sp];
        uint32_t acc0[32], acc1[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * TM_C);
        tmem_ld32(taddr, acc0);
        tmem_ld32(taddr + 32, acc1);
        tmem_ld_wait();
        if (tap >= 0 && P.slice_elems) {
          float4* dst = reinterpret_cast<float4*>(dwp + (size_t)blockIdx.x * P.slice_elems +
                                                  ((size_t)tap * TM_C + a) * TM_C);
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            dst[v] = make_float4(__uint_as_float(acc0[4 * v]), __uint_as_float(acc0[4 * v + 1]),
                                 __uint_as_float(acc0[4 * v + 2]), __uint_as_float(acc0[4 * v + 3]));
            dst[8 + v] = make_float4(__uint_as_float(acc1[4 * v]), __uint_as_float(acc1[4 * v + 1]),
                                     __uint_as_float(acc1[4 * v + 2]), __uint_as_float(acc1[4 * v + 3]));
          }
        } else if (tap >= 0) {
          float* dst = dwp + ((size_t)tap * TM_C + a) * TM_C;
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            red_add_v4(dst + 4 * v, __uint_as_float(acc0[4 * v]), __uint_as_float(acc0[4 * v + 1]),
                       __uint_as_float(acc0[4 * v + 2]), __uint_as_float(acc0[4 * v + 3]));
            red_add_v4(dst + 32 + 4 * v, __uint_as_float(acc1[4 * v]), __uint_as_float(acc1[4 * v + 1]),
                       __uint_as_float(acc1[4 * v + 2]), __uint_as_float(acc1[4 * v + 3]));
          }
        } else if (tap == -2 && colsum) {
          // every row of the ones block holds the column sums of Q; row a contributes column a
          float v = 0.f;
#pragma unroll
          for (int c = 0; c < 32; ++c) {
            if (a == c) v = __uint_as_float(acc0[c]);
            if (a == 32 + c) v = __uint_as_float(acc1[c]);
          }
          atomicAdd(colsum + a, v);
        }
      }
    }
  }
  if (tid == 64) tm_stamp(stamps, 25);                       // epilogue of warp 2 done
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
  if (tid == 0) tm_stamp(stamps, 26);
}

// dw[t + a*a_stride + b*b_stride] += sum_s packed[s][t][a][b]  for every listed layer, one launch
struct UnpackItem {
  const float* src;
  float* dw;
  int ntaps, nslices;
  long long a_stride, b_stride;
};
struct UnpackList {
  int n;
  int first_block[25];           // 16 blocks of 256 threads per tap
  UnpackItem it[24];
};
__global__ void __launch_bounds__(256) wgrad_unpack_kernel(const __grid_constant__ UnpackList L) {
  pdl_wait();
  pdl_trigger();
  int li = 0;
  while (li + 1 < L.n && (int)blockIdx.x >= L.first_block[li + 1]) ++li;
  const UnpackItem& u = L.it[li];
  // block = 256 consecutive elements of one tap image = 64 float4 columns; thread = (column, slice residue mod 4).
  // Slices are summed in a fixed order (residue classes, then the four classes): no dependence on CTA timing.
  if (u.nslices == 1) {      // one image (the atomics form): a plain strided read-modify-write, nothing to reduce
    const int gid = ((int)blockIdx.x - L.first_block[li]) * 256 + threadIdx.x;
    const int b = gid & 63, a = (gid >> 6) & 63, t = gid >> 12;
    if (t < u.ntaps) u.dw[t + (long long)a * u.a_stride + (long long)b * u.b_stride] += __ldg(u.src + gid);
    return;
  }
  __shared__ float4 part[4][64];
  const int col = threadIdx.x & 63, sg = threadIdx.x >> 6;
  const int e0 = ((int)blockIdx.x - L.first_block[li]) * 256 + col * 4;           // (t, a, b), b fastest
  const size_t pitch = (size_t)u.ntaps * TM_C * TM_C;
  const float4* src = reinterpret_cast<const float4*>(u.src + e0);
  const size_t pitch4 = pitch / 4;
  float4 s0 = make_float4(0.f, 0.f, 0.f, 0.f), s1 = s0;
  int sidx = sg;
  for (; sidx + 4 < u.nslices; sidx += 8) {
    const float4 v0 = __ldg(src + (size_t)sidx * pitch4);
    const float4 v1 = __ldg(src + (size_t)(sidx + 4) * pitch4);
    s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
    s1.x += v1.x; s1.y += v1.y; s1.z += v1.z; s1.w += v1.w;
  }
  if (sidx < u.nslices) {
    const float4 v0 = __ldg(src + (size_t)sidx * pitch4);
    s0.x += v0.x; s0.y += v0.y; s0.z += v0.z; s0.w += v0.w;
  }
  part[sg][col] = make_float4(s0.x + s1.x, s0.y + s1.y, s0.z + s1.z, s0.w + s1.w);
  __syncthreads();
  {
    const int e = threadIdx.x;                       // one element per thread for the strided store
    const int c = e >> 2, k = e & 3;
    const float* p0 = reinterpret_cast<const float*>(&part[0][c]) + k;
    const float v = (p0[0] + p0[256]) + (p0[512] + p0[768]);
    const int gid = ((int)blockIdx.x - L.first_block[li]) * 256 + e;
    const int b = gid & 63, a = (gid >> 6) & 63, t = gid >> 12;
    if (t < u.ntaps) u.dw[t + (long long)a * u.a_stride + (long long)b * u.b_stride] += v;
  }
}

template <int TMEM_COLS>
static int launch_wgrad_tma(dim3 grid, size_t smem, const CUtensorMap* tq, const CUtensorMap* tp, float* dwp,
                            float* colsum, const TmaWgradParams& P, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(wgrad_tma_kernel<TMEM_COLS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)TM_SMEM_LIMIT);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(wgrad_tma_kernel)");
    configured = true;
  }
  launch_pdl(wgrad_tma_kernel<TMEM_COLS>, dim3(grid), dim3(WG_THREADS), smem, stream, *tq, tp[0], tp[1], tp[2], tp[3], dwp, colsum, P);
  NSG_CHECK_LAUNCH("wgrad_tma_kernel");
  return NSG_OK;
}

// dwp (packed fp32 [k*k][64][64], ACCUMULATED) += sum_m P[gather(m, tap), a] * Q[m, b]; q_colsum[b] += sum_m Q[m, b]
//
// Split: blockIdx.y = kernel column kw (k groups), blockIdx.x = num_sms / k CTAs striding over the 64-pixel tiles.
// History of the split, measured warm on B200 (3x3, 64 x 20 x 32 pixels): every CTA serving all 9 taps from 9 separate
// boxes left room for only 2 stages of 80 KB, 24.4 us; 3 taps per group (5 stages), 15.5 us; the 3 taps of a COLUMN
// from one halo box (this version) cuts the P bytes by a further third.  The kernel is bound by the L2 -> SM box
// traffic, not by its partials: replacing the atomics by plain stores gains 1-2 us (nsg_conv_wgrad_tma_sliced).
static int wgrad_tiles(const nsg_conv_geom* cg) {
  int bw, bh, l2;
  pick_box_w(cg->out_w, WG_KP, &bw, &bh, &l2);
  return ((cg->out_w + bw - 1) / bw) * ((cg->out_h + bh - 1) / bh) * cg->batch;
}
static int wgrad_grid_x(const nsg_conv_geom* cg, int total_tiles) {
  int x = num_sms() / cg->ksize;
  if (x < 1) x = 1;
  if (x > total_tiles) x = total_tiles;
  return x;
}

int conv_wgrad_slices(const nsg_conv_geom* cg) {
  if (tma_shape_ok(cg, "conv_wgrad_slices")) return 0;
  const int tiles = wgrad_tiles(cg);
  if (tiles <= 0) return 1;
  return wgrad_grid_x(cg, tiles);
}

int conv_wgrad_tma(const nsg_conv_geom* cg, const void* p, const void* q, float* dwp, float* q_colsum, int sliced,
                   cudaStream_t stream) {
  if (int rc = tma_shape_ok(cg, "conv_wgrad_tma")) return rc;
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(dwp) & 15) == 0, "conv_wgrad_tma: packed gradient image must be 16-byte aligned");
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (reinterpret_cast<uintptr_t>(q) & 15) == 0,
              "conv_wgrad_tma: tensors must be 16-byte aligned");
  const int k = cg->ksize, stride = cg->stride;
  NSG_REQUIRE(k >= 1 && k <= 4 && (stride == 1 || stride == 2), "conv_wgrad_tma: kernel %d stride %d not served", k, stride);
  const long long M = (long long)cg->batch * cg->out_h * cg->out_w;
  if (M == 0) {
    if (sliced) {          // the one slice conv_wgrad_slices promised must still be defined
      cudaError_t err = cudaMemsetAsync(dwp, 0, (size_t)k * k * TM_C * TM_C * sizeof(float), stream);
      if (err != cudaSuccess) return cuda_fail(err, "cudaMemsetAsync(wgrad slice)");
    }
    return NSG_OK;
  }
  TmaWgradParams P{};
  P.batch = cg->batch; P.q_h = cg->out_h; P.q_w = cg->out_w;
  int l2;
  pick_box_w(cg->out_w, WG_KP, &P.box_w, &P.box_h, &l2);
  P.tiles_w = (cg->out_w + P.box_w - 1) / P.box_w;
  P.tiles_h = (cg->out_h + P.box_h - 1) / P.box_h;
  P.total_tiles = P.tiles_w * P.tiles_h * cg->batch;
  P.slice_elems = sliced ? k * k * TM_C * TM_C : 0;
  // tuning aid behind profiles/r1_final_wgrad_timeline.txt: 1 = issue no MMAs, 2 = load no P boxes (results are wrong)
  static const int wgrad_dbg = getenv("NSG_WGRAD_DBG") ? atoi(getenv("NSG_WGRAD_DBG")) : 0;
  P.dbg = wgrad_dbg;
  static const int wg_issuers = [] { const char* v = getenv("NSG_WGRAD_ISSUERS"); return v && v[0] == '1' ? 1 : 2; }();
  P.issuers = wg_issuers;
  P.stamps = g_timeline;

  // vertical shift of tap kh inside its (row-parity) view, and the halo the tallest class needs
  int sh[4], cls[4], halo = 0, nboxes_max = 0;
  for (int kh = 0; kh < k; ++kh) {
    const int dh = kh - cg->pad;
    sh[kh] = stride == 1 ? dh : (dh >> 1);
    cls[kh] = stride == 1 ? 0 : (dh & 1);
  }
  int cmin[2] = {1 << 20, 1 << 20}, cmax[2] = {-(1 << 20), -(1 << 20)};
  for (int kh = 0; kh < k; ++kh) {
    cmin[cls[kh]] = sh[kh] < cmin[cls[kh]] ? sh[kh] : cmin[cls[kh]];
    cmax[cls[kh]] = sh[kh] > cmax[cls[kh]] ? sh[kh] : cmax[cls[kh]];
  }
  for (int c = 0; c < 2; ++c)
    if (cmax[c] >= cmin[c]) {
      ++nboxes_max;
      if (cmax[c] - cmin[c] > halo) halo = cmax[c] - cmin[c];
    }
  const uint32_t row_bytes = (uint32_t)P.box_w * 128u;
  P.box_bytes = (uint32_t)(P.box_h + halo) * row_bytes;
  P.stage_bytes = (uint32_t)nboxes_max * P.box_bytes + WG_SLOT;
  int max_ops = 0;
  for (int kw = 0; kw < k; ++kw) {
    WgGroup& G = P.g[kw];
    const int dw = kw - cg->pad;
    int box_of_cls[2] = {-1, -1};
    G.nboxes = 0;
    for (int c = 0; c < 2; ++c)
      if (cmax[c] >= cmin[c]) {
        const int b = G.nboxes++;
        box_of_cls[c] = b;
        G.bmap[b] = stride == 1 ? 0 : c * 2 + (dw & 1);
        G.bdw[b] = stride == 1 ? dw : (dw >> 1);
        G.bdh[b] = cmin[c];
      }
    // operands in address order: box 0's taps by shift, then box 1's, then the constant blocks
    G.nops = 0;
    for (int b = 0; b < G.nboxes; ++b)
      for (int s_ = -8; s_ <= 8; ++s_)
        for (int kh = 0; kh < k; ++kh)
          if (box_of_cls[cls[kh]] == b && sh[kh] == s_) {
            G.off[G.nops] = (uint32_t)b * P.box_bytes + (uint32_t)(sh[kh] - cmin[cls[kh]]) * row_bytes;
            G.tap[G.nops] = kh * k + kw;
            ++G.nops;
          }
    if (q_colsum && kw == 0) {
      G.off[G.nops] = WG_CONST_FLAG | 0u;
      G.tap[G.nops] = -2;
      ++G.nops;
    }
    if (G.nops & 1) {
      G.off[G.nops] = WG_CONST_FLAG | WG_SLOT;
      G.tap[G.nops] = -1;
      ++G.nops;
    }
    if (G.nops > max_ops) max_ops = G.nops;
  }
  const uint32_t tail = 2 * WG_SLOT + (2 * WG_MAX_STAGES + 2) * 8 + 64;
  int stages = (int)((TM_SMEM_LIMIT - 1024 - tail) / P.stage_bytes);
  if (stages > WG_MAX_STAGES) stages = WG_MAX_STAGES;
  NSG_REQUIRE(stages >= 2, "conv_wgrad_tma: stage does not fit twice in shared memory");
  P.stages = stages;
  const size_t smem = 1024 + (size_t)stages * P.stage_bytes + tail;
  CUtensorMap tq, tp[4];
  if (int rc = tma::encode_nhwc64(&tq, q, cg->out_w, cg->out_h, cg->batch, TM_C, (long long)cg->out_w * TM_C,
                                  (long long)cg->out_h * cg->out_w * TM_C, P.box_w, P.box_h))
    return rc;
  if (int rc = encode_input_maps(tp, p, cg->batch, cg->in_h, cg->in_w, stride, P.box_w, P.box_h + halo)) return rc;
  const dim3 grid((unsigned)wgrad_grid_x(cg, P.total_tiles), (unsigned)k);
  const int cols = max_ops / 2 * TM_C;
  if (cols <= 64) return launch_wgrad_tma<64>(grid, smem, &tq, tp, dwp, q_colsum, P, stream);
  if (cols <= 128) return launch_wgrad_tma<128>(grid, smem, &tq, tp, dwp, q_colsum, P, stream);
  return launch_wgrad_tma<256>(grid, smem, &tq, tp, dwp, q_colsum, P, stream);
}

int wgrad_unpack_many(int n, const float* const* src, float* const* dw, const int* ntaps, const int* nslices,
                      const long long* a_stride, const long long* b_stride, cudaStream_t stream) {
  NSG_REQUIRE(n >= 0 && n <= 24, "wgrad_unpack_many: at most 24 layers per call (got %d)", n);
  if (n == 0) return NSG_OK;
  UnpackList L{};
  L.n = n;
  int blocks = 0;
  for (int i = 0; i < n; ++i) {
    NSG_REQUIRE(src[i] && dw[i] && ntaps[i] >= 1 && ntaps[i] <= TM_MAX_TAPS, "wgrad_unpack_many: bad item %d", i);
    const int ns = nslices ? nslices[i] : 1;
    NSG_REQUIRE(ns >= 1 && ns <= 4096, "wgrad_unpack_many: item %d has %d slices", i, ns);
    L.first_block[i] = blocks;
    L.it[i] = UnpackItem{src[i], dw[i], ntaps[i], ns, a_stride[i], b_stride[i]};
    blocks += ntaps[i] * 16;
  }
  L.first_block[n] = blocks;
  launch_pdl(wgrad_unpack_kernel, dim3(blocks), dim3(256), 0, stream, L);
  NSG_CHECK_LAUNCH("wgrad_unpack_kernel");
  return NSG_OK;
}


// ============================================================================================== single-channel ends
// The two single-channel layers (models.py:165 Conv2d(1 -> 64, 4, 2, 1); models.py:182 ConvTranspose2d(64 -> 1, 4, 2, 1))
// on the tensor cores.  Their 64-channel side makes every per-element instruction 64x more expensive than the
// arithmetic of the other side, so the 64-channel reduction / expansion belongs in ONE tcgen05.mma per 128 pixels:
//   c1t_conv   out[128 px, 64] = patches[128 px, K = 16 taps] x W[64, 16]^T   (encoder.0 forward, decoder.6 dgrad)
//   c1t_convT  p[128 px, 16 taps] = v[128 px, K = 64] x W[16, 64]^T, then every output pixel sums its <= 4 products
//   c1t_wgrad  dW[16 taps, 64] = patches^T[16, pixels] x Q[pixels, 64]        (pixels = GEMM K, as in wgrad_tma)
// The 16-wide operands live in the verified 128-byte-row SWIZZLE_128B layouts with only their first 32 bytes
// (K-major: one K = 16 step) or first 16 channels (MN-major) meaningful.

struct C1tConvParams {
  int batch, in_h, in_w, out_h, out_w;
  int box_w, box_h, log2_bw;
  int tiles_h, tiles_w, total_tiles;
};

__global__ void __launch_bounds__(320, 1)
c1t_conv_kernel(const __nv_bfloat16* __restrict__ in1, const uint8_t* __restrict__ wimg,
                __nv_bfloat16* __restrict__ out, EpiArgs epi, BnFuse bn, const __grid_constant__ C1tConvParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  uint8_t* sW = sm;                                  // one 64 x 64 image (K >= 16 zero)
  uint8_t* sA = sm + TM_W_BYTES;                     // four patch tiles (tile i uses buffer i & 3)
  __shared__ __align__(16) float sPar[6 * TM_C];
  __shared__ float sStat[4 * 2 * TM_C];
  __shared__ __align__(8) uint64_t bars[9];          // wfull, full[4], acc_full[4]
  __shared__ uint32_t tmem_slot;
  uint64_t* wfull = bars;
  uint64_t* full = bars + 1;
  uint64_t* acc_full = bars + 5;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  if (tid == 0) {
    mbar_init(wfull, 1);
    for (int b = 0; b < 4; ++b) {
      mbar_init(&full[b], 256);
      mbar_init(&acc_full[b], 1);
    }
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc<256>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;
  const int ntiles_mine = ((int)blockIdx.x < P.total_tiles)
                              ? (P.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == 0) {
    if (lane == 0 && ntiles_mine > 0) {
      pdl_wait();            // encoder.0's image is written by the kernel right before this one
      pdl_trigger();
      mbar_arrive_expect_tx(wfull, TM_W_BYTES);
      bulk_g2s(sW, wimg, TM_W_BYTES, wfull);
    }
  } else if (warp == 1) {
    if (lane == 0 && ntiles_mine > 0) {
      const uint32_t idesc = idesc_bf16(TM_M, TM_C);
      const uint64_t b_desc = sw128_desc(smem_u32(sW));
      mbar_wait(wfull, 0);
      for (int i = 0; i < ntiles_mine; ++i) {
        const int b = i & 3;
        mbar_wait(&full[b], (uint32_t)((i >> 2) & 1));
        tc_fence_after();
        umma_bf16(tmem_base + (uint32_t)b * TM_C, sw128_desc(smem_u32(sA + (size_t)b * TM_A_BYTES)), b_desc, idesc, 0);
        umma_commit(&acc_full[b]);
      }
    }
  } else {
    // all 8 warps work on every tile: thread = (output pixel, 32-channel half).  The per-channel statistics are carried
    // in registers across the CTA's ~9 tiles and reduced once at the end (they used to cost two 32 x 32 shuffle
    // transposes per tile and half -- this kernel is bound by its epilogue's instruction count).
    const int half = (warp - 2) >> 2, q = warp & 3;
    const int r = q * 32 + lane;
    const int ch0 = half * 32;
    const int bw = r & (P.box_w - 1), bh = r >> P.log2_bw;
    const int e = tid - 64;
    const unsigned short* in16 = reinterpret_cast<const unsigned short*>(in1);
    pdl_wait();
    pdl_trigger();
    if (e < TM_C) {
      const int c = e;
      const float bias = epi.bias ? __ldg(epi.bias + c) : 0.f;
      float osc = 1.f, osh = bias;
      if (bn.mode == NSG_BNF_EVAL) {                 // BatchNorm in eval mode folds into the output affine
        const float is = 1.0f / sqrtf(__ldg(bn.running_var + c) + bn.eps);
        osc = __ldg(bn.gamma + c) * is;
        osh = fmaf(bias - __ldg(bn.running_mean + c), osc, __ldg(bn.beta + c));
      }
      sPar[0 * TM_C + c] = osh;
      sPar[1 * TM_C + c] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mscale + c) : 1.f;
      sPar[2 * TM_C + c] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mshift + c) : 0.f;
      sPar[3 * TM_C + c] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.mean + c) : 0.f;
      sPar[4 * TM_C + c] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.invstd + c) : 0.f;
      sPar[5 * TM_C + c] = osc;
    }
    named_bar(2, 256);
    // Per tile the chain is: 8 scattered 2-byte loads -> half a patch row into shared memory -> one MMA round trip ->
    // epilogue.  It is software-pipelined over the four buffers: the patch of tile i+2 is loaded into registers and
    // the patch of tile i+1 is handed to the tensor core BEFORE the epilogue of tile i.
    struct TileRow {
      int n, oh, ow;
      bool valid;
    };
    auto row_of = [&](int i) {
      int t = (int)blockIdx.x + i * (int)gridDim.x;
      const int tw = t % P.tiles_w;
      t /= P.tiles_w;
      TileRow R;
      R.n = t / P.tiles_h;
      R.oh = (t % P.tiles_h) * P.box_h + bh;
      R.ow = tw * P.box_w + bw;
      R.valid = R.oh < P.out_h && R.ow < P.out_w;
      return R;
    };
    uint32_t pk[4];
    auto load_patch = [&](int i) {        // this thread's half (kernel rows 2*half, 2*half + 1) of the 4 x 4 input patch
      const TileRow R = row_of(i);
#pragma unroll
      for (int kk = 0; kk < 2; ++kk) {
        const int kh = half * 2 + kk;
        const int ih = 2 * R.oh - 1 + kh;
        const bool rok = R.valid && ih >= 0 && ih < P.in_h;
        const unsigned short* rowp = in16 + ((long long)R.n * P.in_h + ih) * P.in_w;
        unsigned short v[4];
#pragma unroll
        for (int kw = 0; kw < 4; ++kw) {
          const int iw = 2 * R.ow - 1 + kw;
          v[kw] = (rok && iw >= 0 && iw < P.in_w) ? __ldg(rowp + iw) : (unsigned short)0;
        }
        pk[kk * 2 + 0] = (uint32_t)v[0] | ((uint32_t)v[1] << 16);
        pk[kk * 2 + 1] = (uint32_t)v[2] | ((uint32_t)v[3] << 16);
      }
    };
    auto submit = [&](int buf) {          // patch registers -> operand tile `buf` -> tensor core (taps kh*4 + kw along K)
      uint8_t* myA = sA + (size_t)buf * TM_A_BYTES;
      *reinterpret_cast<uint4*>(myA + sw128_offset(r, half)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
      fence_proxy_async_smem();
      tc_fence_before();                  // orders this thread's earlier tcgen05.ld of that accumulator before the arrive
      mbar_arrive(&full[buf]);
    };
    if (ntiles_mine > 0) {
      load_patch(0);
      submit(0);
      if (1 < ntiles_mine) load_patch(1);
    }
    const bool do_stats = epi.stats_mode != NSG_STATS_NONE;
    const bool bwd_stats = epi.stats_mode == NSG_STATS_BNBWD;
    float s1[32], s2[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) s1[j] = s2[j] = 0.f;
    for (int i = 0; i < ntiles_mine; ++i) {
      if (i + 1 < ntiles_mine) {
        submit((i + 1) & 3);
        if (i + 2 < ntiles_mine) load_patch(i + 2);
      }
      const TileRow R = row_of(i);
      const bool valid = R.valid;
      const long long oh = valid ? (((long long)R.n * P.out_h + R.oh) * P.out_w + R.ow) * TM_C + ch0 : 0;
      const bool use_add = valid && epi.add_src != nullptr;
      uint4 radd[4];
      if (use_add) ld_half(epi.add_src, oh, radd);
      mbar_wait(&acc_full[i & 3], (uint32_t)((i >> 2) & 1));
      tc_fence_after();
      uint32_t acc[32];
      tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(i & 3) * TM_C + (uint32_t)ch0, acc);
      tmem_ld_wait();
      float x[32];
      epi_half(x, acc, epi, sPar, oh, valid, ch0, use_add, radd, bn.mode == NSG_BNF_EVAL);
      if (valid) {
        st16bf(out + oh, &x[0]);
        st16bf(out + oh + 16, &x[16]);
      }
      if (do_stats) {
        if (bwd_stats) {
          if (valid) {
            uint4 rx[4];
            ld_half(epi.xhat_src, oh, rx);
            float mu[32], is[32];
            ld32s(sPar + 3 * TM_C + ch0, mu);
            ld32s(sPar + 4 * TM_C + ch0, is);
#pragma unroll
            for (int g8 = 0; g8 < 4; ++g8) {
              float hx[8];
              unpack8bf(rx[g8], hx);
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                s1[g8 * 8 + j] += x[g8 * 8 + j];
                s2[g8 * 8 + j] = fmaf(x[g8 * 8 + j], (hx[j] - mu[g8 * 8 + j]) * is[g8 * 8 + j], s2[g8 * 8 + j]);
              }
            }
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            s1[j] += x[j];
            s2[j] = fmaf(x[j], x[j], s2[j]);
          }
        }
      }
    }
    if (do_stats) {
      const float t1 = transpose_sum32(s1, lane);
      const float t2 = transpose_sum32(s2, lane);
      // one slot per (TMEM lane quarter, statistic, channel), summed in a FIXED order: the CTA's contribution does not
      // depend on which warp finishes first, so the forward pass is reproducible run to run (the fp64 atomics across
      // CTAs differ at 1e-16, far below the fp32 coefficients derived from them)
      sStat[q * 2 * TM_C + ch0 + lane] = t1;
      sStat[q * 2 * TM_C + TM_C + ch0 + lane] = t2;
      named_bar(1, 256);
      if (e < 2 * TM_C)
        atomicAdd(epi.stats + e, (double)((sStat[e] + sStat[2 * TM_C + e]) + (sStat[4 * TM_C + e] + sStat[6 * TM_C + e])));
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_base);
}

static bool c1t_geom_ok(const nsg_conv_geom* cg) {
  return cg->ksize == 4 && cg->stride == 2 && cg->pad == 1 && cg->dtype == NSG_BF16;
}

// Conv2d(1 -> 64) forward / ConvTranspose2d(64 -> 1) data gradient; packed = 8 KB image from conv_pack_many(ntaps = -1)
int conv_c1_tc(const nsg_conv_geom* cg, const void* in, const void* packed, const nsg_epilogue* epi,
               const nsg_bn_fuse* bnf, void* out, cudaStream_t stream) {
  if (!(c1t_geom_ok(cg) && cg->gather == 0 && cg->cin == 1 && cg->cout == TM_C &&
        cg->out_h == (cg->in_h + 2 - 4) / 2 + 1 && cg->out_w == (cg->in_w + 2 - 4) / 2 + 1)) {
    set_error("conv_c1_tc serves bf16 Conv2d(1 -> 64, k4, s2, p1)-shaped gathers only");
    return NSG_E_UNSUPPORTED;
  }
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 127) == 0, "conv_c1_tc: packed weights must be 128-byte aligned");
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(out) & 31) == 0, "conv_c1_tc: the output must be 32-byte aligned");
  EpiArgs e{};
  if (int rc = make_epi(epi, &e)) return rc;
  NSG_REQUIRE(e.tanh_out == 0 || e.tanh_out == 2, "conv_c1_tc: output activation must be none or ReLU");
  BnFuse f{};
  if (int rc = make_bnfuse(bnf, e, &f)) return rc;
  NSG_REQUIRE(f.mode == NSG_BNF_NONE || f.mode == NSG_BNF_EVAL, "conv_c1_tc: only the eval-mode BatchNorm fold is fused here");
  C1tConvParams P{};
  P.batch = cg->batch; P.in_h = cg->in_h; P.in_w = cg->in_w; P.out_h = cg->out_h; P.out_w = cg->out_w;
  pick_box_w(cg->out_w, TM_M, &P.box_w, &P.box_h, &P.log2_bw);
  P.tiles_w = (cg->out_w + P.box_w - 1) / P.box_w;
  P.tiles_h = (cg->out_h + P.box_h - 1) / P.box_h;
  P.total_tiles = P.tiles_w * P.tiles_h * cg->batch;
  if (P.total_tiles == 0) return NSG_OK;
  const size_t smem = 1024 + TM_W_BYTES + 4 * TM_A_BYTES;
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(c1t_conv_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(c1t_conv_kernel)");
    configured = true;
  }
  const int grid = P.total_tiles < num_sms() ? P.total_tiles : num_sms();
  launch_pdl(c1t_conv_kernel, dim3(grid), dim3(320), smem, stream, static_cast<const __nv_bfloat16*>(in), static_cast<const uint8_t*>(packed),
                                               static_cast<__nv_bfloat16*>(out), e, f, P);
  NSG_CHECK_LAUNCH("c1t_conv_kernel");
  return NSG_OK;
}


// ---------------------------------------------------------------------------------------------- 64 -> 1 transposed forward
struct C1tConvTParams {
  int batch, in_h, in_w, out_h, out_w;
  int R;                 // input rows per work item (its output rows 2*r0 .. 2*(r0+R)-1)
  int rblocks, cblocks;  // work items per image
  int cstep;             // interior input columns per column block (in_w when cblocks == 1, else 126)
  int stages;
  int pstride;
  int tanh_out;
};
constexpr int C1T_MAX_BOXES = 16;

__device__ __forceinline__ void tmem_ld16_(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

__global__ void __launch_bounds__(320, 1)
c1t_convT_kernel(const __grid_constant__ CUtensorMap tm, const uint8_t* __restrict__ wimg,
                 __nv_bfloat16* __restrict__ out, const float* __restrict__ bias,
                 const __grid_constant__ C1tConvTParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  uint8_t* sW = sm;                                              // 16 tap rows x 64 channels (rows >= 16 zero)
  uint8_t* sA = sm + TM_W_BYTES;                                 // ring of 128-pixel boxes
  float* sP = reinterpret_cast<float*>(sA + (size_t)P.stages * TM_A_BYTES);   // [boxes][16][pstride]
  __shared__ __align__(8) uint64_t full[TM_MAX_STAGES], empty[TM_MAX_STAGES], acc_full[C1T_MAX_BOXES], wfull;
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nboxes = P.R + 2;
  int item = blockIdx.x;
  const int cb = item % P.cblocks;
  item /= P.cblocks;
  const int rb = item % P.rblocks, n = item / P.rblocks;
  const int r0 = rb * P.R;
  const int ic0 = cb * P.cstep, ic1 = min(P.in_w, ic0 + P.cstep);
  const int bx0 = P.cblocks == 1 ? 0 : ic0 - 1;

  if (tid == 0) {
    for (int s = 0; s < TM_MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int b = 0; b < C1T_MAX_BOXES; ++b) mbar_init(&acc_full[b], 1);
    mbar_init(&wfull, 1);
    fence_mbar_init();
    tma::prefetch_desc(&tm);
  }
  if (warp == 1) tmem_alloc<256>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (warp == 0) {
    if (lane == 0) {
      pdl_wait();            // set-up above touches no global memory (common.cuh, PDL)
      pdl_trigger();
      mbar_arrive_expect_tx(&wfull, TM_W_BYTES);
      bulk_g2s(sW, wimg, TM_W_BYTES, &wfull);
      for (int b = 0; b < nboxes; ++b) {
        const int s = b % P.stages, round = b / P.stages;
        if (round > 0) mbar_wait(&empty[s], (uint32_t)((round - 1) & 1));
        mbar_arrive_expect_tx(&full[s], TM_A_BYTES);
        tma::load_4d(sA + (size_t)s * TM_A_BYTES, &tm, 0, bx0, r0 - 1 + b, n, &full[s]);
      }
    }
  } else if (warp == 1) {
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(TM_M, 16);
      const uint64_t bdesc = sw128_desc(smem_u32(sW));
      mbar_wait(&wfull, 0);
      for (int b = 0; b < nboxes; ++b) {
        const int s = b % P.stages, round = b / P.stages;
        mbar_wait(&full[s], (uint32_t)(round & 1));
        tc_fence_after();
        const uint64_t a = sw128_desc(smem_u32(sA + (size_t)s * TM_A_BYTES));
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem_base + (uint32_t)b * 16, a + 2 * ks, bdesc + 2 * ks, idesc, ks != 0);
        umma_commit(&empty[s]);
        umma_commit(&acc_full[b]);
      }
    }
  } else {
    pdl_wait();              // bias read, output writes
    if (warp < 6) {
      // tap products of box b, pixel px = q*32 + lane: TMEM -> sP[b][t][px]
      const int q = warp & 3;
      const int px = q * 32 + lane;
      for (int b = 0; b < nboxes; ++b) {
        mbar_wait(&acc_full[b], 0);
        tc_fence_after();
        uint32_t v[16];
        tmem_ld16_(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)b * 16, v);
        tmem_ld_wait();
        float* dst = sP + (size_t)b * 16 * P.pstride + px;
#pragma unroll
        for (int t = 0; t < 16; ++t) dst[(size_t)t * P.pstride] = __uint_as_float(v[t]);
      }
    }
    named_bar(1, 256);
    // every output pixel sums its (up to) four tap products: kh == (oh+1) mod 2, kw == (ow+1) mod 2
    const float b0 = bias ? __ldg(bias) : 0.f;
    const int oh_end = min(2 * (r0 + P.R), P.out_h);
    const int ncols = 2 * (ic1 - ic0);
    const int npix = (oh_end - 2 * r0) * ncols;
    for (int i = tid - 64; i < npix; i += 256) {
      const int oh = 2 * r0 + i / ncols, ow = 2 * ic0 + i % ncols;
      float acc = b0;
#pragma unroll
      for (int a = 0; a < 2; ++a) {
        const int kh = ((oh + 1) & 1) + 2 * a;
        const int ih = (oh + 1 - kh) >> 1;
        if (ih < 0 || ih >= P.in_h) continue;
        const int lr = ih - (r0 - 1);
#pragma unroll
        for (int b = 0; b < 2; ++b) {
          const int kw = ((ow + 1) & 1) + 2 * b;
          const int iw = (ow + 1 - kw) >> 1;
          if (iw < 0 || iw >= P.in_w) continue;
          acc += sP[((size_t)lr * 16 + kh * 4 + kw) * P.pstride + (iw - bx0)];
        }
      }
      out[((long long)n * P.out_h + oh) * P.out_w + ow] = __float2bfloat16_rn(P.tanh_out ? tanhf(acc) : acc);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<256>(tmem_base);
}

// ConvTranspose2d(64 -> 1, 4, 2, 1) forward; packed = 8 KB image from conv_pack_many(ntaps = -2)
int convT_c1_tc(const nsg_conv_geom* cg, const void* in, const void* packed, const float* bias, int tanh_out, void* out,
                cudaStream_t stream) {
  if (!(c1t_geom_ok(cg) && cg->gather == 1 && cg->cin == TM_C && cg->cout == 1 && cg->out_h == 2 * cg->in_h &&
        cg->out_w == 2 * cg->in_w)) {
    set_error("convT_c1_tc serves bf16 ConvTranspose2d(64 -> 1, k4, s2, p1) only");
    return NSG_E_UNSUPPORTED;
  }
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 127) == 0, "convT_c1_tc: packed weights must be 128-byte aligned");
  C1tConvTParams P{};
  P.batch = cg->batch; P.in_h = cg->in_h; P.in_w = cg->in_w; P.out_h = cg->out_h; P.out_w = cg->out_w;
  P.tanh_out = tanh_out;
  P.cblocks = cg->in_w <= 128 ? 1 : (cg->in_w + 125) / 126;
  P.cstep = P.cblocks == 1 ? cg->in_w : 126;
  P.pstride = 130;
  long long per_img = num_sms() / ((long long)cg->batch * P.cblocks);     // row blocks per image for one wave
  if (per_img < 1) per_img = 1;
  if (per_img > cg->in_h) per_img = cg->in_h;
  int R = (int)((cg->in_h + per_img - 1) / per_img);
  if (R > 11) R = 11;                                                     // TMEM columns and shared memory
  P.R = R;
  P.rblocks = (cg->in_h + R - 1) / R;
  P.stages = R + 2 < 4 ? R + 2 : 4;
  const size_t smem = 1024 + TM_W_BYTES + (size_t)P.stages * TM_A_BYTES + (size_t)(R + 2) * 16 * P.pstride * sizeof(float);
  CUtensorMap tm;
  if (int rc = tma::encode_nhwc64(&tm, in, cg->in_w, cg->in_h, cg->batch, TM_C, (long long)cg->in_w * TM_C,
                                  (long long)cg->in_h * cg->in_w * TM_C, 128, 1))
    return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(c1t_convT_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)TM_SMEM_LIMIT);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(c1t_convT_kernel)");
    configured = true;
  }
  const long long grid = (long long)cg->batch * P.rblocks * P.cblocks;
  launch_pdl(c1t_convT_kernel, dim3((unsigned)grid), dim3(320), smem, stream, tm, static_cast<const uint8_t*>(packed),
                                                          static_cast<__nv_bfloat16*>(out), bias, P);
  NSG_CHECK_LAUNCH("c1t_convT_kernel");
  return NSG_OK;
}


// ---------------------------------------------------------------------------------------------- weight gradient, one side single-channel
struct C1tWgradParams {
  int batch, in_h, in_w, q_h, q_w;     // single-channel grid P (in) and 64-channel grid Q
  int box_w, box_h, log2_bw;           // 64-pixel boxes of Q
  int tiles_h, tiles_w, total_tiles;
  int stages;
  long long w_tap_stride_c;            // elements between consecutive 64-side channels in dw (taps are fastest)
};

__global__ void __launch_bounds__(192, 1)
c1t_wgrad_kernel(const __grid_constant__ CUtensorMap tq, const __nv_bfloat16* __restrict__ p1,
                 float* __restrict__ dw, float* __restrict__ q_colsum, const __grid_constant__ C1tWgradParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  constexpr uint32_t STAGE = 3 * C1W_SLOT;            // patch tile (16 taps of 64 "channels"), ones tile, Q box
  __shared__ __align__(8) uint64_t fullQ[WG_MAX_STAGES], fullP[WG_MAX_STAGES], empty[WG_MAX_STAGES], acc_full;
  __shared__ uint32_t tmem_slot;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles_mine = ((int)blockIdx.x < P.total_tiles)
                              ? (P.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
  // patch tiles start as zeros (taps 16..63 of the padded operand never change), ones tiles as bf16 1.0
  for (int s = 0; s < P.stages; ++s) {
    uint4* d0 = reinterpret_cast<uint4*>(sm + (size_t)s * STAGE);
    for (int i = tid; i < (int)(C1W_SLOT / 16); i += 192) {
      d0[i] = make_uint4(0u, 0u, 0u, 0u);
      d0[C1W_SLOT / 16 + i] = make_uint4(0x3F803F80u, 0x3F803F80u, 0x3F803F80u, 0x3F803F80u);
    }
  }
  fence_proxy_async_smem();
  if (tid == 0) {
    for (int s = 0; s < WG_MAX_STAGES; ++s) {
      mbar_init(&fullQ[s], 1);
      mbar_init(&fullP[s], 128);
      mbar_init(&empty[s], 1);
    }
    mbar_init(&acc_full, 1);
    fence_mbar_init();
    tma::prefetch_desc(&tq);
  }
  if (warp == 1) tmem_alloc<64>(&tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = tmem_slot;

  if (ntiles_mine > 0) {
    if (warp == 0) {
      if (lane == 0) {
        pdl_wait();          // set-up above touches no global memory (common.cuh, PDL)
        pdl_trigger();
        for (int i = 0; i < ntiles_mine; ++i) {
          int t = (int)blockIdx.x + i * (int)gridDim.x;
          const int tw = t % P.tiles_w;
          t /= P.tiles_w;
          const int th = t % P.tiles_h, n = t / P.tiles_h;
          const int s = i % P.stages, round = i / P.stages;
          if (round > 0) mbar_wait(&empty[s], (uint32_t)((round - 1) & 1));
          mbar_arrive_expect_tx(&fullQ[s], C1W_SLOT);
          tma::load_4d(sm + (size_t)s * STAGE + 2 * C1W_SLOT, &tq, 0, tw * P.box_w, th * P.box_h, n, &fullQ[s]);
        }
      }
    } else if (warp == 1) {
      if (lane == 0) {
        const uint32_t idesc = idesc_bf16_mnmajor(128, TM_C);
        for (int i = 0; i < ntiles_mine; ++i) {
          const int s = i % P.stages, round = i / P.stages;
          mbar_wait(&fullQ[s], (uint32_t)(round & 1));
          mbar_wait(&fullP[s], (uint32_t)(round & 1));
          tc_fence_after();
          const uint32_t base = smem_u32(sm + (size_t)s * STAGE);
#pragma unroll
          for (int ks = 0; ks < C1W_KP / 16; ++ks) {
            const uint64_t a = sw128_desc_mnmajor(base + ks * 2048, C1W_SLOT);
            const uint64_t b = sw128_desc_mnmajor(base + 2 * C1W_SLOT + ks * 2048, C1W_SLOT);
            umma_bf16(tmem_base, a, b, idesc, (i | ks) != 0);
          }
          umma_commit(&empty[s]);
        }
        umma_commit(&acc_full);
      }
    } else {
      // ---------------------------------------------------------------- patch builders: thread = (pixel, tap half)
      const int t128 = tid - 64;
      const int px = t128 & 63, half = t128 >> 6;
      const int bw = px & (P.box_w - 1), bh = px >> P.log2_bw;
      const unsigned short* in16 = reinterpret_cast<const unsigned short*>(p1);
      pdl_wait();
      for (int i = 0; i < ntiles_mine; ++i) {
        int t = (int)blockIdx.x + i * (int)gridDim.x;
        const int tw = t % P.tiles_w;
        t /= P.tiles_w;
        const int th = t % P.tiles_h, n = t / P.tiles_h;
        const int oh = th * P.box_h + bh, ow = tw * P.box_w + bw;
        const bool valid = oh < P.q_h && ow < P.q_w;
        uint32_t pk[4];
#pragma unroll
        for (int k2 = 0; k2 < 2; ++k2) {
          const int ih = 2 * oh - 1 + half * 2 + k2;
          const bool rok = valid && ih >= 0 && ih < P.in_h;
          const unsigned short* rowp = in16 + ((long long)n * P.in_h + ih) * P.in_w;
          unsigned short v[4];
#pragma unroll
          for (int kw = 0; kw < 4; ++kw) {
            const int iw = 2 * ow - 1 + kw;
            v[kw] = (rok && iw >= 0 && iw < P.in_w) ? __ldg(rowp + iw) : (unsigned short)0;
          }
          pk[k2 * 2 + 0] = (uint32_t)v[0] | ((uint32_t)v[1] << 16);
          pk[k2 * 2 + 1] = (uint32_t)v[2] | ((uint32_t)v[3] << 16);
        }
        const int s = i % P.stages, round = i / P.stages;
        if (round > 0) mbar_wait(&empty[s], (uint32_t)((round - 1) & 1));
        *reinterpret_cast<uint4*>(sm + (size_t)s * STAGE + sw128_offset(px, half)) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
        fence_proxy_async_smem();
        mbar_arrive(&fullP[s]);
      }
      // ---------------------------------------------------------------- epilogue: rows 0..15 = taps, rows 64.. = column sums
      mbar_wait(&acc_full, 0);
      tc_fence_after();
      const int q = warp & 3;
      const int row = q * 32 + lane;
      uint32_t acc0[32], acc1[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16);
      tmem_ld32(taddr, acc0);
      tmem_ld32(taddr + 32, acc1);
      tmem_ld_wait();
      if (row < 16) {
#pragma unroll
        for (int b = 0; b < 32; ++b) {
          atomicAdd(dw + row + (long long)b * P.w_tap_stride_c, __uint_as_float(acc0[b]));
          atomicAdd(dw + row + (long long)(b + 32) * P.w_tap_stride_c, __uint_as_float(acc1[b]));
        }
      } else if (row >= 64 && q_colsum) {
        const int a = row - 64;
        float v = 0.f;
#pragma unroll
        for (int c = 0; c < 32; ++c) {
          if (a == c) v = __uint_as_float(acc0[c]);
          if (a == 32 + c) v = __uint_as_float(acc1[c]);
        }
        atomicAdd(q_colsum + a, v);
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<64>(tmem_base);
}

// dw[t + b*w_cout_stride] += sum_m p1[gather(m, t)] * q[m, b];  q_colsum[b] += sum_m q[m, b]   (cin == 1 side is P)
int wgrad_c1_tc(const nsg_conv_geom* cg, const void* p, const void* q, float* dw, float* q_colsum, cudaStream_t stream) {
  if (!(c1t_geom_ok(cg) && cg->cin == 1 && cg->cout == TM_C && cg->out_h == (cg->in_h + 2 - 4) / 2 + 1 &&
        cg->out_w == (cg->in_w + 2 - 4) / 2 + 1)) {
    set_error("wgrad_c1_tc serves bf16 (1-channel P, 64-channel Q, k4, s2, p1) weight gradients only");
    return NSG_E_UNSUPPORTED;
  }
  C1tWgradParams P{};
  P.batch = cg->batch; P.in_h = cg->in_h; P.in_w = cg->in_w; P.q_h = cg->out_h; P.q_w = cg->out_w;
  P.w_tap_stride_c = cg->w_cout_stride;
  pick_box_w(cg->out_w, C1W_KP, &P.box_w, &P.box_h, &P.log2_bw);
  P.tiles_w = (cg->out_w + P.box_w - 1) / P.box_w;
  P.tiles_h = (cg->out_h + P.box_h - 1) / P.box_h;
  P.total_tiles = P.tiles_w * P.tiles_h * cg->batch;
  if (P.total_tiles == 0) return NSG_OK;
  P.stages = 6;
  const size_t smem = 1024 + (size_t)P.stages * 3 * C1W_SLOT;
  CUtensorMap tq;
  if (int rc = tma::encode_nhwc64(&tq, q, cg->out_w, cg->out_h, cg->batch, TM_C, (long long)cg->out_w * TM_C,
                                  (long long)cg->out_h * cg->out_w * TM_C, P.box_w, P.box_h))
    return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(c1t_wgrad_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(c1t_wgrad_kernel)");
    configured = true;
  }
  const int grid = P.total_tiles < num_sms() ? P.total_tiles : num_sms();
  launch_pdl(c1t_wgrad_kernel, dim3(grid), dim3(192), smem, stream, tq, static_cast<const __nv_bfloat16*>(p), dw, q_colsum, P);
  NSG_CHECK_LAUNCH("c1t_wgrad_kernel");
  return NSG_OK;
}

}  // namespace nsg

namespace nsg {
int conv_tma_hang_init(unsigned int* host_mapped_words) {
  cudaError_t e = cudaMemcpyToSymbol(tc::g_hang_words, &host_mapped_words, sizeof(host_mapped_words));
  return e == cudaSuccess ? NSG_OK : cuda_fail(e, "cudaMemcpyToSymbol(g_hang_words)");
}
}  // namespace nsg
