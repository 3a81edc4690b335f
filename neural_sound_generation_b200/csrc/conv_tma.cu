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
#include "common.cuh"
#include "conv_geom.cuh"
#include "tc_common.cuh"
#include "tma.cuh"

namespace nsg {

using namespace tc;

constexpr int TM_M = 128;                       // pixels per tile
constexpr int TM_C = 64;                        // channels
constexpr uint32_t TM_A_BYTES = TM_M * 128;     // one activation box
constexpr uint32_t TM_W_BYTES = TM_C * 128;     // one tap's weight image (bf16, K-major SWIZZLE_128B)
constexpr int TM_MAX_STAGES = 8;
constexpr int TM_MAX_TAPS = 16;
constexpr uint32_t TM_SMEM_LIMIT = 227 * 1024;

struct TmaPhase {
  int sub_h, sub_w;              // output sub-grid of this stride phase
  int tiles_h, tiles_w, tile0;   // tile grid and index of its first tile
  int ntaps;
  int out_stride, off_h, off_w;  // oh = ohs * out_stride + off_h
  signed char dh[TM_MAX_TAPS], dw[TM_MAX_TAPS];
  unsigned char wslot[TM_MAX_TAPS];
};

struct TmaConvParams {
  int batch, full_h, full_w;
  int in_stride;                 // 1: one tensor map; 2: four parity maps
  int box_w, box_h, log2_bw;
  int nphases, total_tiles, nw, stages;
  TmaPhase ph[4];
};

// ---------------------------------------------------------------------------------------------- packing
struct PackItem {
  const float* w;
  uint8_t* dst;
  int ntaps;
  long long in_stride, out_stride;
};
struct PackList {
  int n;
  int first_block[25];           // prefix of blocks (2 per tap)
  PackItem it[24];
};

// bf16 image per tap: rows = GEMM output channel, 64 GEMM input channels along K, K-major SWIZZLE_128B
__global__ void __launch_bounds__(256) conv_pack_many_kernel(const __grid_constant__ PackList L) {
  int li = 0;
  while (li + 1 < L.n && (int)blockIdx.x >= L.first_block[li + 1]) ++li;
  const PackItem& p = L.it[li];
  const int gid = ((int)blockIdx.x - L.first_block[li]) * 256 + threadIdx.x;
  const int c = gid & 7, r = (gid >> 3) & 63, t = gid >> 9;
  if (t >= p.ntaps) return;
  float v[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) v[i] = __ldg(p.w + t + (long long)(c * 8 + i) * p.in_stride + (long long)r * p.out_stride);
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
__device__ __forceinline__ void st8bf(__nv_bfloat16* p, const float* v) {
  *reinterpret_cast<uint4*>(p) =
      make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
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

__global__ void __launch_bounds__(320, 1)
conv_tma_kernel(const __grid_constant__ CUtensorMap tm0, const __grid_constant__ CUtensorMap tm1,
                const __grid_constant__ CUtensorMap tm2, const __grid_constant__ CUtensorMap tm3,
                const uint8_t* __restrict__ wimg, __nv_bfloat16* __restrict__ out, EpiArgs epi,
                const __grid_constant__ TmaConvParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sW = sm;                                         // nw weight images
  uint8_t* sA = sm + (size_t)P.nw * TM_W_BYTES;             // ring of activation boxes
  uint8_t* tail = sA + (size_t)P.stages * TM_A_BYTES;
  float* sPar = reinterpret_cast<float*>(tail);             // [5][64]: bias, mscale, mshift, mean, invstd
  float* sStat = sPar + 5 * TM_C;                           // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStat + 2 * TM_C);
  uint64_t* full = bars;                                    // [8]
  uint64_t* empty = bars + TM_MAX_STAGES;                   // [8]
  uint64_t* wfull = bars + 2 * TM_MAX_STAGES;               // [1]
  uint64_t* acc_full = wfull + 1;                           // [2]
  uint64_t* acc_empty = acc_full + 2;                       // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int stages = P.stages;

  if (tid < TM_C) {
    sPar[0 * TM_C + tid] = epi.bias ? __ldg(epi.bias + tid) : 0.f;
    sPar[1 * TM_C + tid] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mscale + tid) : 1.f;
    sPar[2 * TM_C + tid] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mshift + tid) : 0.f;
    sPar[3 * TM_C + tid] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.mean + tid) : 0.f;
    sPar[4 * TM_C + tid] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.invstd + tid) : 0.f;
    sStat[tid] = 0.f;
    sStat[TM_C + tid] = 0.f;
  }
  if (tid == 0) {
    for (int s = 0; s < TM_MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(wfull, 1);
    for (int a = 0; a < 2; ++a) {
      mbar_init(&acc_full[a], 1);
      mbar_init(&acc_empty[a], 128);
    }
    fence_mbar_init();
    tma::prefetch_desc(&tm0);
    if (P.in_stride == 2) {
      tma::prefetch_desc(&tm1);
      tma::prefetch_desc(&tm2);
      tma::prefetch_desc(&tm3);
    }
  }
  if (warp == 1) tmem_alloc<128>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const int ntiles_mine = ((int)blockIdx.x < P.total_tiles)
                              ? (P.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0 && ntiles_mine > 0) {
      mbar_arrive_expect_tx(wfull, (uint32_t)P.nw * TM_W_BYTES);
      for (int t = 0; t < P.nw; ++t)
        bulk_g2s(sW + (size_t)t * TM_W_BYTES, wimg + (size_t)t * TM_W_BYTES, TM_W_BYTES, wfull);
      int it = 0;
      for (int i = 0; i < ntiles_mine; ++i) {
        const TmaTile tl = decode_tile(P, (int)blockIdx.x + i * (int)gridDim.x);
        const TmaPhase& ph = P.ph[tl.p];
        const int w0 = tl.tw * P.box_w, h0 = tl.th * P.box_h;
        for (int t = 0; t < ph.ntaps; ++t, ++it) {
          const int s = it % stages, round = it / stages;
          if (round > 0) mbar_wait(&empty[s], (uint32_t)((round - 1) & 1));
          mbar_arrive_expect_tx(&full[s], TM_A_BYTES);
          const int dh = ph.dh[t], dw = ph.dw[t];
          uint8_t* dst = sA + (size_t)s * TM_A_BYTES;
          if (P.in_stride == 1) {
            tma::load_4d(dst, &tm0, 0, w0 + dw, h0 + dh, tl.n, &full[s]);
          } else {
            const int par = (dh & 1) * 2 + (dw & 1);
            const CUtensorMap* tm = par == 0 ? &tm0 : (par == 1 ? &tm1 : (par == 2 ? &tm2 : &tm3));
            tma::load_4d(dst, tm, 0, w0 + (dw >> 1), h0 + (dh >> 1), tl.n, &full[s]);
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0 && ntiles_mine > 0) {
      const uint32_t idesc = idesc_bf16(TM_M, TM_C);
      mbar_wait(wfull, 0);
      int it = 0;
      for (int i = 0; i < ntiles_mine; ++i) {
        const TmaTile tl = decode_tile(P, (int)blockIdx.x + i * (int)gridDim.x);
        const TmaPhase& ph = P.ph[tl.p];
        const int as = i & 1, use = i >> 1;
        if (use > 0) mbar_wait(&acc_empty[as], (uint32_t)((use - 1) & 1));
        tc_fence_after();
        const uint32_t d = tmem_base + (uint32_t)as * TM_C;
        for (int t = 0; t < ph.ntaps; ++t, ++it) {
          const int s = it % stages, round = it / stages;
          mbar_wait(&full[s], (uint32_t)(round & 1));
          tc_fence_after();
          const uint64_t a = sw128_desc(smem_u32(sA + (size_t)s * TM_A_BYTES));
          const uint64_t b = sw128_desc(smem_u32(sW + (size_t)ph.wslot[t] * TM_W_BYTES));
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) umma_bf16(d, a + 2 * ks, b + 2 * ks, idesc, (t | ks) != 0);
          umma_commit(&empty[s]);
        }
        umma_commit(&acc_full[as]);
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue groups: thread = output row
    const int grp = (warp - 2) >> 2;                 // accumulator / tile parity served by this group
    const int q = warp & 3;                          // TMEM lane quarter this warp may read
    const int r = q * 32 + lane;
    const int bw = r & (P.box_w - 1), bh = r >> P.log2_bw;
    const __nv_bfloat16* add_src = static_cast<const __nv_bfloat16*>(epi.add_src);
    const __nv_bfloat16* mask_src = static_cast<const __nv_bfloat16*>(epi.mask_src);
    const __nv_bfloat16* xhat_src = static_cast<const __nv_bfloat16*>(epi.xhat_src);
    for (int i = grp; i < ntiles_mine; i += 2) {
      const TmaTile tl = decode_tile(P, (int)blockIdx.x + i * (int)gridDim.x);
      const TmaPhase& ph = P.ph[tl.p];
      const int ohs = tl.th * P.box_h + bh, ows = tl.tw * P.box_w + bw;
      const bool valid = ohs < ph.sub_h && ows < ph.sub_w;
      const long long o = valid ? ((((long long)tl.n * P.full_h + (ohs * ph.out_stride + ph.off_h)) * P.full_w +
                                    (ows * ph.out_stride + ph.off_w)) * TM_C) : 0;
      mbar_wait(&acc_full[grp], (uint32_t)((i >> 1) & 1));
      tc_fence_after();
      uint32_t acc0[32], acc1[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)grp * TM_C;
      tmem_ld32(taddr, acc0);
      tmem_ld32(taddr + 32, acc1);
      tmem_ld_wait();
      tc_fence_before();
      mbar_arrive(&acc_empty[grp]);
#pragma unroll
      for (int chunk = 0; chunk < 2; ++chunk) {
        const int ch0 = chunk * 32;
        float x[32], y[32];
#pragma unroll
        for (int j = 0; j < 32; ++j)
          x[j] = valid ? __uint_as_float(chunk == 0 ? acc0[j] : acc1[j]) + sPar[ch0 + j] : 0.f;
        if (valid && add_src) {
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {
            float a[8];
            ld8bf(add_src + o + ch0 + g8 * 8, a);
#pragma unroll
            for (int j = 0; j < 8; ++j) x[g8 * 8 + j] += a[j];
          }
        }
        if (valid && epi.mask_mode != NSG_MASK_NONE) {
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) {
            float mk[8];
            ld8bf(mask_src + o + ch0 + g8 * 8, mk);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const int ch = ch0 + g8 * 8 + j;
              const float mv = epi.mask_mode == NSG_MASK_AFFINE ? fmaf(sPar[TM_C + ch], mk[j], sPar[2 * TM_C + ch]) : mk[j];
              x[g8 * 8 + j] = mv > 0.f ? x[g8 * 8 + j] : 0.f;
            }
          }
        }
        if (epi.tanh_out == 2) {
#pragma unroll
          for (int j = 0; j < 32; ++j) x[j] = fmaxf(x[j], 0.f);
        }
        if (valid) {
#pragma unroll
          for (int g8 = 0; g8 < 4; ++g8) st8bf(out + o + ch0 + g8 * 8, &x[g8 * 8]);
        }
        if (epi.stats_mode != NSG_STATS_NONE) {
          if (epi.stats_mode == NSG_STATS_MOMENTS) {
#pragma unroll
            for (int j = 0; j < 32; ++j) y[j] = x[j] * x[j];
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) y[j] = 0.f;
            if (valid) {
#pragma unroll
              for (int g8 = 0; g8 < 4; ++g8) {
                float hx[8];
                ld8bf(xhat_src + o + ch0 + g8 * 8, hx);
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  const int ch = ch0 + g8 * 8 + j;
                  y[g8 * 8 + j] = x[g8 * 8 + j] * ((hx[j] - sPar[3 * TM_C + ch]) * sPar[4 * TM_C + ch]);
                }
              }
            }
          }
          const float s1 = transpose_sum32(x, lane);
          const float s2 = transpose_sum32(y, lane);
          atomicAdd(&sStat[ch0 + lane], s1);
          atomicAdd(&sStat[TM_C + ch0 + lane], s2);
        }
      }
    }
    if (epi.stats_mode != NSG_STATS_NONE) {
      named_bar(1, 256);
      const int e = tid - 64;
      if (e < 2 * TM_C) atomicAdd(epi.stats + e, (double)sStat[e]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<128>(tmem_base);
}

// ---------------------------------------------------------------------------------------------- conv_tma host side
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

static void pick_box(int sub_w, int pixels, int* box_w, int* box_h, int* log2_bw) {
  int bw = 8, l2 = 3;
  while (bw < pixels && bw < sub_w) {
    bw <<= 1;
    ++l2;
  }
  *box_w = bw;
  *box_h = pixels / bw;
  *log2_bw = l2;
}

// tensor maps over `in` (N, in_h, in_w, 64) bf16: plain view (in_stride 1) or the four parity views (in_stride 2)
static int encode_input_maps(CUtensorMap* tm, const void* in, int batch, int in_h, int in_w, int in_stride, int box_w,
                             int box_h) {
  const long long C = TM_C;
  if (in_stride == 1) {
    if (int rc = tma::encode_nhwc64(&tm[0], in, in_w, in_h, batch, C, (long long)in_w * C, (long long)in_h * in_w * C,
                                    box_w, box_h))
      return rc;
    tm[1] = tm[2] = tm[3] = tm[0];
    return NSG_OK;
  }
  for (int ph = 0; ph < 2; ++ph)
    for (int pw = 0; pw < 2; ++pw) {
      const __nv_bfloat16* base = static_cast<const __nv_bfloat16*>(in) + ((long long)ph * in_w + pw) * C;
      const long long dw_ = (in_w - pw + 1) / 2, dh_ = (in_h - ph + 1) / 2;
      if (dw_ < 1 || dh_ < 1) {       // degenerate parity view (1-pixel-wide tensor): never addressed in range
        if (int rc = tma::encode_nhwc64(&tm[ph * 2 + pw], in, 1, 1, batch, 2 * C, 2LL * in_w * C,
                                        (long long)in_h * in_w * C, box_w, box_h))
          return rc;
        continue;
      }
      if (int rc = tma::encode_nhwc64(&tm[ph * 2 + pw], base, dw_, dh_, batch, 2 * C, 2LL * in_w * C,
                                      (long long)in_h * in_w * C, box_w, box_h))
        return rc;
    }
  return NSG_OK;
}

int conv_gemm_tma(const nsg_conv_geom* cg, const void* in, const void* packed, const nsg_epilogue* epi, void* out,
                  cudaStream_t stream) {
  if (int rc = tma_shape_ok(cg, "conv_gemm_tma")) return rc;
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 127) == 0, "conv_gemm_tma: packed weights must be 128-byte aligned");
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(in) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0,
              "conv_gemm_tma: tensors must be 16-byte aligned");
  EpiArgs e{};
  if (int rc = make_epi(epi, &e)) return rc;
  NSG_REQUIRE(e.tanh_out == 0 || e.tanh_out == 2,
              "conv_gemm_tma: output activation must be none (0) or ReLU (2); tanh belongs to the cout == 1 layer");
  TapGeom phases[4];
  const int np = build_phases(cg, phases);
  TmaConvParams P{};
  P.batch = cg->batch; P.full_h = cg->out_h; P.full_w = cg->out_w;
  P.in_stride = phases[0].in_stride;
  NSG_REQUIRE(P.in_stride == 1 || P.in_stride == 2, "conv_gemm_tma: input stride %d", P.in_stride);
  int max_sub_w = 1;
  for (int i = 0; i < np; ++i) max_sub_w = phases[i].sub_w > max_sub_w ? phases[i].sub_w : max_sub_w;
  pick_box(max_sub_w, TM_M, &P.box_w, &P.box_h, &P.log2_bw);
  P.nw = cg->ksize * cg->ksize;
  int tiles = 0, nph = 0;
  for (int i = 0; i < np; ++i) {
    const TapGeom& g = phases[i];
    if (g.sub_h <= 0 || g.sub_w <= 0 || cg->batch == 0) continue;
    NSG_REQUIRE(g.ntaps > 0, "conv_gemm_tma: a stride phase without taps (ksize < stride) is not supported");
    TmaPhase& ph = P.ph[nph++];
    ph.sub_h = g.sub_h; ph.sub_w = g.sub_w;
    ph.tiles_w = (g.sub_w + P.box_w - 1) / P.box_w;
    ph.tiles_h = (g.sub_h + P.box_h - 1) / P.box_h;
    ph.tile0 = tiles;
    tiles += ph.tiles_w * ph.tiles_h * cg->batch;
    ph.ntaps = g.ntaps;
    ph.out_stride = g.out_stride; ph.off_h = g.off_h; ph.off_w = g.off_w;
    for (int t = 0; t < g.ntaps; ++t) {
      ph.dh[t] = (signed char)g.dh[t];
      ph.dw[t] = (signed char)g.dw[t];
      ph.wslot[t] = (unsigned char)g.woff[t];
    }
  }
  if (tiles == 0) return NSG_OK;
  P.nphases = nph;
  P.total_tiles = tiles;
  const uint32_t tail = 7 * TM_C * 4 + (2 * TM_MAX_STAGES + 5) * 8 + 64;
  int stages = (int)((TM_SMEM_LIMIT - 1024 - tail - (uint32_t)P.nw * TM_W_BYTES) / TM_A_BYTES);
  if (stages > TM_MAX_STAGES) stages = TM_MAX_STAGES;
  NSG_REQUIRE(stages >= 2, "conv_gemm_tma: weights do not leave room for the activation ring");
  P.stages = stages;
  const size_t smem = 1024 + (size_t)P.nw * TM_W_BYTES + (size_t)stages * TM_A_BYTES + tail;
  CUtensorMap tm[4];
  if (int rc = encode_input_maps(tm, in, cg->batch, cg->in_h, cg->in_w, P.in_stride, P.box_w, P.box_h)) return rc;
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(conv_tma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)TM_SMEM_LIMIT);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(conv_tma_kernel)");
    configured = true;
  }
  const int grid = tiles < num_sms() ? tiles : num_sms();
  conv_tma_kernel<<<grid, 320, smem, stream>>>(tm[0], tm[1], tm[2], tm[3], static_cast<const uint8_t*>(packed),
                                               static_cast<__nv_bfloat16*>(out), e, P);
  NSG_CHECK_LAUNCH("conv_tma_kernel");
  return NSG_OK;
}

int conv_pack_many(int n, const float* const* w, void* const* dst, const int* ntaps, const long long* in_stride,
                   const long long* out_stride, cudaStream_t stream) {
  NSG_REQUIRE(n >= 0 && n <= 24, "conv_pack_many: at most 24 images per call (got %d)", n);
  if (n == 0) return NSG_OK;
  PackList L{};
  L.n = n;
  int blocks = 0;
  for (int i = 0; i < n; ++i) {
    NSG_REQUIRE(w[i] && dst[i] && ntaps[i] >= 1 && ntaps[i] <= TM_MAX_TAPS, "conv_pack_many: bad item %d", i);
    NSG_REQUIRE((reinterpret_cast<uintptr_t>(dst[i]) & 127) == 0, "conv_pack_many: image %d must be 128-byte aligned", i);
    L.first_block[i] = blocks;
    L.it[i] = PackItem{w[i], static_cast<uint8_t*>(dst[i]), ntaps[i], in_stride[i], out_stride[i]};
    blocks += ntaps[i] * 2;
  }
  L.first_block[n] = blocks;
  conv_pack_many_kernel<<<blocks, 256, 0, stream>>>(L);
  NSG_CHECK_LAUNCH("conv_pack_many_kernel");
  return NSG_OK;
}

// ---------------------------------------------------------------------------------------------- weight gradient
constexpr int WG_KP = 64;                          // pixels per box (GEMM K per stage)
constexpr uint32_t WG_SLOT = WG_KP * 128;          // one 64-pixel x 64-channel bf16 box
constexpr int WG_MAX_STAGES = 8;

struct TmaWgradParams {
  int batch, q_h, q_w;
  int in_stride;                 // P gather: ip = q * in_stride + d
  int box_w, box_h;
  int tiles_h, tiles_w, total_tiles;
  int ntaps, taps_per_group;     // all taps; taps served by one blockIdx.y group
  int nslots;                    // even number of P slots per stage (taps of the group [+ ones slot] [+ zero slot])
  int ones;                      // 1: group 0 carries an all-ones slot right after its taps (column sums of Q)
  int stages;
  signed char dh[TM_MAX_TAPS], dw[TM_MAX_TAPS];
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

template <int TMEM_COLS>
__global__ void __launch_bounds__(192, 1)
wgrad_tma_kernel(const __grid_constant__ CUtensorMap tq, const __grid_constant__ CUtensorMap tp0,
                 const __grid_constant__ CUtensorMap tp1, const __grid_constant__ CUtensorMap tp2,
                 const __grid_constant__ CUtensorMap tp3, float* __restrict__ dwp, float* __restrict__ colsum,
                 const __grid_constant__ TmaWgradParams P) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  const int NS = P.nslots;
  const uint32_t stage_bytes = (uint32_t)(NS + 1) * WG_SLOT;            // P slots, then Q
  uint64_t* bars = reinterpret_cast<uint64_t*>(sm + (size_t)P.stages * stage_bytes);
  uint64_t* full = bars;                         // [8]
  uint64_t* empty = bars + WG_MAX_STAGES;        // [8]
  uint64_t* acc_full = bars + 2 * WG_MAX_STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tap_base = (int)blockIdx.y * P.taps_per_group;
  const int ntaps = min(P.taps_per_group, P.ntaps - tap_base);
  const bool ones = P.ones && blockIdx.y == 0;
  const int ntiles_mine = ((int)blockIdx.x < P.total_tiles)
                              ? (P.total_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  // constant slots (all-ones for the bias gradient, zeros for an odd slot out) are written once per stage
  for (int sl = ntaps; sl < NS; ++sl) {
    const uint32_t fill = (sl == ntaps && ones) ? 0x3F803F80u : 0u;
    for (int s = 0; s < P.stages; ++s) {
      uint4* dst = reinterpret_cast<uint4*>(sm + (size_t)s * stage_bytes + (size_t)sl * WG_SLOT);
      for (int i = tid; i < (int)(WG_SLOT / 16); i += 192) dst[i] = make_uint4(fill, fill, fill, fill);
    }
  }
  fence_proxy_async_smem();
  if (tid == 0) {
    for (int s = 0; s < WG_MAX_STAGES; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(acc_full, 1);
    fence_mbar_init();
    tma::prefetch_desc(&tq);
    tma::prefetch_desc(&tp0);
  }
  if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (ntiles_mine > 0) {
    if (warp == 0) {
      // ---------------------------------------------------------------- TMA producer
      if (lane == 0) {
        for (int i = 0; i < ntiles_mine; ++i) {
          int local = (int)blockIdx.x + i * (int)gridDim.x;
          const int tw = local % P.tiles_w;
          local /= P.tiles_w;
          const int th = local % P.tiles_h, n = local / P.tiles_h;
          const int w0 = tw * P.box_w, h0 = th * P.box_h;
          const int s = i % P.stages, round = i / P.stages;
          if (round > 0) mbar_wait(&empty[s], (uint32_t)((round - 1) & 1));
          uint8_t* stage = sm + (size_t)s * stage_bytes;
          mbar_arrive_expect_tx(&full[s], (uint32_t)(ntaps + 1) * WG_SLOT);
          tma::load_4d(stage + (size_t)NS * WG_SLOT, &tq, 0, w0, h0, n, &full[s]);
          for (int t = 0; t < ntaps; ++t) {
            const int dh = P.dh[tap_base + t], dw = P.dw[tap_base + t];
            if (P.in_stride == 1) {
              tma::load_4d(stage + (size_t)t * WG_SLOT, &tp0, 0, w0 + dw, h0 + dh, n, &full[s]);
            } else {
              const int par = (dh & 1) * 2 + (dw & 1);
              const CUtensorMap* tm = par == 0 ? &tp0 : (par == 1 ? &tp1 : (par == 2 ? &tp2 : &tp3));
              tma::load_4d(stage + (size_t)t * WG_SLOT, tm, 0, w0 + (dw >> 1), h0 + (dh >> 1), n, &full[s]);
            }
          }
        }
      }
    } else if (warp == 1) {
      // ---------------------------------------------------------------- MMA issuer
      if (lane == 0) {
        const uint32_t idesc = idesc_bf16_mnmajor(128, TM_C);
        const int npairs = NS / 2;
        for (int i = 0; i < ntiles_mine; ++i) {
          const int s = i % P.stages, round = i / P.stages;
          mbar_wait(&full[s], (uint32_t)(round & 1));
          tc_fence_after();
          const uint32_t base = smem_u32(sm + (size_t)s * stage_bytes);
          for (int j = 0; j < npairs; ++j) {
            const uint32_t d = tmem_base + (uint32_t)j * TM_C;
#pragma unroll
            for (int ks = 0; ks < WG_KP / 16; ++ks) {
              const uint64_t a = sw128_desc_mnmajor(base + (uint32_t)(2 * j) * WG_SLOT + ks * 2048, WG_SLOT);
              const uint64_t b = sw128_desc_mnmajor(base + (uint32_t)NS * WG_SLOT + ks * 2048, WG_SLOT);
              umma_bf16(d, a, b, idesc, (i | ks) != 0);
            }
          }
          umma_commit(&empty[s]);
        }
        umma_commit(acc_full);
      }
    } else {
      // ---------------------------------------------------------------- epilogue: thread = (slot parity, channel a)
      mbar_wait(acc_full, 0);
      tc_fence_after();
      const int q = warp & 3;
      const int row = q * 32 + lane;
      const int a = row & 63, sp = row >> 6;
      for (int j = 0; j < NS / 2; ++j) {
        const int sl = 2 * j + sp;
        uint32_t acc0[32], acc1[32];
        const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * TM_C);
        tmem_ld32(taddr, acc0);
        tmem_ld32(taddr + 32, acc1);
        tmem_ld_wait();
        if (sl < ntaps) {
          float* dst = dwp + ((size_t)(tap_base + sl) * TM_C + a) * TM_C;
#pragma unroll
          for (int v = 0; v < 8; ++v) {
            red_add_v4(dst + 4 * v, __uint_as_float(acc0[4 * v]), __uint_as_float(acc0[4 * v + 1]),
                       __uint_as_float(acc0[4 * v + 2]), __uint_as_float(acc0[4 * v + 3]));
            red_add_v4(dst + 32 + 4 * v, __uint_as_float(acc1[4 * v]), __uint_as_float(acc1[4 * v + 1]),
                       __uint_as_float(acc1[4 * v + 2]), __uint_as_float(acc1[4 * v + 3]));
          }
        } else if (sl == ntaps && ones && colsum) {
          // every row of the ones slot holds the column sums of Q; row a contributes column a
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
  tc_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc<TMEM_COLS>(tmem_base);
}

// dw[t + a*a_stride + b*b_stride] += packed[t][a][b]  for every listed layer, one launch
struct UnpackItem {
  const float* src;
  float* dw;
  int ntaps;
  long long a_stride, b_stride;
};
struct UnpackList {
  int n;
  int first_block[25];           // 16 blocks of 256 threads per tap
  UnpackItem it[24];
};
__global__ void __launch_bounds__(256) wgrad_unpack_kernel(const __grid_constant__ UnpackList L) {
  int li = 0;
  while (li + 1 < L.n && (int)blockIdx.x >= L.first_block[li + 1]) ++li;
  const UnpackItem& u = L.it[li];
  const int gid = ((int)blockIdx.x - L.first_block[li]) * 256 + threadIdx.x;     // (t, a, b), b fastest
  const int b = gid & 63, a = (gid >> 6) & 63, t = gid >> 12;
  if (t >= u.ntaps) return;
  float* d = u.dw + t + (long long)a * u.a_stride + (long long)b * u.b_stride;
  *d += u.src[gid];
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
  wgrad_tma_kernel<TMEM_COLS><<<grid, 192, smem, stream>>>(*tq, tp[0], tp[1], tp[2], tp[3], dwp, colsum, P);
  NSG_CHECK_LAUNCH("wgrad_tma_kernel");
  return NSG_OK;
}

// dwp (packed fp32 [k*k][64][64], ACCUMULATED) += sum_m P[gather(m, tap), a] * Q[m, b]; q_colsum[b] += sum_m Q[m, b]
int conv_wgrad_tma(const nsg_conv_geom* cg, const void* p, const void* q, float* dwp, float* q_colsum,
                   cudaStream_t stream) {
  if (int rc = tma_shape_ok(cg, "conv_wgrad_tma")) return rc;
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(dwp) & 15) == 0, "conv_wgrad_tma: packed gradient image must be 16-byte aligned");
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(p) & 15) == 0 && (reinterpret_cast<uintptr_t>(q) & 15) == 0,
              "conv_wgrad_tma: tensors must be 16-byte aligned");
  const long long M = (long long)cg->batch * cg->out_h * cg->out_w;
  if (M == 0) return NSG_OK;
  TmaWgradParams P{};
  P.batch = cg->batch; P.q_h = cg->out_h; P.q_w = cg->out_w;
  P.in_stride = cg->stride;
  int l2;
  pick_box(cg->out_w, WG_KP, &P.box_w, &P.box_h, &l2);
  P.tiles_w = (cg->out_w + P.box_w - 1) / P.box_w;
  P.tiles_h = (cg->out_h + P.box_h - 1) / P.box_h;
  P.total_tiles = P.tiles_w * P.tiles_h * cg->batch;
  const int k = cg->ksize;
  P.ntaps = k * k;
  for (int kh = 0; kh < k; ++kh)
    for (int kw = 0; kw < k; ++kw) {
      P.dh[kh * k + kw] = (signed char)(kh - cg->pad);
      P.dw[kh * k + kw] = (signed char)(kw - cg->pad);
    }
  P.ones = q_colsum ? 1 : 0;
  P.taps_per_group = P.ntaps + P.ones <= 10 ? P.ntaps : 8;
  const int groups = (P.ntaps + P.taps_per_group - 1) / P.taps_per_group;
  P.nslots = (P.taps_per_group + P.ones + 1) / 2 * 2;
  const uint32_t stage_bytes = (uint32_t)(P.nslots + 1) * WG_SLOT;
  const uint32_t tail = (2 * WG_MAX_STAGES + 2) * 8 + 64;
  int stages = (int)((TM_SMEM_LIMIT - 1024 - tail) / stage_bytes);
  if (stages > WG_MAX_STAGES) stages = WG_MAX_STAGES;
  NSG_REQUIRE(stages >= 2, "conv_wgrad_tma: stage does not fit twice in shared memory");
  P.stages = stages;
  const size_t smem = 1024 + (size_t)stages * stage_bytes + tail;
  CUtensorMap tq, tp[4];
  if (int rc = tma::encode_nhwc64(&tq, q, cg->out_w, cg->out_h, cg->batch, TM_C, (long long)cg->out_w * TM_C,
                                  (long long)cg->out_h * cg->out_w * TM_C, P.box_w, P.box_h))
    return rc;
  if (int rc = encode_input_maps(tp, p, cg->batch, cg->in_h, cg->in_w, P.in_stride, P.box_w, P.box_h)) return rc;
  int gx = num_sms() / groups;
  if (gx < 1) gx = 1;
  if (gx > P.total_tiles) gx = P.total_tiles;
  const dim3 grid((unsigned)gx, (unsigned)groups);
  const int cols = P.nslots / 2 * TM_C;
  if (cols <= 64) return launch_wgrad_tma<64>(grid, smem, &tq, tp, dwp, q_colsum, P, stream);
  if (cols <= 128) return launch_wgrad_tma<128>(grid, smem, &tq, tp, dwp, q_colsum, P, stream);
  if (cols <= 256) return launch_wgrad_tma<256>(grid, smem, &tq, tp, dwp, q_colsum, P, stream);
  return launch_wgrad_tma<512>(grid, smem, &tq, tp, dwp, q_colsum, P, stream);
}

int wgrad_unpack_many(int n, const float* const* src, float* const* dw, const int* ntaps, const long long* a_stride,
                      const long long* b_stride, cudaStream_t stream) {
  NSG_REQUIRE(n >= 0 && n <= 24, "wgrad_unpack_many: at most 24 layers per call (got %d)", n);
  if (n == 0) return NSG_OK;
  UnpackList L{};
  L.n = n;
  int blocks = 0;
  for (int i = 0; i < n; ++i) {
    NSG_REQUIRE(src[i] && dw[i] && ntaps[i] >= 1 && ntaps[i] <= TM_MAX_TAPS, "wgrad_unpack_many: bad item %d", i);
    L.first_block[i] = blocks;
    L.it[i] = UnpackItem{src[i], dw[i], ntaps[i], a_stride[i], b_stride[i]};
    blocks += ntaps[i] * 16;
  }
  L.first_block[n] = blocks;
  wgrad_unpack_kernel<<<blocks, 256, 0, stream>>>(L);
  NSG_CHECK_LAUNCH("wgrad_unpack_kernel");
  return NSG_OK;
}

}  // namespace nsg
