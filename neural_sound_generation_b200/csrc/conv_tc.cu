// 64-channel convolutions on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// tapgemm_tc: out[m, co] = sum_tap sum_ci pro(in[gather(m, tap), ci]) * W_tap[co, ci]  -- Conv2d / ConvTranspose2d
// forward and both data gradients for cin = cout = 64 (models.py:150,153,168,179 and their ATen backward).
// One CTA = 128 output pixels (UMMA M = 128, N = 64); every tap is one K = 64 block:
//   * 4 producer warps gather the 128 input rows of the tap from NHWC global memory (128 B of bf16 or 256 B of
//     fp32 per row), apply the BatchNorm/ReLU prologue in registers, and store bf16 (hi [+ lo]) into the
//     K-major SWIZZLE_128B operand image of a ring stage; the tap's packed weight image arrives by a 1-D bulk
//     async copy on the same mbarrier;
//   * one MMA thread issues 4 x PASSES tcgen05.mma (K = 16 each) per tap into a 64-column TMEM accumulator and
//     releases the stage with tcgen05.commit;
//   * the producer warps then become the epilogue: tcgen05.ld the accumulator row per thread, bias / residual
//     add / ReLU-mask / per-channel statistics (BatchNorm forward moments or backward reductions), store.
// PASSES: 1 = bf16 x bf16; 2 = activation x (W_hi + W_lo) (exact products for bf16 activations, fp32 weights);
// 3 = split activations as well (fp32 activations: ~2^-16 relative, "fp32-accuracy mode" on tensor cores).
//
// wgrad_tc: dW_tap[a, b] = sum_m P[gather(m, tap), a] * Q[m, b] with the pixel axis as GEMM K: both operands are
// used as MN-major SWIZZLE_128B tiles, which is byte-for-byte the pixel-major image the forward kernel builds.
#include "common.cuh"
#include "conv_geom.cuh"
#define NSG_TU_ID 1
#include "tc_common.cuh"

namespace nsg {

using namespace tc;

constexpr int CT_M = 128;                 // output pixels per CTA
constexpr int CT_C = 64;                  // channels (cin = cout)
constexpr uint32_t CT_A_BYTES = CT_M * 128;   // one bf16 A image (128 rows x 128 B)
constexpr uint32_t CT_B_BYTES = CT_C * 128;   // one bf16 W image (64 rows x 128 B)
constexpr uint32_t CT_WTAP_BYTES = 2 * CT_B_BYTES;   // packed weights per tap: hi image, lo image

// ---------------------------------------------------------------------------------------------- weight packing
// w (reference layout, addressed by the two strides) -> per tap t = kh*k + kw: bf16 hi image, bf16 lo image,
// rows = GEMM output channel, 64 GEMM input channels along K, K-major SWIZZLE_128B.
__global__ void __launch_bounds__(256)
conv_pack_kernel(const float* __restrict__ w, int ntaps, long long in_stride, long long out_stride,
                 uint8_t* __restrict__ packed) {
  const int gid = blockIdx.x * 256 + threadIdx.x;      // (tap, row, chunk)
  const int c = gid & 7, r = (gid >> 3) & 63, t = gid >> 9;
  if (t >= ntaps) return;
  float hi[8], lo[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const float v = __ldg(w + t + (long long)(c * 8 + i) * in_stride + (long long)r * out_stride);
    split_bf16(v, hi[i], lo[i]);
  }
  uint8_t* dst = packed + (size_t)t * CT_WTAP_BYTES + sw128_offset(r, c);
  *reinterpret_cast<uint4*>(dst) =
      make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]), pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
  *reinterpret_cast<uint4*>(dst + CT_B_BYTES) =
      make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
}

// ---------------------------------------------------------------------------------------------- small helpers
// 8 consecutive channels of one pixel as floats
template <typename T>
__device__ __forceinline__ void ld8(const T* p, float (&v)[8]);
template <>
__device__ __forceinline__ void ld8<float>(const float* p, float (&v)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p + 4));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void ld8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
template <typename T>
__device__ __forceinline__ void st8(T* p, const float (&v)[8]);
template <>
__device__ __forceinline__ void st8<float>(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
template <>
__device__ __forceinline__ void st8<__nv_bfloat16>(__nv_bfloat16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) =
      make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]), pack_bf16(v[6], v[7]));
}

// 8 channels of one pixel as loaded (kept raw in registers while the load is in flight)
template <typename T>
struct Raw8;
template <>
struct Raw8<float> {
  float4 a, b;
  __device__ __forceinline__ void load(const float* p) {
    a = __ldg(reinterpret_cast<const float4*>(p));
    b = __ldg(reinterpret_cast<const float4*>(p + 4));
  }
  __device__ __forceinline__ void zero() { a = b = make_float4(0.f, 0.f, 0.f, 0.f); }
  __device__ __forceinline__ void unpack(float (&v)[8]) const {
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
};
template <>
struct Raw8<__nv_bfloat16> {
  uint4 a;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { a = __ldg(reinterpret_cast<const uint4*>(p)); }
  __device__ __forceinline__ void zero() { a = make_uint4(0u, 0u, 0u, 0u); }
  __device__ __forceinline__ void unpack(float (&v)[8]) const {
    const uint32_t w[4] = {a.x, a.y, a.z, a.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
};

// v[i] of lane l -> returns sum over lanes of v[lane] (a 32x32 transpose-reduce in 31 shuffles); v is destroyed
__device__ __forceinline__ float warp_transpose_sum(float (&v)[32], int lane) {
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

// 32 lanes x 16 consecutive fp32 columns
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
}

// ---------------------------------------------------------------------------------------------- tap GEMM
template <int PASSES>
struct TcCfg {
  static constexpr int A_IMGS = PASSES == 3 ? 2 : 1;
  static constexpr int B_IMGS = PASSES >= 2 ? 2 : 1;
  static constexpr uint32_t STAGE_BYTES = A_IMGS * CT_A_BYTES + B_IMGS * CT_B_BYTES;
  static constexpr int STAGES = PASSES == 1 ? 4 : (PASSES == 2 ? 3 : 2);
  static constexpr uint32_t B_OFF = A_IMGS * CT_A_BYTES;
  static constexpr uint32_t TAIL_BYTES = 3 * CT_M * 4 + 7 * CT_C * 4 + 128;
  static constexpr uint32_t SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + TAIL_BYTES;
};

template <typename T, int PASSES>
__global__ void __launch_bounds__(160)
tapgemm_tc_kernel(const T* __restrict__ in, const uint8_t* __restrict__ wpacked, T* __restrict__ out, Pro pro,
                  EpiArgs epi, TapGeom g) {
  using Cfg = TcCfg<PASSES>;
  constexpr int STAGES = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  uint8_t* tail = sm + STAGES * Cfg::STAGE_BYTES;
  int* rowN = reinterpret_cast<int*>(tail);
  int* rowH = rowN + CT_M;
  int* rowW = rowH + CT_M;
  float* sPar = reinterpret_cast<float*>(rowW + CT_M);     // [5][64]: bias, mscale, mshift, mean, invstd
  float* sStat = sPar + 5 * CT_C;                          // [2][64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sStat + 2 * CT_C);   // full[STAGES], empty[STAGES], acc_full
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* acc_full = bars + 2 * STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long M = (long long)g.batch * g.sub_h * g.sub_w;
  const long long m0 = (long long)blockIdx.x * CT_M;

  if (tid < CT_M) {
    const long long m = m0 + tid;
    if (m < M) {
      rowW[tid] = (int)(m % g.sub_w);
      const long long t = m / g.sub_w;
      rowH[tid] = (int)(t % g.sub_h);
      rowN[tid] = (int)(t / g.sub_h);
    } else {
      rowN[tid] = -1;
      rowH[tid] = 0;
      rowW[tid] = 0;
    }
    if (tid < CT_C) {
      sPar[0 * CT_C + tid] = epi.bias ? __ldg(epi.bias + tid) : 0.f;
      sPar[1 * CT_C + tid] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mscale + tid) : 1.f;
      sPar[2 * CT_C + tid] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mshift + tid) : 0.f;
      sPar[3 * CT_C + tid] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.mean + tid) : 0.f;
      sPar[4 * CT_C + tid] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.invstd + tid) : 0.f;
    }
    sStat[tid] = 0.f;
  }
  if (tid == 128) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], CT_M + 1);
      mbar_init(&empty[s], 1);
    }
    mbar_init(acc_full, 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc<64>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16(CT_M, CT_C);
      for (int t = 0; t < g.ntaps; ++t) {
        const int s = t % STAGES;
        mbar_wait(&full[s], (uint32_t)((t / STAGES) & 1));
        tc_fence_after();
        const uint32_t base = smem_u32(sm + (size_t)s * Cfg::STAGE_BYTES);
        const uint64_t a_hi = sw128_desc(base);
        const uint64_t a_lo = sw128_desc(base + CT_A_BYTES);
        const uint64_t b_hi = sw128_desc(base + Cfg::B_OFF);
        const uint64_t b_lo = sw128_desc(base + Cfg::B_OFF + CT_B_BYTES);
#pragma unroll
        for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem_base, a_hi + 2 * ks, b_hi + 2 * ks, idesc, (t | ks) != 0);
        if (PASSES >= 2) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem_base, a_hi + 2 * ks, b_lo + 2 * ks, idesc, 1);
        }
        if (PASSES == 3) {
#pragma unroll
          for (int ks = 0; ks < 4; ++ks) umma_bf16(tmem_base, a_lo + 2 * ks, b_hi + 2 * ks, idesc, 1);
        }
        umma_commit(&empty[s]);
      }
      umma_commit(acc_full);
    }
  } else {
    // ------------------------------------------------------------------ producers: one tap per ring stage
    // Each thread owns the 16-byte channel chunk c of rows rbase + 16*j.  All 8 row loads of a tap are issued
    // back to back, and the loads of tap t+1 are in flight while tap t is transformed and stored.
    const int c = tid & 7;            // channels c*8 .. c*8+7
    const int rbase = tid >> 3;
    float sc[8], sh[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sc[i] = pro.mode >= NSG_PRO_AFFINE_RELU ? __ldg(pro.scale + c * 8 + i) : 1.f;
      sh[i] = pro.mode >= NSG_PRO_AFFINE_RELU ? __ldg(pro.shift + c * 8 + i) : 0.f;
    }
    int rowPix[8], rowHs[8], rowWs[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int r = rbase + 16 * j;
      const int n = rowN[r];
      rowPix[j] = n < 0 ? -1 : n * g.in_h * g.in_w;
      rowHs[j] = rowH[r] * g.in_stride;
      rowWs[j] = rowW[r] * g.in_stride;
    }
    const T* inc = in + c * 8;
    auto issue = [&](int t, Raw8<T> (&dst)[8]) -> unsigned {
      const int dh = g.dh[t], dw = g.dw[t];
      unsigned mask = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int ih = rowHs[j] + dh, iw = rowWs[j] + dw;
        const bool ok = rowPix[j] >= 0 && ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w;
        if (ok) {
          dst[j].load(inc + (long long)(rowPix[j] + ih * g.in_w + iw) * CT_C);
          mask |= 1u << j;
        } else {
          dst[j].zero();
        }
      }
      return mask;
    };
    Raw8<T> cur[8], nxt[8];
    unsigned okc = issue(0, cur), okn = 0;
    for (int t = 0; t < g.ntaps; ++t) {
      const int s = t % STAGES;
      if (t + 1 < g.ntaps) okn = issue(t + 1, nxt);
      if (t >= STAGES) mbar_wait(&empty[s], (uint32_t)(((t / STAGES) - 1) & 1));
      uint8_t* stage = sm + (size_t)s * Cfg::STAGE_BYTES;
      if (tid == 0) {
        mbar_arrive_expect_tx(&full[s], Cfg::B_IMGS * CT_B_BYTES);
        bulk_g2s(stage + Cfg::B_OFF, wpacked + (size_t)g.woff[t] * CT_WTAP_BYTES, Cfg::B_IMGS * CT_B_BYTES, &full[s]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float v[8], hi[8], lo[8];
        cur[j].unpack(v);
        const bool ok = (okc >> j) & 1u;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float x = ok ? pro_apply(pro.mode, v[i], sc[i], sh[i]) : 0.f;
          split_bf16(x, hi[i], lo[i]);
        }
        const uint32_t off = sw128_offset(rbase + 16 * j, c);
        *reinterpret_cast<uint4*>(stage + off) = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]),
                                                            pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
        if (PASSES == 3)
          *reinterpret_cast<uint4*>(stage + CT_A_BYTES + off) = make_uint4(
              pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
      }
      fence_proxy_async_smem();
      mbar_arrive(&full[s]);
#pragma unroll
      for (int j = 0; j < 8; ++j) cur[j] = nxt[j];
      okc = okn;
    }

    // ------------------------------------------------------------------ epilogue: thread = output row
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int r = tid;
    const int n = rowN[r];
    const bool valid = n >= 0;
    long long o = 0;
    if (valid) {
      const int oh = rowH[r] * g.out_stride + g.off_h;
      const int ow = rowW[r] * g.out_stride + g.off_w;
      o = (((long long)n * g.full_h + oh) * g.full_w + ow) * CT_C;
    }
    const T* add_src = static_cast<const T*>(epi.add_src);
    const T* mask_src = static_cast<const T*>(epi.mask_src);
    const T* xhat_src = static_cast<const T*>(epi.xhat_src);
#pragma unroll 1
    for (int chunk = 0; chunk < 2; ++chunk) {
      const int ch0 = chunk * 32;
      uint32_t acc[32];
      tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)ch0, acc);
      tmem_ld_wait();
      float x[32], y[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) x[i] = valid ? __uint_as_float(acc[i]) + sPar[ch0 + i] : 0.f;
      if (valid && add_src) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float a[8];
          ld8<T>(add_src + o + ch0 + q * 8, a);
#pragma unroll
          for (int i = 0; i < 8; ++i) x[q * 8 + i] += a[i];
        }
      }
      if (valid && epi.mask_mode != NSG_MASK_NONE) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float mk[8];
          ld8<T>(mask_src + o + ch0 + q * 8, mk);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int ch = ch0 + q * 8 + i;
            const float mv = epi.mask_mode == NSG_MASK_AFFINE ? fmaf(sPar[CT_C + ch], mk[i], sPar[2 * CT_C + ch]) : mk[i];
            x[q * 8 + i] = mv > 0.f ? x[q * 8 + i] : 0.f;
          }
        }
      }
      if (valid) {
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float a[8];
#pragma unroll
          for (int i = 0; i < 8; ++i) a[i] = x[q * 8 + i];
          st8<T>(out + o + ch0 + q * 8, a);
        }
      }
      if (epi.stats_mode != NSG_STATS_NONE) {
        if (epi.stats_mode == NSG_STATS_MOMENTS) {
#pragma unroll
          for (int i = 0; i < 32; ++i) y[i] = x[i] * x[i];
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) y[i] = 0.f;
          if (valid) {
#pragma unroll
            for (int q = 0; q < 4; ++q) {
              float hx[8];
              ld8<T>(xhat_src + o + ch0 + q * 8, hx);
#pragma unroll
              for (int i = 0; i < 8; ++i) {
                const int ch = ch0 + q * 8 + i;
                y[q * 8 + i] = x[q * 8 + i] * ((hx[i] - sPar[3 * CT_C + ch]) * sPar[4 * CT_C + ch]);
              }
            }
          }
        }
        const float s1 = warp_transpose_sum(x, lane);
        const float s2 = warp_transpose_sum(y, lane);
        atomicAdd(&sStat[ch0 + lane], s1);
        atomicAdd(&sStat[CT_C + ch0 + lane], s2);
      }
    }
    if (epi.stats_mode != NSG_STATS_NONE) {
      asm volatile("bar.sync 1, 128;" ::: "memory");
      atomicAdd(epi.stats + tid, (double)sStat[tid]);     // tid < 128 = [sum(64) | second moment(64)]
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<64>(tmem_base);
}

// ---------------------------------------------------------------------------------------------- host side
static int tc_shape_ok(const nsg_conv_geom* cg, const char* what) {
  if (cg->cin != CT_C || cg->cout != CT_C) {
    set_error("%s: the tcgen05 kernels serve cin == cout == 64 (got %d -> %d)", what, cg->cin, cg->cout);
    return NSG_E_UNSUPPORTED;
  }
  NSG_REQUIRE(cg->ksize >= 1 && cg->ksize <= 4, "%s: ksize %d unsupported (1..4)", what, cg->ksize);
  NSG_REQUIRE(cg->stride >= 1 && cg->stride <= 2, "%s: stride %d unsupported (1..2)", what, cg->stride);
  return NSG_OK;
}

long long conv_pack_bytes(const nsg_conv_geom* cg) {
  if (cg->cin != CT_C || cg->cout != CT_C || cg->ksize < 1 || cg->ksize > 4) return -1;
  return (long long)cg->ksize * cg->ksize * CT_WTAP_BYTES;
}

int conv_pack_weights(const nsg_conv_geom* cg, const float* w, void* packed, cudaStream_t stream) {
  if (int rc = tc_shape_ok(cg, "conv_pack_weights")) return rc;
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 127) == 0, "conv_pack_weights: packed buffer must be 128-byte aligned");
  const int ntaps = cg->ksize * cg->ksize;
  conv_pack_kernel<<<ntaps * 2, 256, 0, stream>>>(w, ntaps, cg->w_cin_stride, cg->w_cout_stride,
                                                   static_cast<uint8_t*>(packed));
  NSG_CHECK_LAUNCH("conv_pack_kernel");
  return NSG_OK;
}

template <typename T, int PASSES>
static int launch_tapgemm_tc(const TapGeom& g, const void* in, const void* packed, void* out, const Pro& p,
                             const EpiArgs& e, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(tapgemm_tc_kernel<T, PASSES>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)TcCfg<PASSES>::SMEM_BYTES);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(tapgemm_tc_kernel)");
    configured = true;
  }
  const long long M = (long long)g.batch * g.sub_h * g.sub_w;
  const unsigned grid = (unsigned)((M + CT_M - 1) / CT_M);
  tapgemm_tc_kernel<T, PASSES><<<grid, 160, TcCfg<PASSES>::SMEM_BYTES, stream>>>(
      (const T*)in, (const uint8_t*)packed, (T*)out, p, e, g);
  NSG_CHECK_LAUNCH("tapgemm_tc_kernel");
  return NSG_OK;
}

int conv_gemm_tc(const nsg_conv_geom* cg, const void* in, const void* packed, const nsg_prologue* pro,
                 const nsg_epilogue* epi, void* out, int passes, cudaStream_t stream) {
  if (int rc = tc_shape_ok(cg, "conv_gemm_tc")) return rc;
  NSG_REQUIRE(passes >= 1 && passes <= 3, "conv_gemm_tc: passes must be 1, 2 or 3");
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(packed) & 127) == 0, "conv_gemm_tc: packed weights must be 128-byte aligned");
  TapGeom phases[4];
  const int np = build_phases(cg, phases);
  Pro p{NSG_PRO_NONE, nullptr, nullptr};
  if (pro) p = Pro{pro->mode, pro->scale, pro->shift};
  if (p.mode >= NSG_PRO_AFFINE_RELU) NSG_REQUIRE(p.scale && p.shift, "conv: affine prologue needs scale/shift");
  EpiArgs e{};
  if (int rc = make_epi(epi, &e)) return rc;
  NSG_REQUIRE(!e.tanh_out, "conv_gemm_tc: tanh epilogue is served by the CUDA-core kernel (cout == 1 layer)");
  if (cg->dtype == NSG_BF16 && passes == 3) passes = 2;     // bf16 activations have no low part
  for (int i = 0; i < np; ++i) {
    const TapGeom& g = phases[i];
    if ((long long)g.batch * g.sub_h * g.sub_w == 0) continue;
    NSG_REQUIRE(g.ntaps > 0, "conv_gemm_tc: a stride phase without taps (ksize < stride) is not supported");
    int rc;
    if (cg->dtype == NSG_F32) {
      rc = passes == 1 ? launch_tapgemm_tc<float, 1>(g, in, packed, out, p, e, stream)
           : passes == 2 ? launch_tapgemm_tc<float, 2>(g, in, packed, out, p, e, stream)
                         : launch_tapgemm_tc<float, 3>(g, in, packed, out, p, e, stream);
    } else if (cg->dtype == NSG_BF16) {
      rc = passes == 1 ? launch_tapgemm_tc<__nv_bfloat16, 1>(g, in, packed, out, p, e, stream)
                       : launch_tapgemm_tc<__nv_bfloat16, 2>(g, in, packed, out, p, e, stream);
    } else {
      set_error("unknown dtype %d", cg->dtype);
      rc = NSG_E_INVALID;
    }
    if (rc) return rc;
  }
  return NSG_OK;
}

// ---------------------------------------------------------------------------------------------- weight gradient
// D_pair[a, b] (+)= P_pair^T (128 x pixels) * Q (pixels x 64): pixels are the GEMM K axis, so the pixel-major
// NHWC tiles are MN-major operands: canonical layout ((8,8,m),(8,k)):((1,8,LBO),(64,SBO)) in elements, i.e. one
// 128-byte row per pixel, 1024-byte groups of 8 pixels (SBO), and the second 64 channels of a 128-row A operand
// LBO bytes away -- which is where the NEXT TAP's tile lives: two taps share one M = 128 instruction.
struct WgradTcGeom {
  int batch, p_h, p_w, q_h, q_w;
  int ksize, stride, pad;
  int tap0, ntaps;               // taps of this launch's blockIdx.y group: tap0 + blockIdx.y * ntaps_per_group
  int taps_per_group, nslots;    // nslots = even number of operand slots per stage (>= taps_per_group)
  long long w_a_stride, w_b_stride;
  int pix_per_cta;
};

// instruction descriptor, kind::f16: D = f32, A = B = bf16, both MN-major
__device__ __host__ constexpr uint32_t idesc_bf16_mn(uint32_t m, uint32_t n) {
  return idesc_bf16(m, n) | (1u << 15) | (1u << 16);
}
// MN-major SWIZZLE_128B descriptor: LBO = bytes between 64-element blocks along M/N, SBO = 1024 (8-pixel groups)
__device__ __forceinline__ uint64_t sw128_desc_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3FFFu) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16) | (64ull << 32) |
         (1ull << 46) | (2ull << 61);
}

template <typename T, bool SPLIT, int TMEM_COLS>
__global__ void __launch_bounds__(160, 1)
wgrad_tc_kernel(const T* __restrict__ P, const T* __restrict__ Q, float* __restrict__ dw,
                float* __restrict__ q_colsum, Pro proP, Pro proQ, WgradTcGeom g) {
  constexpr int KP = SPLIT ? 32 : 64;                 // pixels per stage
  constexpr uint32_t SLOT = KP * 128;                 // one bf16 tile: KP pixels x 64 channels
  constexpr int STAGES = 2;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  const int NS = g.nslots;
  const uint32_t half_bytes = (uint32_t)(NS + 1) * SLOT;                 // P slots + Q, one precision half
  const uint32_t stage_bytes = half_bytes * (SPLIT ? 2 : 1);
  uint8_t* tail = sm + STAGES * stage_bytes;
  float* sCol = reinterpret_cast<float*>(tail);                          // [64]
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail + 256);              // full[2], empty[2], acc_full
  uint64_t* full = bars;
  uint64_t* empty = bars + STAGES;
  uint64_t* acc_full = bars + 2 * STAGES;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int tap_base = g.tap0 + blockIdx.y * g.taps_per_group;
  const int ntaps = min(g.taps_per_group, g.ntaps - blockIdx.y * g.taps_per_group);
  const long long Mtot = (long long)g.batch * g.q_h * g.q_w;
  const long long mbeg = (long long)blockIdx.x * g.pix_per_cta;
  const long long mend = min(Mtot, mbeg + (long long)g.pix_per_cta);
  const int nstages = mend > mbeg ? (int)((mend - mbeg + KP - 1) / KP) : 0;

  if (tid < CT_C) sCol[tid] = 0.f;
  if (tid == 128) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full[s], 128);
      mbar_init(&empty[s], 1);
    }
    mbar_init(acc_full, 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (nstages == 0) {           // whole CTA leaves together (no barrier below is needed)
    if (warp == 4) tmem_dealloc<TMEM_COLS>(tmem_base);
    return;
  }

  if (warp == 4) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = idesc_bf16_mn(128, CT_C);
      const int npairs = NS / 2;
      for (int it = 0; it < nstages; ++it) {
        const int s = it & 1;
        mbar_wait(&full[s], (uint32_t)((it >> 1) & 1));
        tc_fence_after();
        const uint32_t base = smem_u32(sm + (size_t)s * stage_bytes);
        for (int j = 0; j < npairs; ++j) {
          const uint32_t d = tmem_base + (uint32_t)j * CT_C;
#pragma unroll
          for (int ks = 0; ks < KP / 16; ++ks) {
            const uint64_t a_hi = sw128_desc_mn(base + (uint32_t)(2 * j) * SLOT + ks * 2048, SLOT);
            const uint64_t b_hi = sw128_desc_mn(base + (uint32_t)NS * SLOT + ks * 2048, SLOT);
            umma_bf16(d, a_hi, b_hi, idesc, (it | ks) != 0);
            if (SPLIT) {
              const uint64_t a_lo = sw128_desc_mn(base + half_bytes + (uint32_t)(2 * j) * SLOT + ks * 2048, SLOT);
              const uint64_t b_lo = sw128_desc_mn(base + half_bytes + (uint32_t)NS * SLOT + ks * 2048, SLOT);
              umma_bf16(d, a_hi, b_lo, idesc, 1);
              umma_bf16(d, a_lo, b_hi, idesc, 1);
            }
          }
        }
        umma_commit(&empty[s]);
      }
      umma_commit(acc_full);
    }
  } else {
    // ------------------------------------------------------------------ producers
    const int c = tid & 7;
    const int prow = tid >> 3;      // pixel rows prow + 16*j of the stage
    float scP[8], shP[8], scQ[8], shQ[8], colsum[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      scP[i] = proP.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proP.scale + c * 8 + i) : 1.f;
      shP[i] = proP.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proP.shift + c * 8 + i) : 0.f;
      scQ[i] = proQ.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proQ.scale + c * 8 + i) : 1.f;
      shQ[i] = proQ.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proQ.shift + c * 8 + i) : 0.f;
      colsum[i] = 0.f;
    }
    for (int it = 0; it < nstages; ++it) {
      const int s = it & 1;
      if (it >= STAGES) mbar_wait(&empty[s], (uint32_t)(((it >> 1) - 1) & 1));
      uint8_t* stage = sm + (size_t)s * stage_bytes;
#pragma unroll 1
      for (int j = 0; j < KP / 16; ++j) {
        const int pr = prow + 16 * j;
        const long long m = mbeg + (long long)it * KP + pr;
        const bool mv = m < mend;
        int ow = 0, oh = 0, n = 0;
        if (mv) {
          ow = (int)(m % g.q_w);
          const long long t = m / g.q_w;
          oh = (int)(t % g.q_h);
          n = (int)(t / g.q_h);
        }
        const uint32_t off = sw128_offset(pr, c);
        // issue the Q load and every tap's P load before touching any of them
        Raw8<T> rq, rp[10];
        unsigned okp = 0;
        if (mv) rq.load(Q + m * CT_C + c * 8); else rq.zero();
        const int ih0 = oh * g.stride - g.pad, iw0 = ow * g.stride - g.pad;
#pragma unroll
        for (int tl = 0; tl < 10; ++tl) {
          if (tl < ntaps) {
            const int tap = tap_base + tl;
            const int ih = ih0 + tap / g.ksize, iw = iw0 + tap % g.ksize;
            const bool ok = mv && ih >= 0 && ih < g.p_h && iw >= 0 && iw < g.p_w;
            if (ok) {
              rp[tl].load(P + (((long long)n * g.p_h + ih) * g.p_w + iw) * CT_C + c * 8);
              okp |= 1u << tl;
            } else {
              rp[tl].zero();
            }
          }
        }
        {
          float v[8], hi[8], lo[8];
          rq.unpack(v);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float x = mv ? pro_apply(proQ.mode, v[i], scQ[i], shQ[i]) : 0.f;
            colsum[i] += x;
            split_bf16(x, hi[i], lo[i]);
          }
          uint8_t* dst = stage + (uint32_t)NS * SLOT + off;
          *reinterpret_cast<uint4*>(dst) = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]),
                                                      pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
          if (SPLIT)
            *reinterpret_cast<uint4*>(dst + half_bytes) = make_uint4(
                pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
        }
#pragma unroll
        for (int tl = 0; tl < 10; ++tl) {
          if (tl < ntaps) {
            float v[8], hi[8], lo[8];
            rp[tl].unpack(v);
            const bool ok = (okp >> tl) & 1u;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float x = ok ? pro_apply(proP.mode, v[i], scP[i], shP[i]) : 0.f;
              split_bf16(x, hi[i], lo[i]);
            }
            uint8_t* dst = stage + (uint32_t)tl * SLOT + off;
            *reinterpret_cast<uint4*>(dst) = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]),
                                                        pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
            if (SPLIT)
              *reinterpret_cast<uint4*>(dst + half_bytes) = make_uint4(
                  pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
          }
        }
        if (ntaps < NS) {       // the odd slot out: keep it finite (its TMEM lanes are never read)
          uint8_t* dst = stage + (uint32_t)ntaps * SLOT + off;
          *reinterpret_cast<uint4*>(dst) = make_uint4(0u, 0u, 0u, 0u);
          if (SPLIT) *reinterpret_cast<uint4*>(dst + half_bytes) = make_uint4(0u, 0u, 0u, 0u);
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(&full[s]);
    }
    if (q_colsum && blockIdx.y == 0) {
#pragma unroll
      for (int i = 0; i < 8; ++i) atomicAdd(&sCol[c * 8 + i], colsum[i]);
    }

    // ------------------------------------------------------------------ epilogue: thread = (slot parity, channel a)
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int a = tid & 63, sp = tid >> 6;
    for (int j = 0; j < NS / 2; ++j) {
      const int tl = 2 * j + sp;
#pragma unroll 1
      for (int chunk = 0; chunk < 2; ++chunk) {
        uint32_t acc[32];
        tmem_ld32(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(j * CT_C + chunk * 32), acc);
        tmem_ld_wait();
        if (tl < ntaps) {
          float* dst = dw + (tap_base + tl) + (long long)a * g.w_a_stride + (long long)(chunk * 32) * g.w_b_stride;
#pragma unroll
          for (int i = 0; i < 32; ++i) atomicAdd(dst + (long long)i * g.w_b_stride, __uint_as_float(acc[i]));
        }
      }
    }
    if (q_colsum && blockIdx.y == 0) {
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (tid < CT_C) atomicAdd(q_colsum + tid, sCol[tid]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc<TMEM_COLS>(tmem_base);
}

template <typename T, bool SPLIT, int TMEM_COLS>
static int launch_wgrad_tc(const WgradTcGeom& g, unsigned gx, unsigned gy, size_t smem, const void* p, const void* q,
                           float* dw, float* q_colsum, const Pro& pp, const Pro& pq, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    cudaError_t err = cudaFuncSetAttribute(wgrad_tc_kernel<T, SPLIT, TMEM_COLS>,
                                           cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(wgrad_tc_kernel)");
    configured = true;
  }
  wgrad_tc_kernel<T, SPLIT, TMEM_COLS><<<dim3(gx, gy), 160, smem, stream>>>((const T*)p, (const T*)q, dw, q_colsum,
                                                                            pp, pq, g);
  NSG_CHECK_LAUNCH("wgrad_tc_kernel");
  return NSG_OK;
}

template <typename T, bool SPLIT>
static int dispatch_wgrad_cols(int cols, const WgradTcGeom& g, unsigned gx, unsigned gy, size_t smem, const void* p,
                               const void* q, float* dw, float* q_colsum, const Pro& pp, const Pro& pq,
                               cudaStream_t stream) {
  if (cols <= 64) return launch_wgrad_tc<T, SPLIT, 64>(g, gx, gy, smem, p, q, dw, q_colsum, pp, pq, stream);
  if (cols <= 256) return launch_wgrad_tc<T, SPLIT, 256>(g, gx, gy, smem, p, q, dw, q_colsum, pp, pq, stream);
  return launch_wgrad_tc<T, SPLIT, 512>(g, gx, gy, smem, p, q, dw, q_colsum, pp, pq, stream);
}

int conv_wgrad_tc(const nsg_conv_geom* cg, const void* p, const nsg_prologue* pro_p, const void* q,
                  const nsg_prologue* pro_q, float* dw, float* q_colsum, int passes, cudaStream_t stream) {
  if (int rc = tc_shape_ok(cg, "conv_wgrad_tc")) return rc;
  NSG_REQUIRE(passes >= 1 && passes <= 3, "conv_wgrad_tc: passes must be 1, 2 or 3");
  const bool split = passes >= 2;
  const int kp = split ? 32 : 64;
  WgradTcGeom g{};
  g.batch = cg->batch; g.p_h = cg->in_h; g.p_w = cg->in_w; g.q_h = cg->out_h; g.q_w = cg->out_w;
  g.ksize = cg->ksize; g.stride = cg->stride; g.pad = cg->pad;
  g.w_a_stride = cg->w_cin_stride; g.w_b_stride = cg->w_cout_stride;
  const int taps = cg->ksize * cg->ksize;
  g.tap0 = 0; g.ntaps = taps;
  g.taps_per_group = taps <= 10 ? taps : 8;
  g.nslots = (g.taps_per_group + 1) / 2 * 2;
  const unsigned groups = (unsigned)((taps + g.taps_per_group - 1) / g.taps_per_group);
  const long long M = (long long)g.batch * g.q_h * g.q_w;
  if (M == 0) return NSG_OK;
  long long nctas = num_sms() / (long long)groups;
  if (nctas < 1) nctas = 1;
  long long ppc = (M + nctas - 1) / nctas;
  ppc = (ppc + kp - 1) / kp * kp;
  if (ppc < 2 * kp) ppc = 2 * kp;
  g.pix_per_cta = (int)ppc;
  const unsigned gx = (unsigned)((M + ppc - 1) / ppc);
  const size_t slot = (size_t)kp * 128;
  const size_t smem = 1024 + 2 * (size_t)(g.nslots + 1) * slot * (split ? 2 : 1) + 512;
  Pro pp{NSG_PRO_NONE, nullptr, nullptr}, pq{NSG_PRO_NONE, nullptr, nullptr};
  if (pro_p) pp = Pro{pro_p->mode, pro_p->scale, pro_p->shift};
  if (pro_q) pq = Pro{pro_q->mode, pro_q->scale, pro_q->shift};
  if (pp.mode >= NSG_PRO_AFFINE_RELU) NSG_REQUIRE(pp.scale && pp.shift, "wgrad: affine prologue needs scale/shift");
  if (pq.mode >= NSG_PRO_AFFINE_RELU) NSG_REQUIRE(pq.scale && pq.shift, "wgrad: affine prologue needs scale/shift");
  const int cols = g.nslots / 2 * CT_C;
  if (cg->dtype == NSG_F32) {
    return split ? dispatch_wgrad_cols<float, true>(cols, g, gx, groups, smem, p, q, dw, q_colsum, pp, pq, stream)
                 : dispatch_wgrad_cols<float, false>(cols, g, gx, groups, smem, p, q, dw, q_colsum, pp, pq, stream);
  } else if (cg->dtype == NSG_BF16) {
    return split ? dispatch_wgrad_cols<__nv_bfloat16, true>(cols, g, gx, groups, smem, p, q, dw, q_colsum, pp, pq, stream)
                 : dispatch_wgrad_cols<__nv_bfloat16, false>(cols, g, gx, groups, smem, p, q, dw, q_colsum, pp, pq, stream);
  }
  set_error("unknown dtype %d", cg->dtype);
  return NSG_E_INVALID;
}

}  // namespace nsg

namespace nsg {
int conv_tc_hang_init(unsigned int* host_mapped_words) {
  cudaError_t e = cudaMemcpyToSymbol(tc::g_hang_words, &host_mapped_words, sizeof(host_mapped_words));
  return e == cudaSuccess ? NSG_OK : cuda_fail(e, "cudaMemcpyToSymbol(g_hang_words)");
}
}  // namespace nsg
