// Streaming kernels for the two single-channel ends of the VQ-VAE at their actual geometry (kernel 4, stride 2,
// pad 1, 64 channels on the wide side): models.py:165 Conv2d(1 -> 64) and models.py:182 ConvTranspose2d(64 -> 1),
// with their data and weight gradients.  They are HBM-bound over the 64-channel tensor (one read or one write of
// it is the whole algorithmic traffic), so the design goal is one pass over that tensor with 8/16-byte accesses:
//   c1f_conv   (encoder.0 forward, decoder.6 data gradient): persistent CTAs over output rows; the four input rows a
//              row needs are staged once in shared memory; thread = (4 channels, 1 of 16 pixel lanes) with its
//              16 taps x 4 channels of weights in registers; mask / BatchNorm-statistic epilogues as in conv_tma.
//   c1f_wgrad  (encoder.0 and decoder.6 weight gradients): same staging; 16 taps x 4 channels of accumulators per
//              thread over all rows of the CTA, one shuffle + shared + global reduction at the end.
//   c1f_convT  (decoder.6 forward): per block of input rows, every input pixel's 16 tap products
//              p[t] = sum_c v[c] * w[c][t] are formed ONCE (the 64-channel tensor is read once) into shared memory;
//              each output pixel then sums its (up to) four contributions, adds the bias and applies tanh.
// Other geometries fall back to the generic kernels in conv_c1.cu.
#include "common.cuh"
#include "conv_geom.cuh"

namespace nsg {

constexpr int CF_C = 64;

template <typename T>
__device__ __forceinline__ void cf_ld4(const T* p, float (&v)[4]) {
  const float4 a = Act<T>::ld4(p);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w;
}
template <typename T>
__device__ __forceinline__ void cf_st4(T* p, const float (&v)[4]) {
  Act<T>::st4(p, make_float4(v[0], v[1], v[2], v[3]));
}

struct CfGeom {
  int batch, in_h, in_w, out_h, out_w;   // single-channel grid (in) and 64-channel grid (out) for conv / wgrad
  long long w_tap_stride_c;              // elements between consecutive 64-side channels in w (taps are fastest)
};

// Input rows 2*oh-1 .. 2*oh+2 of image n (single channel) are staged as floats with one zero column on each side.
// The loads of the NEXT row of the CTA are issued into registers before the current row is computed.
// (Values stay RAW in registers until they are stored: converting a loaded value would stall the in-order warp on the load.)
__device__ __forceinline__ float cf_to_float(float v) { return v; }
__device__ __forceinline__ float cf_to_float(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T, int NREG>
__device__ __forceinline__ void cf_fetch_rows(const T* __restrict__ in, T (&r)[NREG], int n, int oh, const CfGeom& g) {
  const int pitch = g.in_w + 2;
#pragma unroll
  for (int k = 0; k < NREG; ++k) {
    const int i = threadIdx.x + k * 256;
    T v = T(0.f);
    if (i < 4 * pitch) {
      const int rr = i / pitch, c = i - rr * pitch - 1;
      const int ih = 2 * oh - 1 + rr;
      if (ih >= 0 && ih < g.in_h && c >= 0 && c < g.in_w) v = __ldg(in + ((long long)n * g.in_h + ih) * g.in_w + c);
    }
    r[k] = v;
  }
}
template <typename T, int NREG>
__device__ __forceinline__ void cf_store_rows(float* sIn, const T (&r)[NREG], int pitch) {
#pragma unroll
  for (int k = 0; k < NREG; ++k) {
    const int i = threadIdx.x + k * 256;
    if (i < 4 * pitch) sIn[i] = cf_to_float(r[k]);
  }
}

// four consecutive channels of one pixel, kept as loaded
template <typename T>
struct CfRaw4;
template <>
struct CfRaw4<float> {
  float4 v;
  __device__ __forceinline__ void load(const float* p) { v = __ldg(reinterpret_cast<const float4*>(p)); }
  __device__ __forceinline__ void get(float (&o)[4]) const { o[0] = v.x; o[1] = v.y; o[2] = v.z; o[3] = v.w; }
};
template <>
struct CfRaw4<__nv_bfloat16> {
  uint2 v;
  __device__ __forceinline__ void load(const __nv_bfloat16* p) { v = __ldg(reinterpret_cast<const uint2*>(p)); }
  __device__ __forceinline__ void get(float (&o)[4]) const {
    o[0] = __uint_as_float(v.x << 16); o[1] = __uint_as_float(v.x & 0xffff0000u);
    o[2] = __uint_as_float(v.y << 16); o[3] = __uint_as_float(v.y & 0xffff0000u);
  }
};

// ---------------------------------------------------------------------------------------------- 1 -> 64 forward / dgrad
template <typename T, int NREG>
__global__ void __launch_bounds__(256)
c1f_conv_kernel(const T* __restrict__ in, const float* __restrict__ w, T* __restrict__ out, EpiArgs epi, CfGeom g) {
  extern __shared__ float cf_smem[];
  float* sIn = cf_smem;                       // [4][in_w + 2]
  __shared__ float sStat[2][CF_C];
  const int tid = threadIdx.x;
  const int c0 = (tid & 15) * 4, pg = tid >> 4;
  const int pitch = g.in_w + 2;
  float2 wr[16][2];
  float bias[4], msc[4], msh[4], mu[4], istd[4], s1[4], s2[4];
#pragma unroll
  for (int t = 0; t < 16; ++t) {
    wr[t][0] = make_float2(__ldg(w + t + (long long)(c0 + 0) * g.w_tap_stride_c), __ldg(w + t + (long long)(c0 + 1) * g.w_tap_stride_c));
    wr[t][1] = make_float2(__ldg(w + t + (long long)(c0 + 2) * g.w_tap_stride_c), __ldg(w + t + (long long)(c0 + 3) * g.w_tap_stride_c));
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    bias[j] = epi.bias ? __ldg(epi.bias + c0 + j) : 0.f;
    msc[j] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mscale + c0 + j) : 1.f;
    msh[j] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mshift + c0 + j) : 0.f;
    mu[j] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.mean + c0 + j) : 0.f;
    istd[j] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.invstd + c0 + j) : 0.f;
    s1[j] = s2[j] = 0.f;
  }
  if (tid < 2 * CF_C) (&sStat[0][0])[tid] = 0.f;
  const T* add_src = static_cast<const T*>(epi.add_src);
  const T* mask_src = static_cast<const T*>(epi.mask_src);
  const T* xhat_src = static_cast<const T*>(epi.xhat_src);
  const int rows = g.batch * g.out_h;
  T stage[NREG];
  if ((int)blockIdx.x < rows) cf_fetch_rows<T, NREG>(in, stage, blockIdx.x / g.out_h, blockIdx.x % g.out_h, g);
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    __syncthreads();
    cf_store_rows<T, NREG>(sIn, stage, pitch);
    __syncthreads();
    const int nxt = row + gridDim.x;
    if (nxt < rows) cf_fetch_rows<T, NREG>(in, stage, nxt / g.out_h, nxt % g.out_h, g);
    const long long rowbase = (long long)row * g.out_w * CF_C + c0;
    const bool use_add = add_src != nullptr, use_mask = epi.mask_mode != NSG_MASK_NONE,
               use_xhat = epi.stats_mode == NSG_STATS_BNBWD;
    CfRaw4<T> radd, rmask, rxhat, nadd, nmask, nxhat;
    if (pg < g.out_w) {
      const long long o = rowbase + (long long)pg * CF_C;
      if (use_add) radd.load(add_src + o);
      if (use_mask) rmask.load(mask_src + o);
      if (use_xhat) rxhat.load(xhat_src + o);
    }
    for (int ow = pg; ow < g.out_w; ow += 16) {
      if (ow + 16 < g.out_w) {          // next pixel's operands are in flight during this pixel's FMAs
        const long long o2 = rowbase + (long long)(ow + 16) * CF_C;
        if (use_add) nadd.load(add_src + o2);
        if (use_mask) nmask.load(mask_src + o2);
        if (use_xhat) nxhat.load(xhat_src + o2);
      }
      float2 a01 = make_float2(bias[0], bias[1]), a23 = make_float2(bias[2], bias[3]);
#pragma unroll
      for (int kh = 0; kh < 4; ++kh) {
#pragma unroll
        for (int kw = 0; kw < 4; ++kw) {
          const float x = sIn[kh * pitch + 2 * ow + kw];
          const float2 xx = make_float2(x, x);
          a01 = __ffma2_rn(xx, wr[kh * 4 + kw][0], a01);
          a23 = __ffma2_rn(xx, wr[kh * 4 + kw][1], a23);
        }
      }
      float acc[4] = {a01.x, a01.y, a23.x, a23.y};
      const long long o = rowbase + (long long)ow * CF_C;
      if (use_add) {
        float a[4];
        radd.get(a);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[j] += a[j];
      }
      if (use_mask) {
        float mk[4];
        rmask.get(mk);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float mv = epi.mask_mode == NSG_MASK_AFFINE ? fmaf(msc[j], mk[j], msh[j]) : mk[j];
          acc[j] = mv > 0.f ? acc[j] : 0.f;
        }
      }
      cf_st4<T>(out + o, acc);
      if (epi.stats_mode == NSG_STATS_MOMENTS) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          s1[j] += acc[j];
          s2[j] = fmaf(acc[j], acc[j], s2[j]);
        }
      } else if (use_xhat) {
        float hx[4];
        rxhat.get(hx);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          s1[j] += acc[j];
          s2[j] = fmaf(acc[j], (hx[j] - mu[j]) * istd[j], s2[j]);
        }
      }
      radd = nadd; rmask = nmask; rxhat = nxhat;
    }
  }
  if (epi.stats_mode != NSG_STATS_NONE) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {      // lanes l and l+16 hold the same channels
      s1[j] += __shfl_xor_sync(0xffffffffu, s1[j], 16);
      s2[j] += __shfl_xor_sync(0xffffffffu, s2[j], 16);
    }
    __syncthreads();
    if ((tid & 31) < 16) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        atomicAdd(&sStat[0][c0 + j], s1[j]);
        atomicAdd(&sStat[1][c0 + j], s2[j]);
      }
    }
    __syncthreads();
    if (tid < 2 * CF_C) atomicAdd(epi.stats + tid, (double)(&sStat[0][0])[tid]);
  }
}

// ---------------------------------------------------------------------------------------------- weight gradient
// dw[t + b*w_tap_stride_c] += sum_m p1[gather(m, t)] * pro(q[m, b]);  colsum[b] += sum_m pro(q[m, b])
template <typename T, int NREG>
__global__ void __launch_bounds__(256)
c1f_wgrad_kernel(const T* __restrict__ P, const T* __restrict__ Q, float* __restrict__ dw,
                 float* __restrict__ q_colsum, Pro proQ, CfGeom g) {
  extern __shared__ float cf_smem[];
  float* sIn = cf_smem;                       // [4][in_w + 2]
  __shared__ float sAcc[17][CF_C];
  const int tid = threadIdx.x;
  const int c0 = (tid & 15) * 4, pg = tid >> 4;
  const int pitch = g.in_w + 2;
  for (int i = tid; i < 17 * CF_C; i += 256) (&sAcc[0][0])[i] = 0.f;
  float sc[4], sh[4], col[4];
  float2 acc[16][2];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    sc[j] = proQ.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proQ.scale + c0 + j) : 1.f;
    sh[j] = proQ.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proQ.shift + c0 + j) : 0.f;
    col[j] = 0.f;
  }
#pragma unroll
  for (int t = 0; t < 16; ++t) acc[t][0] = acc[t][1] = make_float2(0.f, 0.f);
  const int rows = g.batch * g.out_h;
  T stage[NREG];
  if ((int)blockIdx.x < rows) cf_fetch_rows<T, NREG>(P, stage, blockIdx.x / g.out_h, blockIdx.x % g.out_h, g);
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    __syncthreads();
    cf_store_rows<T, NREG>(sIn, stage, pitch);
    __syncthreads();
    const int nxt = row + gridDim.x;
    if (nxt < rows) cf_fetch_rows<T, NREG>(P, stage, nxt / g.out_h, nxt % g.out_h, g);
    const T* qrow = Q + (long long)row * g.out_w * CF_C + c0;
    CfRaw4<T> rq, nq;
    if (pg < g.out_w) rq.load(qrow + (long long)pg * CF_C);
    for (int ow = pg; ow < g.out_w; ow += 16) {
      if (ow + 16 < g.out_w) nq.load(qrow + (long long)(ow + 16) * CF_C);
      float q[4];
      rq.get(q);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        q[j] = pro_apply(proQ.mode, q[j], sc[j], sh[j]);
        col[j] += q[j];
      }
      const float2 q01 = make_float2(q[0], q[1]), q23 = make_float2(q[2], q[3]);
#pragma unroll
      for (int kh = 0; kh < 4; ++kh) {
#pragma unroll
        for (int kw = 0; kw < 4; ++kw) {
          const float x = sIn[kh * pitch + 2 * ow + kw];
          const float2 xx = make_float2(x, x);
          acc[kh * 4 + kw][0] = __ffma2_rn(xx, q01, acc[kh * 4 + kw][0]);
          acc[kh * 4 + kw][1] = __ffma2_rn(xx, q23, acc[kh * 4 + kw][1]);
        }
      }
      rq = nq;
    }
  }
  const bool lead = (tid & 31) < 16;
#pragma unroll
  for (int t = 0; t < 17; ++t) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float v = t < 16 ? ((j & 1) ? ((j & 2) ? acc[t & 15][1].y : acc[t & 15][0].y)
                                  : ((j & 2) ? acc[t & 15][1].x : acc[t & 15][0].x))
                       : col[j];
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      if (lead) atomicAdd(&sAcc[t][c0 + j], v);
    }
  }
  __syncthreads();
  for (int i = tid; i < 16 * CF_C; i += 256) {
    const int t = i / CF_C, b = i % CF_C;
    atomicAdd(dw + t + (long long)b * g.w_tap_stride_c, sAcc[t][b]);
  }
  if (q_colsum && tid < CF_C) atomicAdd(q_colsum + tid, sAcc[16][tid]);
}

// ---------------------------------------------------------------------------------------------- 64 -> 1 transposed forward
struct CtGeom {
  int batch, in_h, in_w, out_h, out_w;   // 64-channel grid (in) and single-channel grid (out = 2 * in)
  int rows_per_cta;                      // input rows whose output rows one CTA produces
  int pstride;                           // floats between taps of one staged row (in_w padded, == 2 mod 8)
  long long w_tap_stride_c;
};

template <typename T>
__global__ void __launch_bounds__(256, 2)
c1f_convT_kernel(const T* __restrict__ in, const float* __restrict__ w, T* __restrict__ out, Pro pro,
                 const float* __restrict__ bias, int tanh_out, CtGeom g) {
  extern __shared__ float cf_smem[];
  float* sP = cf_smem;                                   // [rows_per_cta + 2][16][pstride]
  __shared__ __align__(16) float sSc[CF_C], sSh[CF_C];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid < CF_C) {
    sSc[tid] = pro.mode >= NSG_PRO_AFFINE_RELU ? __ldg(pro.scale + tid) : 1.f;
    sSh[tid] = pro.mode >= NSG_PRO_AFFINE_RELU ? __ldg(pro.shift + tid) : 0.f;
  }
  // phase A: tap products of input rows r0-1 .. r0+R.  thread = (4 taps, 16 channels, 1 of 2 pixel slots of the warp)
  const int tg = lane & 3, cq = (lane >> 2) & 3, ps = lane >> 4;
  float2 w2[16][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const float* wc = w + (long long)(cq * 16 + i) * g.w_tap_stride_c + tg * 4;
    w2[i][0] = make_float2(__ldg(wc), __ldg(wc + 1));
    w2[i][1] = make_float2(__ldg(wc + 2), __ldg(wc + 3));
  }
  __syncthreads();
  const int R = g.rows_per_cta;
  const int blocks_per_img = (g.in_h + R - 1) / R;
  const int n = blockIdx.x / blocks_per_img, r0 = (blockIdx.x % blocks_per_img) * R;
  const int total = (R + 2) * g.in_w;
  // one item ahead: the 16 channels of the next pixel are loaded (raw) before the current pixel's FMAs
  CfRaw4<T> cur[4], nxt[4];
  bool cur_ok, nxt_ok = false;
  auto fetch = [&](int base, CfRaw4<T> (&dst)[4]) -> bool {
    const int item = base + ps;
    const int lr = item / g.in_w, iw = item - lr * g.in_w;
    const int ih = r0 - 1 + lr;
    const bool ok = item < total && ih >= 0 && ih < g.in_h;
    if (ok) {
      const T* px = in + (((long long)n * g.in_h + ih) * g.in_w + iw) * CF_C + cq * 16;
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) dst[q4].load(px + q4 * 4);
    }
    return ok;
  };
  cur_ok = fetch(warp * 2, cur);
  for (int base = warp * 2; base < total; base += 16) {
    if (base + 16 < total) nxt_ok = fetch(base + 16, nxt);
    const int item = base + ps;
    const int lr = item / g.in_w, iw = item - lr * g.in_w;
    float2 p01 = make_float2(0.f, 0.f), p23 = make_float2(0.f, 0.f);
    if (cur_ok) {
      float v[16];
#pragma unroll
      for (int q4 = 0; q4 < 4; ++q4) {
        float t4[4];
        cur[q4].get(t4);
        v[q4 * 4 + 0] = t4[0]; v[q4 * 4 + 1] = t4[1]; v[q4 * 4 + 2] = t4[2]; v[q4 * 4 + 3] = t4[3];
      }
      if (pro.mode != NSG_PRO_NONE) {
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = pro_apply(pro.mode, v[i], sSc[cq * 16 + i], sSh[cq * 16 + i]);
      }
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        const float2 xx = make_float2(v[i], v[i]);
        p01 = __ffma2_rn(xx, w2[i][0], p01);
        p23 = __ffma2_rn(xx, w2[i][1], p23);
      }
    }
    float p4[4] = {p01.x, p01.y, p23.x, p23.y};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      p4[j] += __shfl_xor_sync(0xffffffffu, p4[j], 4);
      p4[j] += __shfl_xor_sync(0xffffffffu, p4[j], 8);
    }
    if (cq == 0 && item < total) {
      float* dst = sP + ((size_t)lr * 16 + tg * 4) * g.pstride + iw;
#pragma unroll
      for (int j = 0; j < 4; ++j) dst[(size_t)j * g.pstride] = p4[j];
    }
#pragma unroll
    for (int q4 = 0; q4 < 4; ++q4) cur[q4] = nxt[q4];
    cur_ok = nxt_ok;
  }
  __syncthreads();
  // phase B: output rows 2*r0 .. 2*(r0+R)-1; out(oh, ow) = bias + sum over kh == (oh+1) mod 2, kw == (ow+1) mod 2
  const float b0 = bias ? __ldg(bias) : 0.f;
  const int oh_end = min(2 * (r0 + R), g.out_h);
  const int npix = (oh_end - 2 * r0) * g.out_w;
  for (int i = tid; i < npix; i += 256) {
    const int oh = 2 * r0 + i / g.out_w, ow = i % g.out_w;
    float acc = b0;
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int kh = ((oh + 1) & 1) + 2 * a;
      const int ih = (oh + 1 - kh) >> 1;
      if (ih < 0 || ih >= g.in_h) continue;
      const int lr = ih - (r0 - 1);
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int kw = ((ow + 1) & 1) + 2 * b;
        const int iw = (ow + 1 - kw) >> 1;
        if (iw < 0 || iw >= g.in_w) continue;
        acc += sP[((size_t)lr * 16 + kh * 4 + kw) * g.pstride + iw];
      }
    }
    Act<T>::st(out + ((long long)n * g.out_h + oh) * g.out_w + ow, tanh_out ? tanhf(acc) : acc);
  }
}

// ---------------------------------------------------------------------------------------------- dispatch
// 1 when served here, 0 when the geometry is not (k4, s2, p1, 64 channels), < 0 on error
static bool cf_geom_ok(const nsg_conv_geom* cg) { return cg->ksize == 4 && cg->stride == 2 && cg->pad == 1; }

int conv_gemm_c1_fast(const nsg_conv_geom* cg, const void* in, const float* w, const Pro& pro, const EpiArgs& e,
                      void* out, cudaStream_t stream) {
  if (!cf_geom_ok(cg)) return 0;
  if (cg->gather == 0 && cg->cin == 1 && cg->cout == CF_C && pro.mode == NSG_PRO_NONE && !e.tanh_out &&
      cg->out_h == (cg->in_h + 2 - 4) / 2 + 1 && cg->out_w == (cg->in_w + 2 - 4) / 2 + 1) {
    CfGeom g{cg->batch, cg->in_h, cg->in_w, cg->out_h, cg->out_w, cg->w_cout_stride};
    const size_t smem = (size_t)4 * (cg->in_w + 2) * sizeof(float);
    if (4 * (cg->in_w + 2) > 16 * 256) return 0;
    int blocks = cg->batch * cg->out_h;
    if (blocks > num_sms() * 2) blocks = num_sms() * 2;
    const int need = (4 * (cg->in_w + 2) + 255) / 256;
    return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
      if (need <= 5)
        c1f_conv_kernel<T, 5><<<blocks, 256, smem, stream>>>((const T*)in, w, (T*)out, e, g);
      else
        c1f_conv_kernel<T, 16><<<blocks, 256, smem, stream>>>((const T*)in, w, (T*)out, e, g);
      NSG_CHECK_LAUNCH("c1f_conv_kernel");
      return 1;
    });
  }
  if (cg->gather == 1 && cg->cin == CF_C && cg->cout == 1 && !e.add_src && e.mask_mode == NSG_MASK_NONE &&
      e.stats_mode == NSG_STATS_NONE && cg->out_h == 2 * cg->in_h && cg->out_w == 2 * cg->in_w) {
    CtGeom g{};
    g.batch = cg->batch; g.in_h = cg->in_h; g.in_w = cg->in_w; g.out_h = cg->out_h; g.out_w = cg->out_w;
    g.w_tap_stride_c = cg->w_cin_stride;
    g.pstride = cg->in_w + ((2 - cg->in_w % 8) + 8) % 8;
    int R = 4;
    while (R > 1 && (size_t)(R + 2) * 16 * g.pstride * sizeof(float) > 96 * 1024) --R;
    const size_t smem = (size_t)(R + 2) * 16 * g.pstride * sizeof(float);
    if (smem > 200 * 1024) return 0;
    g.rows_per_cta = R;
    const int blocks = cg->batch * ((cg->in_h + R - 1) / R);
    return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
      static bool configured = false;
      if (!configured) {
        cudaError_t err = cudaFuncSetAttribute(c1f_convT_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        if (err != cudaSuccess) return cuda_fail(err, "cudaFuncSetAttribute(c1f_convT_kernel)");
        configured = true;
      }
      c1f_convT_kernel<T><<<blocks, 256, smem, stream>>>((const T*)in, w, (T*)out, pro, e.bias, e.tanh_out, g);
      NSG_CHECK_LAUNCH("c1f_convT_kernel");
      return 1;
    });
  }
  return 0;
}

int conv_wgrad_c1_fast(const nsg_conv_geom* cg, const void* p, const Pro& pp, const void* q, const Pro& pq, float* dw,
                       float* q_colsum, cudaStream_t stream) {
  if (!cf_geom_ok(cg)) return 0;
  if (!(cg->cin == 1 && cg->cout == CF_C && pp.mode == NSG_PRO_NONE && cg->out_h == (cg->in_h + 2 - 4) / 2 + 1 &&
        cg->out_w == (cg->in_w + 2 - 4) / 2 + 1))
    return 0;
  CfGeom g{cg->batch, cg->in_h, cg->in_w, cg->out_h, cg->out_w, cg->w_cout_stride};
  const size_t smem = (size_t)4 * (cg->in_w + 2) * sizeof(float);
  if (4 * (cg->in_w + 2) > 16 * 256) return 0;
  int blocks = cg->batch * cg->out_h;
  if (blocks > num_sms() * 2) blocks = num_sms() * 2;
  const int need = (4 * (cg->in_w + 2) + 255) / 256;
  return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
    if (need <= 5)
      c1f_wgrad_kernel<T, 5><<<blocks, 256, smem, stream>>>((const T*)p, (const T*)q, dw, q_colsum, pq, g);
    else
      c1f_wgrad_kernel<T, 16><<<blocks, 256, smem, stream>>>((const T*)p, (const T*)q, dw, q_colsum, pq, g);
    NSG_CHECK_LAUNCH("c1f_wgrad_kernel");
    return 1;
  });
}

}  // namespace nsg
