// The two single-channel ends of the VQ-VAE (models.py:165 Conv2d(1 -> 64, 4, 2, 1) and models.py:182
// ConvTranspose2d(64 -> 1, 4, 2, 1)) and their gradients.  With one channel on one side these are not GEMMs
// worth a tensor core (K = 16 or N = 1): they are HBM-bound streaming kernels over the 64-channel tensor, so
// each gets a direct CUDA-core kernel with 16-byte accesses along the channel axis:
//   c1_conv    out[m, 0..63] = bias + sum_t in1[gather(m, t)] * w[t, 0..63]     (+ mask / BN-statistic epilogues)
//              -- encoder.0 forward and decoder.6 data gradient
//   c1_convT   out1[m] = [tanh](bias + sum_{valid t} sum_c pro(in[gatherT(m, t), c]) * w[c, t])
//              -- decoder.6 forward (and encoder.0 data gradient, unused by the train step)
//   c1_wgrad   dw[t, b] += sum_m p1[gather(m, t)] * pro(q[m, b]);  colsum[b] += sum_m pro(q[m, b])
//              -- encoder.0 and decoder.6 weight gradients
#include "common.cuh"
#include "conv_geom.cuh"

namespace nsg {

constexpr int C1_C = 64;

template <typename T>
__device__ __forceinline__ void c1_ld8(const T* p, float (&v)[8]) {
  const float4 a = Act<T>::ld4(p), b = Act<T>::ld4(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <typename T>
__device__ __forceinline__ void c1_st8(T* p, const float (&v)[8]) {
  Act<T>::st4(p, make_float4(v[0], v[1], v[2], v[3]));
  Act<T>::st4(p + 4, make_float4(v[4], v[5], v[6], v[7]));
}

struct C1Geom {
  int batch, in_h, in_w, out_h, out_w;
  int ksize, stride, pad;
  long long w_tap_stride_c;   // elements between consecutive 64-side channels in w (tap index is the fastest axis)
};

// ---------------------------------------------------------------------------------------------- 1 -> 64, conv-like gather
template <typename T>
__global__ void __launch_bounds__(256)
c1_conv_kernel(const T* __restrict__ in, const float* __restrict__ w, T* __restrict__ out, EpiArgs epi, C1Geom g) {
  __shared__ __align__(16) float sW[16][C1_C];
  __shared__ float sStat[2][C1_C];
  const int tid = threadIdx.x;
  const int ntaps = g.ksize * g.ksize;
  for (int i = tid; i < 16 * C1_C; i += 256) {
    const int t = i / C1_C, co = i % C1_C;
    sW[t][co] = t < ntaps ? __ldg(w + t + (long long)co * g.w_tap_stride_c) : 0.f;
  }
  if (tid < 2 * C1_C) (&sStat[0][0])[tid] = 0.f;
  __syncthreads();
  const int c0 = (tid & 7) * 8;
  float bias[8], msc[8], msh[8], mu[8], istd[8], s1[8], s2[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    bias[i] = epi.bias ? __ldg(epi.bias + c0 + i) : 0.f;
    msc[i] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mscale + c0 + i) : 1.f;
    msh[i] = epi.mask_mode == NSG_MASK_AFFINE ? __ldg(epi.mshift + c0 + i) : 0.f;
    mu[i] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.mean + c0 + i) : 0.f;
    istd[i] = epi.stats_mode == NSG_STATS_BNBWD ? __ldg(epi.invstd + c0 + i) : 0.f;
    s1[i] = s2[i] = 0.f;
  }
  const T* add_src = static_cast<const T*>(epi.add_src);
  const T* mask_src = static_cast<const T*>(epi.mask_src);
  const T* xhat_src = static_cast<const T*>(epi.xhat_src);
  const long long M = (long long)g.batch * g.out_h * g.out_w;
  for (long long m = (long long)blockIdx.x * 32 + (tid >> 3); m < M; m += (long long)gridDim.x * 32) {
    const int ow = (int)(m % g.out_w);
    const long long t2 = m / g.out_w;
    const int oh = (int)(t2 % g.out_h), n = (int)(t2 / g.out_h);
    const int ih0 = oh * g.stride - g.pad, iw0 = ow * g.stride - g.pad;
    const T* base = in + (long long)n * g.in_h * g.in_w;
    float acc[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) acc[i] = bias[i];
#pragma unroll
    for (int kh = 0; kh < 4; ++kh) {
#pragma unroll
      for (int kw = 0; kw < 4; ++kw) {
        if (kh < g.ksize && kw < g.ksize) {
          const int ih = ih0 + kh, iw = iw0 + kw;
          float x = 0.f;
          if (ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w) x = Act<T>::ld(base + (long long)ih * g.in_w + iw);
          const int t = kh * g.ksize + kw;
          const float4 wa = *reinterpret_cast<const float4*>(&sW[t][c0]);
          const float4 wb = *reinterpret_cast<const float4*>(&sW[t][c0 + 4]);
          acc[0] = fmaf(x, wa.x, acc[0]); acc[1] = fmaf(x, wa.y, acc[1]);
          acc[2] = fmaf(x, wa.z, acc[2]); acc[3] = fmaf(x, wa.w, acc[3]);
          acc[4] = fmaf(x, wb.x, acc[4]); acc[5] = fmaf(x, wb.y, acc[5]);
          acc[6] = fmaf(x, wb.z, acc[6]); acc[7] = fmaf(x, wb.w, acc[7]);
        }
      }
    }
    const long long o = m * C1_C + c0;
    if (add_src) {
      float a[8];
      c1_ld8<T>(add_src + o, a);
#pragma unroll
      for (int i = 0; i < 8; ++i) acc[i] += a[i];
    }
    if (epi.mask_mode != NSG_MASK_NONE) {
      float mk[8];
      c1_ld8<T>(mask_src + o, mk);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float mv = epi.mask_mode == NSG_MASK_AFFINE ? fmaf(msc[i], mk[i], msh[i]) : mk[i];
        acc[i] = mv > 0.f ? acc[i] : 0.f;
      }
    }
    c1_st8<T>(out + o, acc);
    if (epi.stats_mode == NSG_STATS_MOMENTS) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s1[i] += acc[i];
        s2[i] = fmaf(acc[i], acc[i], s2[i]);
      }
    } else if (epi.stats_mode == NSG_STATS_BNBWD) {
      float hx[8];
      c1_ld8<T>(xhat_src + o, hx);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        s1[i] += acc[i];
        s2[i] = fmaf(acc[i], (hx[i] - mu[i]) * istd[i], s2[i]);
      }
    }
  }
  if (epi.stats_mode != NSG_STATS_NONE) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {      // lanes l, l+8, l+16, l+24 hold the same channels
      s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], 8);
      s1[i] += __shfl_xor_sync(0xffffffffu, s1[i], 16);
      s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], 8);
      s2[i] += __shfl_xor_sync(0xffffffffu, s2[i], 16);
    }
    if ((tid & 31) < 8) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        atomicAdd(&sStat[0][c0 + i], s1[i]);
        atomicAdd(&sStat[1][c0 + i], s2[i]);
      }
    }
    __syncthreads();
    if (tid < 2 * C1_C) atomicAdd(epi.stats + tid, (double)(&sStat[0][0])[tid]);
  }
}

// ---------------------------------------------------------------------------------------------- 64 -> 1, transposed-like gather
// thread = output pixel; tap (kh, kw) contributes when (oh + pad - kh) and (ow + pad - kw) are multiples of stride
template <typename T>
__global__ void __launch_bounds__(256)
c1_convT_kernel(const T* __restrict__ in, const float* __restrict__ w, T* __restrict__ out, Pro pro,
                const float* __restrict__ bias, int tanh_out, C1Geom g) {
  __shared__ __align__(16) float sW[16][C1_C + 4];       // +4: taps of different parity hit different banks
  __shared__ __align__(16) float sSc[C1_C], sSh[C1_C];
  const int tid = threadIdx.x;
  const int ntaps = g.ksize * g.ksize;
  for (int i = tid; i < 16 * C1_C; i += 256) {
    const int t = i / C1_C, c = i % C1_C;
    sW[t][c] = t < ntaps ? __ldg(w + t + (long long)c * g.w_tap_stride_c) : 0.f;
  }
  if (tid < C1_C) {
    sSc[tid] = pro.mode >= NSG_PRO_AFFINE_RELU ? __ldg(pro.scale + tid) : 1.f;
    sSh[tid] = pro.mode >= NSG_PRO_AFFINE_RELU ? __ldg(pro.shift + tid) : 0.f;
  }
  __syncthreads();
  const float b0 = bias ? __ldg(bias) : 0.f;
  const long long M = (long long)g.batch * g.out_h * g.out_w;
  for (long long m = (long long)blockIdx.x * 256 + tid; m < M; m += (long long)gridDim.x * 256) {
    const int ow = (int)(m % g.out_w);
    const long long t2 = m / g.out_w;
    const int oh = (int)(t2 % g.out_h), n = (int)(t2 / g.out_h);
    const T* base = in + (long long)n * g.in_h * g.in_w * C1_C;
    float acc = b0;
    for (int kh = 0; kh < g.ksize; ++kh) {
      const int nh = oh + g.pad - kh;
      if (nh < 0 || nh % g.stride != 0) continue;
      const int ih = nh / g.stride;
      if (ih >= g.in_h) continue;
      for (int kw = 0; kw < g.ksize; ++kw) {
        const int nw = ow + g.pad - kw;
        if (nw < 0 || nw % g.stride != 0) continue;
        const int iw = nw / g.stride;
        if (iw >= g.in_w) continue;
        const T* px = base + ((long long)ih * g.in_w + iw) * C1_C;
        const float* wt = sW[kh * g.ksize + kw];
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          float v[8];
          c1_ld8<T>(px + q * 8, v);
          const float4 wa = *reinterpret_cast<const float4*>(wt + q * 8);
          const float4 wb = *reinterpret_cast<const float4*>(wt + q * 8 + 4);
          const float4 sa = *reinterpret_cast<const float4*>(sSc + q * 8), sb = *reinterpret_cast<const float4*>(sSc + q * 8 + 4);
          const float4 ha = *reinterpret_cast<const float4*>(sSh + q * 8), hb = *reinterpret_cast<const float4*>(sSh + q * 8 + 4);
          acc = fmaf(pro_apply(pro.mode, v[0], sa.x, ha.x), wa.x, acc);
          acc = fmaf(pro_apply(pro.mode, v[1], sa.y, ha.y), wa.y, acc);
          acc = fmaf(pro_apply(pro.mode, v[2], sa.z, ha.z), wa.z, acc);
          acc = fmaf(pro_apply(pro.mode, v[3], sa.w, ha.w), wa.w, acc);
          acc = fmaf(pro_apply(pro.mode, v[4], sb.x, hb.x), wb.x, acc);
          acc = fmaf(pro_apply(pro.mode, v[5], sb.y, hb.y), wb.y, acc);
          acc = fmaf(pro_apply(pro.mode, v[6], sb.z, hb.z), wb.z, acc);
          acc = fmaf(pro_apply(pro.mode, v[7], sb.w, hb.w), wb.w, acc);
        }
      }
    }
    Act<T>::st(out + m, tanh_out ? tanhf(acc) : acc);
  }
}

// ---------------------------------------------------------------------------------------------- weight gradient, one side single-channel
// thread = (pixel lane, 8 channels of Q): 16 taps x 8 channels of accumulators in registers over a grid-stride
// pixel loop; reduced across the 4 pixel lanes of a warp by shuffles, across warps in shared memory, across CTAs
// with float atomics.
template <typename T>
__global__ void __launch_bounds__(256)
c1_wgrad_kernel(const T* __restrict__ P, const T* __restrict__ Q, float* __restrict__ dw,
                float* __restrict__ q_colsum, Pro proQ, C1Geom g) {
  __shared__ float sAcc[17][C1_C];
  const int tid = threadIdx.x;
  for (int i = tid; i < 17 * C1_C; i += 256) (&sAcc[0][0])[i] = 0.f;
  __syncthreads();
  const int c0 = (tid & 7) * 8;
  float sc[8], sh[8], acc[16][8], col[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sc[i] = proQ.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proQ.scale + c0 + i) : 1.f;
    sh[i] = proQ.mode >= NSG_PRO_AFFINE_RELU ? __ldg(proQ.shift + c0 + i) : 0.f;
    col[i] = 0.f;
#pragma unroll
    for (int t = 0; t < 16; ++t) acc[t][i] = 0.f;
  }
  // geometry: out_* = Q's pixel grid, in_* = P's pixel grid
  const long long M = (long long)g.batch * g.out_h * g.out_w;
  for (long long m = (long long)blockIdx.x * 32 + (tid >> 3); m < M; m += (long long)gridDim.x * 32) {
    const int ow = (int)(m % g.out_w);
    const long long t2 = m / g.out_w;
    const int oh = (int)(t2 % g.out_h), n = (int)(t2 / g.out_h);
    const int ih0 = oh * g.stride - g.pad, iw0 = ow * g.stride - g.pad;
    const T* base = P + (long long)n * g.in_h * g.in_w;
    float q[8];
    c1_ld8<T>(Q + m * C1_C + c0, q);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      q[i] = pro_apply(proQ.mode, q[i], sc[i], sh[i]);
      col[i] += q[i];
    }
#pragma unroll
    for (int kh = 0; kh < 4; ++kh) {
#pragma unroll
      for (int kw = 0; kw < 4; ++kw) {
        if (kh < g.ksize && kw < g.ksize) {
          const int ih = ih0 + kh, iw = iw0 + kw;
          float x = 0.f;
          if (ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w) x = Act<T>::ld(base + (long long)ih * g.in_w + iw);
#pragma unroll
          for (int i = 0; i < 8; ++i) acc[kh * 4 + kw][i] = fmaf(x, q[i], acc[kh * 4 + kw][i]);
        }
      }
    }
  }
  const bool lead = (tid & 31) < 8;
#pragma unroll
  for (int t = 0; t < 17; ++t) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float v = t < 16 ? acc[t & 15][i] : col[i];
      v += __shfl_xor_sync(0xffffffffu, v, 8);
      v += __shfl_xor_sync(0xffffffffu, v, 16);
      if (lead) atomicAdd(&sAcc[t][c0 + i], v);
    }
  }
  __syncthreads();
  for (int i = tid; i < 16 * C1_C; i += 256) {
    const int slot = i / C1_C, b = i % C1_C;          // slot = kh*4 + kw
    const int kh = slot >> 2, kw = slot & 3;
    if (kh < g.ksize && kw < g.ksize)
      atomicAdd(dw + (kh * g.ksize + kw) + (long long)b * g.w_tap_stride_c, sAcc[slot][b]);
  }
  if (q_colsum && tid < C1_C) atomicAdd(q_colsum + tid, sAcc[16][tid]);
}

// ---------------------------------------------------------------------------------------------- dispatch
// 1 when the call was served here, 0 when the shape is not one of the single-channel layers, < 0 on error
int conv_gemm_c1_fast(const nsg_conv_geom* cg, const void* in, const float* w, const Pro& pro, const EpiArgs& e,
                      void* out, cudaStream_t stream);
int conv_wgrad_c1_fast(const nsg_conv_geom* cg, const void* p, const Pro& pp, const void* q, const Pro& pq, float* dw,
                       float* q_colsum, cudaStream_t stream);

int conv_gemm_c1(const nsg_conv_geom* cg, const void* in, const float* w, const Pro& pro, const EpiArgs& e, void* out,
                 cudaStream_t stream) {
  if (cg->ksize > 4) return 0;
  if (const int served = conv_gemm_c1_fast(cg, in, w, pro, e, out, stream)) return served;
  C1Geom g{cg->batch, cg->in_h, cg->in_w, cg->out_h, cg->out_w, cg->ksize, cg->stride, cg->pad, 0};
  const long long M = (long long)cg->batch * cg->out_h * cg->out_w;
  if (cg->gather == 0 && cg->cin == 1 && cg->cout == C1_C && pro.mode == NSG_PRO_NONE && !e.tanh_out) {
    g.w_tap_stride_c = cg->w_cout_stride;
    long long blocks = (M + 31) / 32;
    if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
    return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
      c1_conv_kernel<T><<<(unsigned)blocks, 256, 0, stream>>>((const T*)in, w, (T*)out, e, g);
      NSG_CHECK_LAUNCH("c1_conv_kernel");
      return 1;
    });
  }
  if (cg->gather == 1 && cg->cin == C1_C && cg->cout == 1 && !e.add_src && e.mask_mode == NSG_MASK_NONE &&
      e.stats_mode == NSG_STATS_NONE) {
    g.w_tap_stride_c = cg->w_cin_stride;
    long long blocks = (M + 255) / 256;
    if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
    return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
      c1_convT_kernel<T><<<(unsigned)blocks, 256, 0, stream>>>((const T*)in, w, (T*)out, pro, e.bias, e.tanh_out, g);
      NSG_CHECK_LAUNCH("c1_convT_kernel");
      return 1;
    });
  }
  return 0;
}

int conv_wgrad_c1(const nsg_conv_geom* cg, const void* p, const Pro& pp, const void* q, const Pro& pq, float* dw,
                  float* q_colsum, cudaStream_t stream) {
  if (!(cg->cin == 1 && cg->cout == C1_C && cg->ksize <= 4 && pp.mode == NSG_PRO_NONE)) return 0;
  if (const int served = conv_wgrad_c1_fast(cg, p, pp, q, pq, dw, q_colsum, stream)) return served;
  C1Geom g{cg->batch, cg->in_h, cg->in_w, cg->out_h, cg->out_w, cg->ksize, cg->stride, cg->pad, cg->w_cout_stride};
  const long long M = (long long)cg->batch * cg->out_h * cg->out_w;
  long long blocks = (M + 31) / 32;
  if (blocks > (long long)num_sms() * 2) blocks = (long long)num_sms() * 2;
  return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
    c1_wgrad_kernel<T><<<(unsigned)blocks, 256, 0, stream>>>((const T*)p, (const T*)q, dw, q_colsum, pq, g);
    NSG_CHECK_LAUNCH("c1_wgrad_kernel");
    return 1;
  });
}

}  // namespace nsg
