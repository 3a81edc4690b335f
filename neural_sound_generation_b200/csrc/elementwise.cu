// HBM-bound helpers of the VQ-VAE step: BatchNorm finalize / apply / backward, ReLU backward,
// bias gradients, reconstruction loss, fused Adam.  All vectorised over 4 channels when C % 4 == 0.
#include "common.cuh"

namespace nsg {

static inline int grid_for(long long work, int threads, int per_sm = 8) {
  long long blocks = (work + threads - 1) / threads;
  long long cap = (long long)num_sms() * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

// ---------------------------------------------------------------- BN finalize
__global__ void bn_finalize_kernel(const double* __restrict__ stats, int C, double count,
                                   const float* __restrict__ gamma, const float* __restrict__ beta, float eps,
                                   float momentum, float* running_mean, float* running_var, long long* nbt,
                                   float* mean, float* invstd, float* scale, float* shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c == 0 && nbt) *nbt += 1;
  if (c >= C) return;
  const double m = stats[c] / count;
  double var = stats[C + c] / count - m * m;
  if (var < 0.0) var = 0.0;
  const float is = (float)(1.0 / sqrt(var + (double)eps));
  const float sc = gamma[c] * is;
  mean[c] = (float)m;
  invstd[c] = is;
  scale[c] = sc;
  shift[c] = beta[c] - (float)m * sc;
  if (running_mean) {
    const double unbiased = count > 1.0 ? var * count / (count - 1.0) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * (float)m;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * (float)unbiased;
  }
}

__global__ void bn_eval_coeffs_kernel(int C, const float* gamma, const float* beta, const float* rm,
                                      const float* rv, float eps, float* scale, float* shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float is = 1.0f / sqrtf(rv[c] + eps);
  const float sc = gamma[c] * is;
  scale[c] = sc;
  shift[c] = beta[c] - rm[c] * sc;
}

// ---------------------------------------------------------------- BN apply (+ residual)
template <typename T, int VW>
__global__ void bn_apply_kernel(const T* __restrict__ h, const float* __restrict__ scale,
                                const float* __restrict__ shift, const T* __restrict__ res, int res_relu,
                                int relu_out, T* __restrict__ y, long long total_vec, int C) {
  const int cq = C / VW;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cq) * VW;
    const long long o = i * VW;
    float v[4], r[4] = {0.f, 0.f, 0.f, 0.f};
    if (VW == 4) {
      float4 t = Act<T>::ld4(h + o);
      v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
      if (res) {
        float4 u = Act<T>::ld4(res + o);
        r[0] = u.x; r[1] = u.y; r[2] = u.z; r[3] = u.w;
      }
    } else {
      v[0] = Act<T>::ld(h + o);
      if (res) r[0] = Act<T>::ld(res + o);
    }
#pragma unroll
    for (int e = 0; e < VW; ++e) {
      float x = fmaf(__ldg(scale + c + e), v[e], __ldg(shift + c + e));
      if (relu_out) x = fmaxf(x, 0.f);
      if (res) x += res_relu ? fmaxf(r[e], 0.f) : r[e];
      v[e] = x;
    }
    if (VW == 4)
      Act<T>::st4(y + o, make_float4(v[0], v[1], v[2], v[3]));
    else
      Act<T>::st(y + o, v[0]);
  }
}

// ---------------------------------------------------------------- per-channel reductions over pixels
// MODE 0: colsum(g)                 -> out_f[c]  (float atomics)
// MODE 1: sum g', sum g'*xhat       -> stats[c], stats[C+c]  (double atomics), g' = relu-masked g
template <typename T, int VW, int MODE>
__global__ void __launch_bounds__(256)
chan_reduce_kernel(const T* __restrict__ g, const T* __restrict__ h, const float* __restrict__ mean,
                   const float* __restrict__ invstd, const float* __restrict__ scale,
                   const float* __restrict__ shift, int relu_mask, double* __restrict__ stats,
                   float* __restrict__ out_f, long long pixels, int C) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ float red[];  // [2][lanes][C]
  const int cq = C / VW;
  const int lanes = 256 / cq;
  const int tid = threadIdx.x;
  const int lane = tid / cq, q = tid % cq, c = q * VW;
  float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
  if (lane < lanes) {
    float mu[4], is[4], sc[4], sh[4];
#pragma unroll
    for (int e = 0; e < VW; ++e) {
      mu[e] = MODE == 1 ? __ldg(mean + c + e) : 0.f;
      is[e] = MODE == 1 ? __ldg(invstd + c + e) : 0.f;
      sc[e] = (MODE == 1 && relu_mask) ? __ldg(scale + c + e) : 0.f;
      sh[e] = (MODE == 1 && relu_mask) ? __ldg(shift + c + e) : 0.f;
    }
    for (long long p = (long long)blockIdx.x * lanes + lane; p < pixels; p += (long long)gridDim.x * lanes) {
      const long long o = p * C + c;
      float gv[4], hv[4] = {0.f, 0.f, 0.f, 0.f};
      if (VW == 4) {
        float4 t = Act<T>::ld4(g + o);
        gv[0] = t.x; gv[1] = t.y; gv[2] = t.z; gv[3] = t.w;
        if (MODE == 1) {
          float4 u = Act<T>::ld4(h + o);
          hv[0] = u.x; hv[1] = u.y; hv[2] = u.z; hv[3] = u.w;
        }
      } else {
        gv[0] = Act<T>::ld(g + o);
        if (MODE == 1) hv[0] = Act<T>::ld(h + o);
      }
#pragma unroll
      for (int e = 0; e < VW; ++e) {
        float x = gv[e];
        if (MODE == 1) {
          if (relu_mask && !(fmaf(sc[e], hv[e], sh[e]) > 0.f)) x = 0.f;
          s2[e] = fmaf(x, (hv[e] - mu[e]) * is[e], s2[e]);
        }
        s1[e] += x;
      }
    }
  }
  // reduce over lanes through shared memory
  if (lane < lanes) {
#pragma unroll
    for (int e = 0; e < VW; ++e) {
      red[(0 * lanes + lane) * C + c + e] = s1[e];
      red[(1 * lanes + lane) * C + c + e] = s2[e];
    }
  }
  __syncthreads();
  for (int i = tid; i < 2 * C; i += 256) {
    const int which = i / C, ch = i % C;
    if (MODE == 0 && which == 1) continue;
    double s = 0.0;
    for (int l = 0; l < lanes; ++l) s += (double)red[(which * lanes + l) * C + ch];
    if (MODE == 0)
      atomicAdd(out_f + ch, (float)s);
    else
      atomicAdd(stats + which * C + ch, s);
  }
}

// ---------------------------------------------------------------- BN backward apply
template <typename T, int VW>
__global__ void bn_bwd_apply_kernel(const T* __restrict__ g, const T* __restrict__ h,
                                    const float* __restrict__ mean, const float* __restrict__ invstd,
                                    const float* __restrict__ gamma, const float* __restrict__ scale,
                                    const float* __restrict__ shift, int relu_mask,
                                    const double* __restrict__ stats, double count, T* __restrict__ dh,
                                    float* dgamma, float* dbeta, long long total_vec, int C) {
  const int cq = C / VW;
  if (blockIdx.x == 0) {
    for (int c = threadIdx.x; c < C; c += blockDim.x) {
      if (dbeta) dbeta[c] += (float)stats[c];
      if (dgamma) dgamma[c] += (float)stats[C + c];
    }
  }
  const float inv_count = (float)(1.0 / count);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    const int c = (int)(i % cq) * VW;
    const long long o = i * VW;
    float gv[4], hv[4];
    if (VW == 4) {
      float4 t = Act<T>::ld4(g + o);
      gv[0] = t.x; gv[1] = t.y; gv[2] = t.z; gv[3] = t.w;
      float4 u = Act<T>::ld4(h + o);
      hv[0] = u.x; hv[1] = u.y; hv[2] = u.z; hv[3] = u.w;
    } else {
      gv[0] = Act<T>::ld(g + o);
      hv[0] = Act<T>::ld(h + o);
    }
#pragma unroll
    for (int e = 0; e < VW; ++e) {
      const int ch = c + e;
      float x = gv[e];
      if (relu_mask && !(fmaf(__ldg(scale + ch), hv[e], __ldg(shift + ch)) > 0.f)) x = 0.f;
      const float is = __ldg(invstd + ch);
      const float xhat = (hv[e] - __ldg(mean + ch)) * is;
      const float sg = (float)stats[ch] * inv_count, sgx = (float)stats[C + ch] * inv_count;
      gv[e] = __ldg(gamma + ch) * is * (x - sg - xhat * sgx);
    }
    if (VW == 4)
      Act<T>::st4(dh + o, make_float4(gv[0], gv[1], gv[2], gv[3]));
    else
      Act<T>::st(dh + o, gv[0]);
  }
}

template <typename T, int VW>
__global__ void relu_bwd_add_kernel(const T* __restrict__ g1, const T* __restrict__ g2, const T* __restrict__ x,
                                    T* __restrict__ dx, long long total_vec) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    const long long o = i * VW;
    if (VW == 4) {
      float4 a = Act<T>::ld4(g1 + o);
      if (g2) {
        float4 b = Act<T>::ld4(g2 + o);
        a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
      }
      float4 m = Act<T>::ld4(x + o);
      a.x = m.x > 0.f ? a.x : 0.f;
      a.y = m.y > 0.f ? a.y : 0.f;
      a.z = m.z > 0.f ? a.z : 0.f;
      a.w = m.w > 0.f ? a.w : 0.f;
      Act<T>::st4(dx + o, a);
    } else {
      float a = Act<T>::ld(g1 + o);
      if (g2) a += Act<T>::ld(g2 + o);
      Act<T>::st(dx + o, Act<T>::ld(x + o) > 0.f ? a : 0.f);
    }
  }
}

__global__ void tanh_bwd_kernel(const float* __restrict__ g, const float* __restrict__ y, float* __restrict__ out,
                                long long n) {
  pdl_wait();
  pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const float t = y[i];
    out[i] = g[i] * (1.f - t * t);
  }
}

// ---------------------------------------------------------------- reconstruction loss (train.py:118-129)
__global__ void __launch_bounds__(256)
recon_loss_kernel(const float* __restrict__ xt, const float* __restrict__ c, long long rows, int w, int wt,
                  double* loss_sum, float* __restrict__ dpre, float gscale) {
  pdl_wait();
  pdl_trigger();
  float acc = 0.f;
  const long long total = rows * w;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / w;
    const int wi = (int)(i % w);
    const float t = __ldg(c + i);
    float x = 0.f;
    if (wi < wt) x = __ldg(xt + r * wt + wi);
    const float d = x - t;
    acc = fmaf(d, d, acc);
    if (dpre && wi < wt) dpre[r * wt + wi] = gscale * d * (1.f - x * x);
  }
  __shared__ float part[8];
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < 8; ++i) s += (double)part[i];
    atomicAdd(loss_sum, s);
  }
}

// bf16 decoder output: the same loss with the two dtype conversions around it folded in (x_tilde widened to fp32 for
// the caller, dpre narrowed to the bf16 the decoder's backward kernels read) -- one pass instead of three
__global__ void __launch_bounds__(256)
recon_loss_bf16_kernel(const __nv_bfloat16* __restrict__ xt, const float* __restrict__ c, long long rows, int w, int wt,
                       double* loss_sum, __nv_bfloat16* __restrict__ dpre, float* __restrict__ xt32, float gscale,
                       float* __restrict__ dpre_sum, int vec) {
  pdl_wait();
  pdl_trigger();
  float acc = 0.f, dsum = 0.f;
  const long long total = rows * w;
  if (vec) {
    // x_tilde as wide as the target (T a multiple of 4, the usual case): no padding, so element i of one is element i of
    // the other -- 8 elements per thread through 16-byte accesses, no 64-bit division per element
    for (long long i8 = (long long)blockIdx.x * blockDim.x + threadIdx.x; i8 * 8 < total;
         i8 += (long long)gridDim.x * blockDim.x) {
      const uint4 xr = __ldg(reinterpret_cast<const uint4*>(xt + i8 * 8));
      const float4 t0 = __ldg(reinterpret_cast<const float4*>(c + i8 * 8)), t1 = __ldg(reinterpret_cast<const float4*>(c + i8 * 8 + 4));
      const uint32_t xw[4] = {xr.x, xr.y, xr.z, xr.w};
      const float tv[8] = {t0.x, t0.y, t0.z, t0.w, t1.x, t1.y, t1.z, t1.w};
      float xv[8];
      uint32_t dw[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        xv[2 * k] = __uint_as_float(xw[k] << 16);
        xv[2 * k + 1] = __uint_as_float(xw[k] & 0xffff0000u);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const float d0 = xv[2 * k] - tv[2 * k], d1 = xv[2 * k + 1] - tv[2 * k + 1];
        acc = fmaf(d0, d0, fmaf(d1, d1, acc));
        const __nv_bfloat16 e0 = __float2bfloat16_rn(gscale * d0 * (1.f - xv[2 * k] * xv[2 * k]));
        const __nv_bfloat16 e1 = __float2bfloat16_rn(gscale * d1 * (1.f - xv[2 * k + 1] * xv[2 * k + 1]));
        dsum += __bfloat162float(e0) + __bfloat162float(e1);
        dw[k] = (uint32_t)__bfloat16_as_ushort(e0) | ((uint32_t)__bfloat16_as_ushort(e1) << 16);
      }
      if (dpre) *reinterpret_cast<uint4*>(dpre + i8 * 8) = make_uint4(dw[0], dw[1], dw[2], dw[3]);
      if (xt32) {
        *reinterpret_cast<float4*>(xt32 + i8 * 8) = make_float4(xv[0], xv[1], xv[2], xv[3]);
        *reinterpret_cast<float4*>(xt32 + i8 * 8 + 4) = make_float4(xv[4], xv[5], xv[6], xv[7]);
      }
    }
    if (!dpre) dsum = 0.f;
  } else
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / w;
    const int wi = (int)(i % w);
    const float t = __ldg(c + i);
    float x = 0.f;
    if (wi < wt) {
      x = __bfloat162float(xt[r * wt + wi]);
      if (xt32) xt32[r * wt + wi] = x;
    }
    const float d = x - t;
    acc = fmaf(d, d, acc);
    if (dpre && wi < wt) {
      const __nv_bfloat16 dv = __float2bfloat16_rn(gscale * d * (1.f - x * x));
      dpre[r * wt + wi] = dv;
      dsum += __bfloat162float(dv);        // (the rounded value: what the decoder's backward kernels read)
    }
  }
  __shared__ float part[8], part2[8];
  acc = warp_sum(acc);
  dsum = warp_sum(dsum);
  if ((threadIdx.x & 31) == 0) {
    part[threadIdx.x >> 5] = acc;
    part2[threadIdx.x >> 5] = dsum;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    float s2 = 0.f;
    for (int i = 0; i < 8; ++i) {
      s += (double)part[i];
      s2 += part2[i];
    }
    atomicAdd(loss_sum, s);
    if (dpre_sum) atomicAdd(dpre_sum, s2);   // gradient of the output layer's single bias (models.py:182)
  }
}

// losses[0] = sums[0] * a, losses[1] = losses[2] = sums[1] * b   (train.py:129-134: recon, vq, commit share one sum)
__global__ void loss_finalize_kernel(const double* __restrict__ sums, double a, double b, double* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  if (threadIdx.x == 0) out[0] = sums[0] * a;
  if (threadIdx.x == 1) out[1] = sums[1] * b;
  if (threadIdx.x == 2) out[2] = sums[1] * b;
}

// ---------------------------------------------------------------- Adam (torch.optim.Adam single-tensor maths)
// VEC = 4: 16-byte accesses (n % 4 == 0, 16-byte aligned buffers).  zero != 0 clears each gradient after its use, which
// replaces the next step's zero_grad fill.
template <int VEC>
__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, long long n,
            float lr, float b1, float b2, float eps, float gscale, long long* step_count, int zero,
            unsigned int* ticket) {
  pdl_wait();
  pdl_trigger();
  __shared__ float s_step_size, s_bc2_sqrt;
  const long long nv = n / VEC;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  // The first (at this model's size: the only) batch of loads is requested BEFORE the bias corrections are derived:
  // two double-precision pow() calls by one thread and the barrier behind them are several microseconds during which
  // the tensors would otherwise not even have been asked for.
  float4 g4 = make_float4(0.f, 0.f, 0.f, 0.f), m4 = g4, v4 = g4, p4 = g4;
  float g1 = 0.f, m1 = 0.f, v1 = 0.f, p1 = 0.f;
  auto request = [&](long long j) {
    if (j >= nv) return;
    if (VEC == 4) {
      g4 = reinterpret_cast<const float4*>(g)[j]; m4 = reinterpret_cast<const float4*>(m)[j];
      v4 = reinterpret_cast<const float4*>(v)[j]; p4 = reinterpret_cast<const float4*>(p)[j];
    } else {
      g1 = g[j]; m1 = m[j]; v1 = v[j]; p1 = p[j];
    }
  };
  request(i);
  if (threadIdx.x == 0) {
    // beta ** step by binary exponentiation (the step is an integer: ~20 dependent double multiplications for both
    // powers together, against two calls of the general pow(), a few thousand cycles in front of the CTA's barrier)
    long long t = *step_count + 1;
    double q1 = (double)b1, q2 = (double)b2, r1 = 1.0, r2 = 1.0;
    while (t > 0) {
      if (t & 1) {
        r1 *= q1;
        r2 *= q2;
      }
      q1 *= q1;
      q2 *= q2;
      t >>= 1;
    }
    const double bc1 = 1.0 - r1;
    const double bc2 = 1.0 - r2;
    s_step_size = (float)((double)lr / bc1);
    s_bc2_sqrt = (float)sqrt(bc2);
  }
  __syncthreads();
  const float step_size = s_step_size, bc2_sqrt = s_bc2_sqrt;
  for (; i < nv; i += (long long)gridDim.x * blockDim.x) {
    float gr[VEC], mi[VEC], vi[VEC], pi[VEC];
    if (VEC == 4) {
      gr[0] = g4.x; gr[1 % VEC] = g4.y; gr[2 % VEC] = g4.z; gr[3 % VEC] = g4.w;
      mi[0] = m4.x; mi[1 % VEC] = m4.y; mi[2 % VEC] = m4.z; mi[3 % VEC] = m4.w;
      vi[0] = v4.x; vi[1 % VEC] = v4.y; vi[2 % VEC] = v4.z; vi[3 % VEC] = v4.w;
      pi[0] = p4.x; pi[1 % VEC] = p4.y; pi[2 % VEC] = p4.z; pi[3 % VEC] = p4.w;
    } else {
      gr[0] = g1; mi[0] = m1; vi[0] = v1; pi[0] = p1;
    }
    request(i + (long long)gridDim.x * blockDim.x);
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      const float ge = gr[e] * gscale;
      mi[e] = mi[e] + (ge - mi[e]) * (1.f - b1);          // exp_avg.lerp_(grad, 1 - beta1)
      vi[e] = vi[e] * b2 + (1.f - b2) * ge * ge;          // exp_avg_sq.mul_(beta2).addcmul_(grad, grad, 1 - beta2)
      const float denom = sqrtf(vi[e]) / bc2_sqrt + eps;
      pi[e] = pi[e] - step_size * (mi[e] / denom);
    }
    if (VEC == 4) {
      reinterpret_cast<float4*>(p)[i] = make_float4(pi[0], pi[1 % VEC], pi[2 % VEC], pi[3 % VEC]);
      reinterpret_cast<float4*>(m)[i] = make_float4(mi[0], mi[1 % VEC], mi[2 % VEC], mi[3 % VEC]);
      reinterpret_cast<float4*>(v)[i] = make_float4(vi[0], vi[1 % VEC], vi[2 % VEC], vi[3 % VEC]);
      if (zero) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      p[i] = pi[0]; m[i] = mi[0]; v[i] = vi[0];
      if (zero) g[i] = 0.f;
    }
  }
  // step counter: advanced by the LAST block to finish (every block read it at its start), so the optimiser step
  // is one launch; `ticket` is a zero-initialised counter owned by the caller, left at zero again
  if (ticket) {
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      if (atomicAdd(ticket, 1u) == gridDim.x - 1) {
        *ticket = 0u;
        *step_count += 1;
      }
    }
  }
}
__global__ void tick_kernel(long long* step_count) {
  pdl_wait();
  pdl_trigger(); *step_count += 1; }

}  // namespace nsg

using namespace nsg;

extern "C" {

int nsg_bn_finalize(const double* stats, int channels, double count, const float* gamma, const float* beta,
                    float eps, float momentum, float* running_mean, float* running_var,
                    long long* num_batches_tracked, float* mean, float* invstd, float* scale, float* shift,
                    void* stream) {
  NSG_REQUIRE(stats && gamma && beta && mean && invstd && scale && shift && channels > 0 && count > 0,
              "bn_finalize: bad arguments");
  bn_finalize_kernel<<<(channels + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      stats, channels, count, gamma, beta, eps, momentum, running_mean, running_var, num_batches_tracked, mean,
      invstd, scale, shift);
  NSG_CHECK_LAUNCH("bn_finalize_kernel");
  return NSG_OK;
}

int nsg_bn_eval_coeffs(int channels, const float* gamma, const float* beta, const float* running_mean,
                       const float* running_var, float eps, float* scale, float* shift, void* stream) {
  NSG_REQUIRE(gamma && beta && running_mean && running_var && scale && shift && channels > 0,
              "bn_eval_coeffs: bad arguments");
  bn_eval_coeffs_kernel<<<(channels + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
      channels, gamma, beta, running_mean, running_var, eps, scale, shift);
  NSG_CHECK_LAUNCH("bn_eval_coeffs_kernel");
  return NSG_OK;
}

int nsg_bn_apply(const void* h, const float* scale, const float* shift, const void* res, int res_relu,
                 int relu_out, void* y, long long pixels, int channels, int dtype, void* stream) {
  NSG_REQUIRE(h && scale && shift && y && pixels >= 0 && channels > 0, "bn_apply: bad arguments");
  if (pixels == 0) return NSG_OK;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    if (channels % 4 == 0) {
      const long long tv = pixels * channels / 4;
      bn_apply_kernel<T, 4><<<grid_for(tv, 256), 256, 0, (cudaStream_t)stream>>>(
          (const T*)h, scale, shift, (const T*)res, res_relu, relu_out, (T*)y, tv, channels);
    } else {
      const long long tv = pixels * channels;
      bn_apply_kernel<T, 1><<<grid_for(tv, 256), 256, 0, (cudaStream_t)stream>>>(
          (const T*)h, scale, shift, (const T*)res, res_relu, relu_out, (T*)y, tv, channels);
    }
    NSG_CHECK_LAUNCH("bn_apply_kernel");
    return NSG_OK;
  });
}

}  // extern "C"

template <typename T, int MODE>
static int launch_chan_reduce(const void* g, const void* h, const float* mean, const float* invstd,
                              const float* scale, const float* shift, int relu_mask, double* stats, float* out_f,
                              long long pixels, int C, cudaStream_t stream) {
  const int vw = (C % 4 == 0) ? 4 : 1;
  const int cq = C / vw;
  NSG_REQUIRE(cq <= 256, "channel reduction: too many channels (%d)", C);
  const int lanes = 256 / cq;
  const size_t smem = (size_t)2 * lanes * C * sizeof(float);
  NSG_REQUIRE(smem <= 48 * 1024, "channel reduction: shared memory %zu too large", smem);
  long long blocks = (pixels + (long long)lanes * 8 - 1) / ((long long)lanes * 8);
  const long long cap = (long long)num_sms() * 4;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  if (vw == 4)
    launch_pdl(chan_reduce_kernel<T, 4, MODE>, dim3((int)blocks), dim3(256), smem, stream, 
        (const T*)g, (const T*)h, mean, invstd, scale, shift, relu_mask, stats, out_f, pixels, C);
  else
    launch_pdl(chan_reduce_kernel<T, 1, MODE>, dim3((int)blocks), dim3(256), smem, stream, 
        (const T*)g, (const T*)h, mean, invstd, scale, shift, relu_mask, stats, out_f, pixels, C);
  NSG_CHECK_LAUNCH("chan_reduce_kernel");
  return NSG_OK;
}

extern "C" {

int nsg_bn_bwd_reduce(const void* g, const void* h, const float* mean, const float* invstd, const float* scale,
                      const float* shift, int relu_mask, double* stats, long long pixels, int channels,
                      int dtype, void* stream) {
  NSG_REQUIRE(g && h && mean && invstd && stats && channels > 0, "bn_bwd_reduce: bad arguments");
  if (relu_mask) NSG_REQUIRE(scale && shift, "bn_bwd_reduce: relu mask needs scale/shift");
  if (pixels == 0) return NSG_OK;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    return launch_chan_reduce<T, 1>(g, h, mean, invstd, scale, shift, relu_mask, stats, nullptr, pixels, channels,
                                    (cudaStream_t)stream);
  });
}

int nsg_colsum(const void* g, float* out, long long pixels, int channels, int dtype, void* stream) {
  NSG_REQUIRE(g && out && channels > 0, "colsum: bad arguments");
  if (pixels == 0) return NSG_OK;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    return launch_chan_reduce<T, 0>(g, nullptr, nullptr, nullptr, nullptr, nullptr, 0, nullptr, out, pixels,
                                    channels, (cudaStream_t)stream);
  });
}

int nsg_bn_bwd_apply(const void* g, const void* h, const float* mean, const float* invstd, const float* gamma,
                     const float* scale, const float* shift, int relu_mask, const double* stats, double count,
                     void* dh, float* dgamma, float* dbeta, long long pixels, int channels, int dtype,
                     void* stream) {
  NSG_REQUIRE(g && h && mean && invstd && gamma && stats && dh && channels > 0 && count > 0,
              "bn_bwd_apply: bad arguments");
  if (relu_mask) NSG_REQUIRE(scale && shift, "bn_bwd_apply: relu mask needs scale/shift");
  if (pixels == 0) return NSG_OK;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    if (channels % 4 == 0) {
      const long long tv = pixels * channels / 4;
      bn_bwd_apply_kernel<T, 4><<<grid_for(tv, 256), 256, 0, (cudaStream_t)stream>>>(
          (const T*)g, (const T*)h, mean, invstd, gamma, scale, shift, relu_mask, stats, count, (T*)dh, dgamma,
          dbeta, tv, channels);
    } else {
      const long long tv = pixels * channels;
      bn_bwd_apply_kernel<T, 1><<<grid_for(tv, 256), 256, 0, (cudaStream_t)stream>>>(
          (const T*)g, (const T*)h, mean, invstd, gamma, scale, shift, relu_mask, stats, count, (T*)dh, dgamma,
          dbeta, tv, channels);
    }
    NSG_CHECK_LAUNCH("bn_bwd_apply_kernel");
    return NSG_OK;
  });
}

int nsg_relu_bwd_add(const void* g1, const void* g2, const void* x, void* dx, long long n, int dtype,
                     void* stream) {
  NSG_REQUIRE(g1 && x && dx && n >= 0, "relu_bwd_add: bad arguments");
  if (n == 0) return NSG_OK;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    if (n % 4 == 0) {
      relu_bwd_add_kernel<T, 4><<<grid_for(n / 4, 256), 256, 0, (cudaStream_t)stream>>>(
          (const T*)g1, (const T*)g2, (const T*)x, (T*)dx, n / 4);
    } else {
      relu_bwd_add_kernel<T, 1><<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(
          (const T*)g1, (const T*)g2, (const T*)x, (T*)dx, n);
    }
    NSG_CHECK_LAUNCH("relu_bwd_add_kernel");
    return NSG_OK;
  });
}

int nsg_recon_loss(const float* x_tilde, const float* c, int batch, int h, int w, int wt, double* loss_sum,
                   float* dpre, void* stream) {
  NSG_REQUIRE(x_tilde && c && loss_sum && batch > 0 && h > 0 && w > 0 && wt >= 0 && wt <= w,
              "recon_loss: bad arguments");
  const long long rows = (long long)batch * h;
  const float gscale = (float)(2.0 / ((double)rows * w));
  launch_pdl(recon_loss_kernel, dim3(grid_for(rows * w, 256)), dim3(256), 0, (cudaStream_t)stream, x_tilde, c, rows, w, wt, loss_sum,
                                                                                 dpre, gscale);
  NSG_CHECK_LAUNCH("recon_loss_kernel");
  return NSG_OK;
}

int nsg_recon_loss_bf16(const void* x_tilde, const float* c, int batch, int h, int w, int wt, double* loss_sum,
                        void* dpre, float* x_tilde_f32, float* dpre_sum, void* stream) {
  NSG_REQUIRE(x_tilde && c && loss_sum && batch >= 0 && h >= 0 && w >= 0 && wt >= 0 && wt <= w,
              "recon_loss_bf16: bad arguments");
  const long long rows = (long long)batch * h;
  if (rows * w == 0) return NSG_OK;
  const float gscale = (float)(2.0 / ((double)rows * w));
  const int vec = (wt == w && (w & 7) == 0 &&
                   ((reinterpret_cast<uintptr_t>(x_tilde) | reinterpret_cast<uintptr_t>(c) | reinterpret_cast<uintptr_t>(dpre) |
                     reinterpret_cast<uintptr_t>(x_tilde_f32)) & 15) == 0) ? 1 : 0;
  launch_pdl(recon_loss_bf16_kernel, dim3(grid_for(vec ? rows * w / 8 : rows * w, 256)), dim3(256), 0,
             (cudaStream_t)stream,
             static_cast<const __nv_bfloat16*>(x_tilde), c, rows, w, wt, loss_sum, static_cast<__nv_bfloat16*>(dpre),
             x_tilde_f32, gscale, dpre_sum, vec);
  NSG_CHECK_LAUNCH("recon_loss_bf16_kernel");
  return NSG_OK;
}

int nsg_loss_finalize(const double* sums, double recon_scale, double vq_scale, double* losses3, void* stream) {
  NSG_REQUIRE(sums && losses3, "loss_finalize: null pointer");
  launch_pdl(loss_finalize_kernel, dim3(1), dim3(32), 0, (cudaStream_t)stream, sums, recon_scale, vq_scale, losses3);
  NSG_CHECK_LAUNCH("loss_finalize_kernel");
  return NSG_OK;
}

int nsg_tanh_bwd(const float* g, const float* y, float* out, long long n, void* stream) {
  NSG_REQUIRE(g && y && out && n >= 0, "tanh_bwd: bad arguments");
  if (n == 0) return NSG_OK;
  launch_pdl(tanh_bwd_kernel, dim3(grid_for(n, 256)), dim3(256), 0, (cudaStream_t)stream, g, y, out, n);
  NSG_CHECK_LAUNCH("tanh_bwd_kernel");
  return NSG_OK;
}

int nsg_adam_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                  float beta1, float beta2, float eps, float grad_scale, long long* step_count, int zero_grads,
                  unsigned int* ticket, void* stream) {
  NSG_REQUIRE(params && grads && exp_avg && exp_avg_sq && step_count && n >= 0, "adam_step: bad arguments");
  if (n > 0) {
    const bool vec = n % 4 == 0 && ((reinterpret_cast<uintptr_t>(params) | reinterpret_cast<uintptr_t>(grads) |
                                     reinterpret_cast<uintptr_t>(exp_avg) | reinterpret_cast<uintptr_t>(exp_avg_sq)) & 15) == 0;
    if (vec) {
      int blocks = grid_for(n / 4, 256);
      if (blocks > 2 * num_sms()) blocks = 2 * num_sms();
      launch_pdl(adam_kernel<4>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, params, grads, exp_avg, exp_avg_sq, n,
                 lr, beta1, beta2, eps, grad_scale, step_count, zero_grads, ticket);
    } else {
      launch_pdl(adam_kernel<1>, dim3(grid_for(n, 256)), dim3(256), 0, (cudaStream_t)stream, params, grads, exp_avg,
                 exp_avg_sq, n, lr, beta1, beta2, eps, grad_scale, step_count, zero_grads, ticket);
    }
    NSG_CHECK_LAUNCH("adam_kernel");
    if (ticket) return NSG_OK;
  }
  launch_pdl(tick_kernel, dim3(1), dim3(1), 0, (cudaStream_t)stream, step_count);
  NSG_CHECK_LAUNCH("tick_kernel");
  return NSG_OK;
}

}  // extern "C"
