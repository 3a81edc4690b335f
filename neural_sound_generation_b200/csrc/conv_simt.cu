// fp32 CUDA-core implicit-GEMM convolution family ("accuracy mode" and small-channel layers).
//
// One kernel (tapgemm) serves Conv2d / ConvTranspose2d forward and both data gradients; one kernel
// (wgrad) serves both weight gradients.  See include/nsg_b200.h for the contract.  These are the
// exact-fp32 kernels: 128x64 output tile, K chunks of 16, register-prefetched global loads,
// 8x4 outputs per thread.  The tensor-core (tcgen05) kernels in conv_tc.cu replace them for the
// 64-channel bf16 layers.
#include "common.cuh"
#include "conv_geom.cuh"

namespace nsg {

constexpr int BM = 128, BN = 64, BK = 16, AS_LD = BM + 4;

template <typename T, bool VEC>
__global__ void __launch_bounds__(256, 2)
tapgemm_simt_kernel(const T* __restrict__ in, const float* __restrict__ w, T* __restrict__ out, Pro pro,
                    EpiArgs epi, TapGeom g) {
  __shared__ __align__(16) float As[BK][AS_LD];
  __shared__ __align__(16) float Bs[BK][BN];
  __shared__ int rowN[BM], rowH[BM], rowW[BM];
  __shared__ float sStat[2][BN];

  const int tid = threadIdx.x;
  const long long M = (long long)g.batch * g.sub_h * g.sub_w;
  const long long m0 = (long long)blockIdx.x * BM;
  const int n0 = blockIdx.y * BN;
  const int K = g.ntaps * g.cin;

  if (tid < BM) {
    long long m = m0 + tid;
    if (m < M) {
      int ws = (int)(m % g.sub_w);
      long long t = m / g.sub_w;
      rowW[tid] = ws;
      rowH[tid] = (int)(t % g.sub_h);
      rowN[tid] = (int)(t / g.sub_h);
    } else {
      rowN[tid] = -1;
      rowH[tid] = 0;
      rowW[tid] = 0;
    }
  }
  if (tid < 2 * BN) (&sStat[0][0])[tid] = 0.f;
  __syncthreads();

  // ---- per-thread loader roles
  // VEC:    rows {tid/4, tid/4+64}, channels (tid%4)*4..+3 of the chunk
  // scalar: row tid%128, chunk-k {tid/128 + 2j}
  float ra[8];
  float rb[4];
  const int b_kl = tid >> 4, b_n4 = (tid & 15) * 4;

  auto load_chunk = [&](int kc) {
    const int k0 = kc * BK;
    if (VEC) {
      const int tap = k0 / g.cin;
      const int c = (k0 % g.cin) + (tid & 3) * 4;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int r = (tid >> 2) + 64 * j;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        const int n = rowN[r];
        const int ih = rowH[r] * g.in_stride + g.dh[tap];
        const int iw = rowW[r] * g.in_stride + g.dw[tap];
        if (n >= 0 && ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w) {
          const long long off = (((long long)n * g.in_h + ih) * g.in_w + iw) * g.cin + c;
          v = Act<T>::ld4(in + off);
          if (pro.mode != NSG_PRO_NONE) {
            float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
            if (pro.mode >= NSG_PRO_AFFINE_RELU) {
              sc = __ldg(reinterpret_cast<const float4*>(pro.scale + c));
              sh = __ldg(reinterpret_cast<const float4*>(pro.shift + c));
            }
            v.x = pro_apply(pro.mode, v.x, sc.x, sh.x);
            v.y = pro_apply(pro.mode, v.y, sc.y, sh.y);
            v.z = pro_apply(pro.mode, v.z, sc.z, sh.z);
            v.w = pro_apply(pro.mode, v.w, sc.w, sh.w);
          }
        }
        ra[4 * j + 0] = v.x;
        ra[4 * j + 1] = v.y;
        ra[4 * j + 2] = v.z;
        ra[4 * j + 3] = v.w;
      }
    } else {
      const int r = tid & 127;
      const int n = rowN[r];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int k = k0 + (tid >> 7) + 2 * j;
        float v = 0.f;
        if (k < K && n >= 0) {
          const int tap = k / g.cin, c = k % g.cin;
          const int ih = rowH[r] * g.in_stride + g.dh[tap];
          const int iw = rowW[r] * g.in_stride + g.dw[tap];
          if (ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w) {
            const long long off = (((long long)n * g.in_h + ih) * g.in_w + iw) * g.cin + c;
            v = Act<T>::ld(in + off);
            float sc = 1.f, sh = 0.f;
            if (pro.mode >= NSG_PRO_AFFINE_RELU) {
              sc = __ldg(pro.scale + c);
              sh = __ldg(pro.shift + c);
            }
            v = pro_apply(pro.mode, v, sc, sh);
          }
        }
        ra[j] = v;
      }
    }
    {
      const int k = k0 + b_kl;
      const bool kv = k < K;
      const int tap = kv ? k / g.cin : 0, c = kv ? k % g.cin : 0;
      const long long base = g.woff[tap] + (long long)c * g.w_cin_stride;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int co = n0 + b_n4 + e;
        rb[e] = (kv && co < g.cout) ? __ldg(w + base + (long long)co * g.w_cout_stride) : 0.f;
      }
    }
  };
  auto store_chunk = [&]() {
    if (VEC) {
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int r = (tid >> 2) + 64 * j;
#pragma unroll
        for (int e = 0; e < 4; ++e) As[(tid & 3) * 4 + e][r] = ra[4 * j + e];
      }
    } else {
      const int r = tid & 127;
#pragma unroll
      for (int j = 0; j < 8; ++j) As[(tid >> 7) + 2 * j][r] = ra[j];
    }
    *reinterpret_cast<float4*>(&Bs[b_kl][b_n4]) = make_float4(rb[0], rb[1], rb[2], rb[3]);
  };

  const int tx = tid & 15, ty = tid >> 4;
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;

  const int nchunks = (K + BK - 1) / BK;
  load_chunk(0);
  for (int kc = 0; kc < nchunks; ++kc) {
    store_chunk();
    __syncthreads();
    if (kc + 1 < nchunks) load_chunk(kc + 1);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 8 + 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][e] = fmaf(a[i], bb[e], acc[i][e]);
    }
    __syncthreads();
  }

  // ---- epilogue
  const int co0 = n0 + tx * 4;
  float bias[4], msc[4], msh[4], mu[4], istd[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const int co = co0 + e;
    const bool ok = co < g.cout;
    bias[e] = (ok && epi.bias) ? __ldg(epi.bias + co) : 0.f;
    msc[e] = (ok && epi.mask_mode == NSG_MASK_AFFINE) ? __ldg(epi.mscale + co) : 1.f;
    msh[e] = (ok && epi.mask_mode == NSG_MASK_AFFINE) ? __ldg(epi.mshift + co) : 0.f;
    mu[e] = (ok && epi.stats_mode == NSG_STATS_BNBWD) ? __ldg(epi.mean + co) : 0.f;
    istd[e] = (ok && epi.stats_mode == NSG_STATS_BNBWD) ? __ldg(epi.invstd + co) : 0.f;
  }
  const T* add_src = static_cast<const T*>(epi.add_src);
  const T* mask_src = static_cast<const T*>(epi.mask_src);
  const T* xhat_src = static_cast<const T*>(epi.xhat_src);
  const bool vec_out = (g.cout % 4 == 0) && (co0 + 3 < g.cout);
  float s1[4] = {0.f, 0.f, 0.f, 0.f}, s2[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int r = ty * 8 + i;
    const int n = rowN[r];
    if (n < 0) continue;
    const int oh = rowH[r] * g.out_stride + g.off_h;
    const int ow = rowW[r] * g.out_stride + g.off_w;
    const long long o = (((long long)n * g.full_h + oh) * g.full_w + ow) * g.cout + co0;
    float v[4], ad[4] = {0.f, 0.f, 0.f, 0.f}, mk[4] = {1.f, 1.f, 1.f, 1.f}, hx[4] = {0.f, 0.f, 0.f, 0.f};
    if (vec_out) {
      if (add_src) {
        float4 t = Act<T>::ld4(add_src + o);
        ad[0] = t.x; ad[1] = t.y; ad[2] = t.z; ad[3] = t.w;
      }
      if (epi.mask_mode != NSG_MASK_NONE) {
        float4 t = Act<T>::ld4(mask_src + o);
        mk[0] = t.x; mk[1] = t.y; mk[2] = t.z; mk[3] = t.w;
      }
      if (epi.stats_mode == NSG_STATS_BNBWD) {
        float4 t = Act<T>::ld4(xhat_src + o);
        hx[0] = t.x; hx[1] = t.y; hx[2] = t.z; hx[3] = t.w;
      }
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        if (co0 + e < g.cout) {
          if (add_src) ad[e] = Act<T>::ld(add_src + o + e);
          if (epi.mask_mode != NSG_MASK_NONE) mk[e] = Act<T>::ld(mask_src + o + e);
          if (epi.stats_mode == NSG_STATS_BNBWD) hx[e] = Act<T>::ld(xhat_src + o + e);
        }
      }
    }
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float x = acc[i][e] + bias[e] + ad[e];
      if (epi.mask_mode == NSG_MASK_POS) x = mk[e] > 0.f ? x : 0.f;
      if (epi.mask_mode == NSG_MASK_AFFINE) x = fmaf(msc[e], mk[e], msh[e]) > 0.f ? x : 0.f;
      if (epi.tanh_out) x = tanhf(x);
      v[e] = x;
      if (epi.stats_mode == NSG_STATS_MOMENTS) {
        s1[e] += x;
        s2[e] = fmaf(x, x, s2[e]);
      } else if (epi.stats_mode == NSG_STATS_BNBWD) {
        s1[e] += x;
        s2[e] = fmaf(x, (hx[e] - mu[e]) * istd[e], s2[e]);
      }
    }
    if (vec_out) {
      Act<T>::st4(out + o, make_float4(v[0], v[1], v[2], v[3]));
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e)
        if (co0 + e < g.cout) Act<T>::st(out + o + e, v[e]);
    }
  }
  if (epi.stats_mode != NSG_STATS_NONE) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      atomicAdd(&sStat[0][tx * 4 + e], s1[e]);
      atomicAdd(&sStat[1][tx * 4 + e], s2[e]);
    }
    __syncthreads();
    if (tid < 2 * BN) {
      const int which = tid / BN, c = tid % BN;
      if (n0 + c < g.cout) atomicAdd(epi.stats + which * g.cout + n0 + c, (double)sStat[which][c]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// weight gradient: G[a][b] = sum_m P[gather(m, tap(a)), ch(a)] * Q[m, b],  a = tap*ca + ch
// tile 64(a) x 64(b), 16 pixels per stage, pixel range split over blockIdx.z, float atomics at the end.
struct WgradGeom {
  int batch, p_h, p_w, ca;
  int q_h, q_w, cb;
  int ksize, stride, pad;
  long long w_a_stride, w_b_stride;
  int pix_per_cta;
};

template <typename T, bool VEC>
__global__ void __launch_bounds__(256, 2)
wgrad_simt_kernel(const T* __restrict__ P, const T* __restrict__ Q, float* __restrict__ dw,
                  float* __restrict__ q_colsum, Pro proP, Pro proQ, WgradGeom g) {
  __shared__ __align__(16) float Ps[16][64];
  __shared__ __align__(16) float Qs[16][64];
  const int tid = threadIdx.x;
  const int ntaps = g.ksize * g.ksize;
  const int A_total = ntaps * g.ca;
  const int a0 = blockIdx.x * 64, b0 = blockIdx.y * 64;
  const long long M = (long long)g.batch * g.q_h * g.q_w;
  const long long mbeg = (long long)blockIdx.z * g.pix_per_cta;
  const long long mend = min(M, mbeg + (long long)g.pix_per_cta);

  const int kl = tid >> 4, c4 = (tid & 15) * 4;
  float rp[4], rq[4];

  auto load_stage = [&](long long mb) {
    const long long m = mb + kl;
    const bool mv = m < mend;
    int ow = 0, oh = 0, n = 0;
    if (mv) {
      ow = (int)(m % g.q_w);
      long long t = m / g.q_w;
      oh = (int)(t % g.q_h);
      n = (int)(t / g.q_h);
    }
    // P (gathered)
    if (VEC) {
      const int a = a0 + c4;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (mv && a < A_total) {
        const int tap = a / g.ca, c = a % g.ca;
        const int ih = oh * g.stride - g.pad + tap / g.ksize;
        const int iw = ow * g.stride - g.pad + tap % g.ksize;
        if (ih >= 0 && ih < g.p_h && iw >= 0 && iw < g.p_w) {
          v = Act<T>::ld4(P + (((long long)n * g.p_h + ih) * g.p_w + iw) * g.ca + c);
          if (proP.mode != NSG_PRO_NONE) {
            float4 sc = make_float4(1.f, 1.f, 1.f, 1.f), sh = make_float4(0.f, 0.f, 0.f, 0.f);
            if (proP.mode >= NSG_PRO_AFFINE_RELU) {
              sc = __ldg(reinterpret_cast<const float4*>(proP.scale + c));
              sh = __ldg(reinterpret_cast<const float4*>(proP.shift + c));
            }
            v.x = pro_apply(proP.mode, v.x, sc.x, sh.x);
            v.y = pro_apply(proP.mode, v.y, sc.y, sh.y);
            v.z = pro_apply(proP.mode, v.z, sc.z, sh.z);
            v.w = pro_apply(proP.mode, v.w, sc.w, sh.w);
          }
        }
      }
      rp[0] = v.x; rp[1] = v.y; rp[2] = v.z; rp[3] = v.w;
    } else {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int a = a0 + c4 + e;
        float v = 0.f;
        if (mv && a < A_total) {
          const int tap = a / g.ca, c = a % g.ca;
          const int ih = oh * g.stride - g.pad + tap / g.ksize;
          const int iw = ow * g.stride - g.pad + tap % g.ksize;
          if (ih >= 0 && ih < g.p_h && iw >= 0 && iw < g.p_w) {
            v = Act<T>::ld(P + (((long long)n * g.p_h + ih) * g.p_w + iw) * g.ca + c);
            float sc = 1.f, sh = 0.f;
            if (proP.mode >= NSG_PRO_AFFINE_RELU) {
              sc = __ldg(proP.scale + c);
              sh = __ldg(proP.shift + c);
            }
            v = pro_apply(proP.mode, v, sc, sh);
          }
        }
        rp[e] = v;
      }
    }
    // Q (direct)
#pragma unroll
    for (int e = 0; e < 4; ++e) rq[e] = 0.f;
    if (mv) {
      const int b = b0 + c4;
      if ((g.cb % 4 == 0) && b + 3 < g.cb) {
        float4 v = Act<T>::ld4(Q + m * g.cb + b);
        rq[0] = v.x; rq[1] = v.y; rq[2] = v.z; rq[3] = v.w;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float sc = 1.f, sh = 0.f;
          if (proQ.mode >= NSG_PRO_AFFINE_RELU) {
            sc = __ldg(proQ.scale + b + e);
            sh = __ldg(proQ.shift + b + e);
          }
          rq[e] = pro_apply(proQ.mode, rq[e], sc, sh);
        }
      } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (b + e < g.cb) {
            float sc = 1.f, sh = 0.f;
            if (proQ.mode >= NSG_PRO_AFFINE_RELU) {
              sc = __ldg(proQ.scale + b + e);
              sh = __ldg(proQ.shift + b + e);
            }
            rq[e] = pro_apply(proQ.mode, Act<T>::ld(Q + m * g.cb + b + e), sc, sh);
          }
        }
      }
    }
  };

  const int ta = tid >> 4, tb = tid & 15;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
  float qs[4] = {0.f, 0.f, 0.f, 0.f};

  if (mbeg < mend) load_stage(mbeg);
  for (long long mb = mbeg; mb < mend; mb += 16) {
    *reinterpret_cast<float4*>(&Ps[kl][c4]) = make_float4(rp[0], rp[1], rp[2], rp[3]);
    *reinterpret_cast<float4*>(&Qs[kl][c4]) = make_float4(rq[0], rq[1], rq[2], rq[3]);
    __syncthreads();
    if (mb + 16 < mend) load_stage(mb + 16);
#pragma unroll
    for (int k = 0; k < 16; ++k) {
      const float4 a = *reinterpret_cast<const float4*>(&Ps[k][ta * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Qs[k][tb * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int e = 0; e < 4; ++e) acc[i][e] = fmaf(av[i], bv[e], acc[i][e]);
      if (ta == 0) {
#pragma unroll
        for (int e = 0; e < 4; ++e) qs[e] += bv[e];
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int a = a0 + ta * 4 + i;
    if (a >= A_total) continue;
    const int tap = a / g.ca, c = a % g.ca;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int b = b0 + tb * 4 + e;
      if (b < g.cb) atomicAdd(dw + tap + (long long)c * g.w_a_stride + (long long)b * g.w_b_stride, acc[i][e]);
    }
  }
  if (q_colsum && blockIdx.x == 0 && ta == 0) {
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int b = b0 + tb * 4 + e;
      if (b < g.cb) atomicAdd(q_colsum + b, qs[e]);
    }
  }
}

// ------------------------------------------------------------------------------------------------
int conv_gemm_c1(const nsg_conv_geom* cg, const void* in, const float* w, const Pro& pro, const EpiArgs& e, void* out,
                 cudaStream_t stream);
int conv_wgrad_c1(const nsg_conv_geom* cg, const void* p, const Pro& pp, const void* q, const Pro& pq, float* dw,
                  float* q_colsum, cudaStream_t stream);

int conv_gemm_simt(const nsg_conv_geom* cg, const void* in, const float* w, const nsg_prologue* pro,
                   const nsg_epilogue* epi, void* out, cudaStream_t stream) {
  NSG_REQUIRE(cg->ksize >= 1 && cg->ksize <= 4, "conv: ksize %d unsupported (1..4)", cg->ksize);
  NSG_REQUIRE(cg->stride >= 1 && cg->stride <= 2, "conv: stride %d unsupported (1..2)", cg->stride);
  TapGeom phases[4];
  const int np = build_phases(cg, phases);
  Pro p{NSG_PRO_NONE, nullptr, nullptr};
  if (pro) p = Pro{pro->mode, pro->scale, pro->shift};
  if (p.mode >= NSG_PRO_AFFINE_RELU) NSG_REQUIRE(p.scale && p.shift, "conv: affine prologue needs scale/shift");
  EpiArgs e{};
  if (int rc = make_epi(epi, &e)) return rc;
  {   // the single-channel ends of the network have their own streaming kernels (conv_c1.cu)
    const int served = conv_gemm_c1(cg, in, w, p, e, out, stream);
    if (served != 0) return served < 0 ? served : NSG_OK;
  }
  const bool vec = (cg->cin % 16 == 0);
  return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
    for (int i = 0; i < np; ++i) {
      const TapGeom& g = phases[i];
      const long long M = (long long)g.batch * g.sub_h * g.sub_w;
      if (M == 0 || g.ntaps == 0) continue;
      dim3 grid((unsigned)((M + BM - 1) / BM), (unsigned)((g.cout + BN - 1) / BN));
      if (vec)
        tapgemm_simt_kernel<T, true><<<grid, 256, 0, stream>>>((const T*)in, w, (T*)out, p, e, g);
      else
        tapgemm_simt_kernel<T, false><<<grid, 256, 0, stream>>>((const T*)in, w, (T*)out, p, e, g);
      NSG_CHECK_LAUNCH("tapgemm_simt_kernel");
    }
    return NSG_OK;
  });
}

int conv_wgrad_simt(const nsg_conv_geom* cg, const void* p, const nsg_prologue* pro_p, const void* q,
                    const nsg_prologue* pro_q, float* dw, float* q_colsum, cudaStream_t stream) {
  NSG_REQUIRE(cg->ksize >= 1 && cg->ksize <= 4, "wgrad: ksize %d unsupported", cg->ksize);
  WgradGeom g{};
  g.batch = cg->batch; g.p_h = cg->in_h; g.p_w = cg->in_w; g.ca = cg->cin;
  g.q_h = cg->out_h; g.q_w = cg->out_w; g.cb = cg->cout;
  g.ksize = cg->ksize; g.stride = cg->stride; g.pad = cg->pad;
  g.w_a_stride = cg->w_cin_stride; g.w_b_stride = cg->w_cout_stride;
  const long long M = (long long)g.batch * g.q_h * g.q_w;
  if (M == 0) return NSG_OK;
  const int a_tiles = (g.ksize * g.ksize * g.ca + 63) / 64, b_tiles = (g.cb + 63) / 64;
  // enough pixel chunks to fill the machine ~2x, at least 256 pixels each
  long long want = (2LL * num_sms() * 2 + a_tiles * b_tiles - 1) / (a_tiles * b_tiles);
  if (want < 1) want = 1;
  long long ppc = (M + want - 1) / want;
  if (ppc < 256) ppc = 256;
  ppc = (ppc + 15) / 16 * 16;
  g.pix_per_cta = (int)ppc;
  const unsigned zc = (unsigned)((M + ppc - 1) / ppc);
  Pro pp{NSG_PRO_NONE, nullptr, nullptr}, pq{NSG_PRO_NONE, nullptr, nullptr};
  if (pro_p) pp = Pro{pro_p->mode, pro_p->scale, pro_p->shift};
  if (pro_q) pq = Pro{pro_q->mode, pro_q->scale, pro_q->shift};
  {
    const int served = conv_wgrad_c1(cg, p, pp, q, pq, dw, q_colsum, stream);
    if (served != 0) return served < 0 ? served : NSG_OK;
  }
  const bool vec = (g.ca % 4 == 0);
  return NSG_DISPATCH_DTYPE(cg->dtype, [&]() -> int {
    dim3 grid(a_tiles, b_tiles, zc);
    if (vec)
      wgrad_simt_kernel<T, true><<<grid, 256, 0, stream>>>((const T*)p, (const T*)q, dw, q_colsum, pp, pq, g);
    else
      wgrad_simt_kernel<T, false><<<grid, 256, 0, stream>>>((const T*)p, (const T*)q, dw, q_colsum, pp, pq, g);
    NSG_CHECK_LAUNCH("wgrad_simt_kernel");
    return NSG_OK;
  });
}

}  // namespace nsg
