// General stride-1 convolutions with RECTANGULAR kernels, any channel counts, any zero padding (fp32, CUDA cores):
// the layers of the reference's other model families that the square <= 4 x 4 kernels of
// conv_simt.cu / conv_tma.cu do not cover --
//   /root/reference/src/models.py:74,85     VAE: Conv2d(dim, dim, 5, 1, 0), ConvTranspose2d(dim, dim, 5, 1, 0)
//   /root/reference/src/models.py:241-253   GatedMaskedConv2d: vertical stack (k//2+1, k) pad (k//2, k//2), horizontal stack
//                                           (1, k//2+1) pad (0, k//2), k = 7 and 3
// One implicit-GEMM kernel serves nn.Conv2d forward / nn.ConvTranspose2d data gradient (gather 0: ih = oh - pad_h + kh)
// and nn.ConvTranspose2d forward / nn.Conv2d data gradient (gather 1: ih = oh + pad_h - kh); one kernel the weight
// gradients.  The weight tensor is read in the reference's own layout through (tap, cin, cout) strides, as everywhere
// in this library.  64 x 64 output tile, K chunks of 16, 4 x 4 outputs per thread; weight-gradient partials over pixel
// ranges meet through float atomics.
#include "common.cuh"

namespace nsg {

struct RectGeom {
  int batch, in_h, in_w, cin;
  int out_h, out_w, cout;
  int kh, kw, pad_h, pad_w, gather;
  long long w_cin_stride, w_cout_stride;
};

constexpr int RB = 64, RK = 16;

__global__ void __launch_bounds__(256)
rect_gemm_kernel(const float* __restrict__ in, const float* __restrict__ w, const float* __restrict__ bias,
                 const float* __restrict__ add, float* __restrict__ out, double* __restrict__ stats, RectGeom g, int act) {
  __shared__ __align__(16) float As[RK][RB + 4];
  __shared__ __align__(16) float Bs[RK][RB];
  __shared__ int rowN[RB], rowH[RB], rowW[RB];
  pdl_wait();
  pdl_trigger();
  const int tid = threadIdx.x;
  const long long M = (long long)g.batch * g.out_h * g.out_w;
  const long long m0 = (long long)blockIdx.x * RB;
  const int n0 = blockIdx.y * RB;
  const int K = g.kh * g.kw * g.cin;
  if (tid < RB) {
    const long long m = m0 + tid;
    if (m < M) {
      rowW[tid] = (int)(m % g.out_w);
      const long long t = m / g.out_w;
      rowH[tid] = (int)(t % g.out_h);
      rowN[tid] = (int)(t / g.out_h);
    } else {
      rowN[tid] = -1;
      rowH[tid] = rowW[tid] = 0;
    }
  }
  __syncthreads();
  const int ak = tid & 15, am = tid >> 4;          // A loader: channel-fastest (coalesced within a tap)
  const int bn = tid & 63, bk = tid >> 6;          // B loader
  const int tx = tid & 15, ty = tid >> 4;
  float acc[4][4] = {};
  for (int k0 = 0; k0 < K; k0 += RK) {
    {
      const int k = k0 + ak;
      int tap = 0, c = 0, dh = 0, dw = 0;
      if (k < K) {
        tap = k / g.cin;
        c = k - tap * g.cin;
        const int th = tap / g.kw, tw = tap - th * g.kw;
        dh = g.gather ? g.pad_h - th : th - g.pad_h;
        dw = g.gather ? g.pad_w - tw : tw - g.pad_w;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = am + 16 * j;
        float v = 0.f;
        const int n = rowN[r];
        if (k < K && n >= 0) {
          const int ih = rowH[r] + dh, iw = rowW[r] + dw;
          if (ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w)
            v = __ldg(in + (((long long)n * g.in_h + ih) * g.in_w + iw) * g.cin + c);
        }
        As[ak][r] = v;
      }
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int kk = bk + 4 * j, k = k0 + kk, co = n0 + bn;
      float v = 0.f;
      if (k < K && co < g.cout) {
        const int tap = k / g.cin, c = k - tap * g.cin;
        v = __ldg(w + tap + (long long)c * g.w_cin_stride + (long long)co * g.w_cout_stride);
      }
      Bs[kk][bn] = v;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < RK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
  float s1[4] = {}, s2[4] = {};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long m = m0 + ty * 4 + i;
    if (m >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co >= g.cout) continue;
      float v = acc[i][j] + (bias ? __ldg(bias + co) : 0.f);
      if (add) v += __ldg(add + m * g.cout + co);
      s1[j] += v;
      s2[j] = fmaf(v, v, s2[j]);
      if (act == 1) v = tanhf(v);
      if (act == 2) v = fmaxf(v, 0.f);
      out[m * g.cout + co] = v;
    }
  }
  if (stats) {      // BatchNorm moments of the (pre-activation) output: stats[c] += sum v, stats[cout + c] += sum v^2
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = n0 + tx * 4 + j;
      if (co < g.cout) {
        atomicAdd(stats + co, (double)s1[j]);
        atomicAdd(stats + g.cout + co, (double)s2[j]);
      }
    }
  }
}

// dw[tap + a*w_a_stride + b*w_b_stride] += sum_m P[gather(m, tap), a] * Q[m, b]   (m over Q's pixels; gather 0 only:
// the caller swaps the operands for a transposed convolution, exactly as nsg_conv_wgrad documents)
__global__ void __launch_bounds__(256)
rect_wgrad_kernel(const float* __restrict__ P, const float* __restrict__ Q, float* __restrict__ dw,
                  float* __restrict__ q_colsum, RectGeom g, int pix_per_cta) {
  __shared__ __align__(16) float Ps[RK][RB + 4];
  __shared__ __align__(16) float Qs[RK][RB];
  pdl_wait();
  pdl_trigger();
  const int tid = threadIdx.x;
  const int A_total = g.kh * g.kw * g.cin;           // rows: (tap, P channel)
  const int a0 = blockIdx.x * RB, b0 = blockIdx.y * RB;
  const long long M = (long long)g.batch * g.out_h * g.out_w;
  const long long mbeg = (long long)blockIdx.z * pix_per_cta;
  const long long mend = mbeg + pix_per_cta < M ? mbeg + pix_per_cta : M;
  const int lc = tid & 63, lm = tid >> 6;            // loaders: channel-fastest
  const int tx = tid & 15, ty = tid >> 4;
  const int a = a0 + lc;
  int tap = 0, c = 0, dh = 0, dwc = 0;
  if (a < A_total) {
    tap = a / g.cin;
    c = a - tap * g.cin;
    const int th = tap / g.kw, tw = tap - th * g.kw;
    dh = th - g.pad_h;
    dwc = tw - g.pad_w;
  }
  float acc[4][4] = {};
  float qs = 0.f;
  for (long long mb = mbeg; mb < mend; mb += RK) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int mm = lm + 4 * j;
      const long long m = mb + mm;
      float pv = 0.f, qv = 0.f;
      if (m < mend) {
        const int ow = (int)(m % g.out_w);
        const long long t = m / g.out_w;
        const int oh = (int)(t % g.out_h), n = (int)(t / g.out_h);
        if (a < A_total) {
          const int ih = oh + dh, iw = ow + dwc;
          if (ih >= 0 && ih < g.in_h && iw >= 0 && iw < g.in_w)
            pv = __ldg(P + (((long long)n * g.in_h + ih) * g.in_w + iw) * g.cin + c);
        }
        if (b0 + lc < g.cout) qv = __ldg(Q + m * g.cout + b0 + lc);
      }
      Ps[mm][lc] = pv;
      Qs[mm][lc] = qv;
      if (blockIdx.x == 0) qs += qv;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < RK; ++kk) {
      const float4 pa = *reinterpret_cast<const float4*>(&Ps[kk][ty * 4]);
      const float4 qb = *reinterpret_cast<const float4*>(&Qs[kk][tx * 4]);
      const float av[4] = {pa.x, pa.y, pa.z, pa.w}, bv[4] = {qb.x, qb.y, qb.z, qb.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int aa = a0 + ty * 4 + i;
    if (aa >= A_total) continue;
    const int t2 = aa / g.cin, c2 = aa - t2 * g.cin;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int b = b0 + tx * 4 + j;
      if (b < g.cout) atomicAdd(dw + t2 + (long long)c2 * g.w_cin_stride + (long long)b * g.w_cout_stride, acc[i][j]);
    }
  }
  if (q_colsum && blockIdx.x == 0 && b0 + lc < g.cout) atomicAdd(q_colsum + b0 + lc, qs);
}

static int make_rect(const nsg_rect_geom* r, RectGeom* out, const char* what) {
  NSG_REQUIRE(r, "%s: null geometry", what);
  NSG_REQUIRE(r->batch >= 0 && r->in_h > 0 && r->in_w > 0 && r->cin > 0 && r->out_h > 0 && r->out_w > 0 && r->cout > 0,
              "%s: bad tensor extents", what);
  NSG_REQUIRE(r->kh >= 1 && r->kw >= 1 && r->kh * r->kw <= 4096 && r->pad_h >= 0 && r->pad_w >= 0,
              "%s: kernel %d x %d / padding %d, %d", what, r->kh, r->kw, r->pad_h, r->pad_w);
  NSG_REQUIRE(r->gather == 0 || r->gather == 1, "%s: gather %d", what, r->gather);
  const int eh = r->gather ? r->in_h + r->kh - 1 - 2 * r->pad_h : r->in_h - r->kh + 1 + 2 * r->pad_h;
  const int ew = r->gather ? r->in_w + r->kw - 1 - 2 * r->pad_w : r->in_w - r->kw + 1 + 2 * r->pad_w;
  NSG_REQUIRE(r->out_h <= eh && r->out_w <= ew && r->out_h > 0 && r->out_w > 0,
              "%s: output %d x %d exceeds what a stride-1 %d x %d kernel with padding (%d, %d) yields from %d x %d (%d x %d)",
              what, r->out_h, r->out_w, r->kh, r->kw, r->pad_h, r->pad_w, r->in_h, r->in_w, eh, ew);
  RectGeom g{r->batch, r->in_h, r->in_w, r->cin, r->out_h, r->out_w, r->cout, r->kh, r->kw, r->pad_h, r->pad_w, r->gather,
             r->w_cin_stride, r->w_cout_stride};
  *out = g;
  return NSG_OK;
}

}  // namespace nsg

using namespace nsg;

extern "C" int nsg_conv_rect(const nsg_rect_geom* geom, const float* in, const float* w, const float* bias,
                             const float* add_src, int act, double* stats, float* out, void* stream) {
  RectGeom g;
  if (int rc = make_rect(geom, &g, "conv_rect")) return rc;
  NSG_REQUIRE(in && w && out, "conv_rect: null pointer");
  NSG_REQUIRE(act >= 0 && act <= 2, "conv_rect: output activation %d (0 none, 1 tanh, 2 ReLU)", act);
  NSG_REQUIRE(!stats || act == 0, "conv_rect: BatchNorm moments are those of the un-activated output");
  const long long M = (long long)g.batch * g.out_h * g.out_w;
  if (M == 0) return NSG_OK;
  dim3 grid((unsigned)((M + RB - 1) / RB), (unsigned)((g.cout + RB - 1) / RB));
  launch_pdl(rect_gemm_kernel, grid, dim3(256), 0, static_cast<cudaStream_t>(stream), in, w, bias, add_src, out, stats, g, act);
  NSG_CHECK_LAUNCH("rect_gemm_kernel");
  return NSG_OK;
}

extern "C" int nsg_conv_rect_wgrad(const nsg_rect_geom* geom, const float* p, const float* q, float* dw, float* q_colsum,
                                   void* stream) {
  RectGeom g;
  if (int rc = make_rect(geom, &g, "conv_rect_wgrad")) return rc;
  NSG_REQUIRE(g.gather == 0, "conv_rect_wgrad: gather must be 0 (swap the operands for a transposed convolution)");
  NSG_REQUIRE(p && q && dw, "conv_rect_wgrad: null pointer");
  const long long M = (long long)g.batch * g.out_h * g.out_w;
  if (M == 0) return NSG_OK;
  const int a_tiles = (g.kh * g.kw * g.cin + RB - 1) / RB, b_tiles = (g.cout + RB - 1) / RB;
  long long want = (4LL * num_sms() + (long long)a_tiles * b_tiles - 1) / ((long long)a_tiles * b_tiles);
  if (want < 1) want = 1;
  long long ppc = (M + want - 1) / want;
  if (ppc < 128) ppc = 128;
  ppc = (ppc + RK - 1) / RK * RK;
  const unsigned zc = (unsigned)((M + ppc - 1) / ppc);
  NSG_REQUIRE(zc <= 65535, "conv_rect_wgrad: too many pixel chunks");
  launch_pdl(rect_wgrad_kernel, dim3(a_tiles, b_tiles, zc), dim3(256), 0, static_cast<cudaStream_t>(stream), p, q, dw,
             q_colsum, g, (int)ppc);
  NSG_CHECK_LAUNCH("rect_wgrad_kernel");
  return NSG_OK;
}
