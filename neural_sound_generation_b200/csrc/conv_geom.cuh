// Geometry shared by the CUDA-core and tcgen05 convolution kernels: one "phase" = one launch over an
// output sub-grid whose taps all exist (stride phases of a transposed-like gather), see include/nsg_b200.h.
#pragma once
#include "common.cuh"

namespace nsg {

struct TapGeom {
  int batch, in_h, in_w, cin;
  int sub_h, sub_w;               // output sub-grid iterated by this launch (one stride phase)
  int full_h, full_w, cout;       // output tensor
  int ntaps;
  int in_stride;                  // ih = ohs*in_stride + dh[t]
  int out_stride, off_h, off_w;   // oh = ohs*out_stride + off_h
  int dh[16], dw[16], woff[16];
  long long w_cin_stride, w_cout_stride;
};

struct EpiArgs {
  const float* bias;
  const void* add_src;
  int mask_mode;
  const void* mask_src;
  const float* mscale;
  const float* mshift;
  int stats_mode;
  double* stats;
  const void* xhat_src;
  const float* mean;
  const float* invstd;
  int tanh_out;
};

static inline int build_phases(const nsg_conv_geom* cg, TapGeom* phases) {
  const int k = cg->ksize, s = cg->stride, p = cg->pad;
  TapGeom base{};
  base.batch = cg->batch; base.in_h = cg->in_h; base.in_w = cg->in_w; base.cin = cg->cin;
  base.full_h = cg->out_h; base.full_w = cg->out_w; base.cout = cg->cout;
  base.w_cin_stride = cg->w_cin_stride; base.w_cout_stride = cg->w_cout_stride;
  if (cg->gather == 0) {
    TapGeom g = base;
    g.sub_h = cg->out_h; g.sub_w = cg->out_w;
    g.in_stride = s; g.out_stride = 1; g.off_h = 0; g.off_w = 0;
    g.ntaps = k * k;
    for (int kh = 0; kh < k; ++kh)
      for (int kw = 0; kw < k; ++kw) {
        g.dh[kh * k + kw] = kh - p;
        g.dw[kh * k + kw] = kw - p;
        g.woff[kh * k + kw] = kh * k + kw;
      }
    phases[0] = g;
    return 1;
  }
  int np = 0;
  for (int ph = 0; ph < s; ++ph)
    for (int pw = 0; pw < s; ++pw) {
      TapGeom g = base;
      g.sub_h = (cg->out_h - ph + s - 1) / s;
      g.sub_w = (cg->out_w - pw + s - 1) / s;
      if (g.sub_h <= 0 || g.sub_w <= 0) continue;
      g.in_stride = 1; g.out_stride = s; g.off_h = ph; g.off_w = pw;
      int nt = 0;
      for (int kh = 0; kh < k; ++kh) {
        if (((ph + p - kh) % s + s) % s != 0) continue;
        for (int kw = 0; kw < k; ++kw) {
          if (((pw + p - kw) % s + s) % s != 0) continue;
          // exact division (numerator is a multiple of s, possibly negative)
          g.dh[nt] = (ph + p - kh) / s;
          g.dw[nt] = (pw + p - kw) / s;
          g.woff[nt] = kh * k + kw;
          ++nt;
        }
      }
      g.ntaps = nt;
      phases[np++] = g;
    }
  return np;
}


// validated copy of the public epilogue struct
static inline int make_epi(const nsg_epilogue* epi, EpiArgs* out) {
  EpiArgs e{};
  if (epi) {
    e.bias = epi->bias; e.add_src = epi->add_src; e.mask_mode = epi->mask_mode; e.mask_src = epi->mask_src;
    e.mscale = epi->mscale; e.mshift = epi->mshift; e.stats_mode = epi->stats_mode; e.stats = epi->stats;
    e.xhat_src = epi->xhat_src; e.mean = epi->mean; e.invstd = epi->invstd; e.tanh_out = epi->tanh_out;
  }
  if (e.mask_mode != NSG_MASK_NONE) NSG_REQUIRE(e.mask_src, "conv: mask_src missing");
  if (e.mask_mode == NSG_MASK_AFFINE) NSG_REQUIRE(e.mscale && e.mshift, "conv: mask affine coefficients missing");
  if (e.stats_mode != NSG_STATS_NONE) NSG_REQUIRE(e.stats, "conv: stats buffer missing");
  if (e.stats_mode == NSG_STATS_BNBWD)
    NSG_REQUIRE(e.xhat_src && e.mean && e.invstd, "conv: BN-bwd stats operands missing");
  *out = e;
  return NSG_OK;
}

}  // namespace nsg
