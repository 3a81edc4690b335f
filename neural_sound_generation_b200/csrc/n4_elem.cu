// Streaming kernels of the reference's other model families:
//   gated activation tanh(a) * sigmoid(b) with the class-conditional bias    /root/reference/src/models.py:218-225, 262-270
//   VAE reparameterisation + KL(N(mu, std) || N(0, 1)) and its backward        /root/reference/src/models.py:103-110
//   summed squared error and its gradient (through the decoder's tanh)        /root/reference/src/loss.py:22-24
//   one categorical draw per row by inverse CDF over softmax(logits)           /root/reference/src/models.py:334-339
#include "common.cuh"

namespace nsg {

__global__ void __launch_bounds__(256)
gated_act_kernel(const float* __restrict__ x, const float* __restrict__ add2, const float* __restrict__ cond,
                 long long pixels, int C, long long ppi, float* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  const long long total = pixels * C;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long m = e / C;
    const int c = (int)(e - m * C);
    float a = __ldg(x + m * 2 * C + c), b = __ldg(x + m * 2 * C + C + c);
    if (add2) {
      a += __ldg(add2 + m * 2 * C + c);
      b += __ldg(add2 + m * 2 * C + C + c);
    }
    if (cond) {
      const long long it = m / ppi;
      a += __ldg(cond + it * 2 * C + c);
      b += __ldg(cond + it * 2 * C + C + c);
    }
    out[e] = tanhf(a) * (1.0f / (1.0f + expf(-b)));
  }
}

__global__ void __launch_bounds__(256)
vae_reparam_kl_kernel(const float* __restrict__ st, const float* __restrict__ eps, long long pixels, int Z,
                      float* __restrict__ z, double* __restrict__ kl_sum) {
  pdl_wait();
  pdl_trigger();
  const long long total = pixels * Z;
  double part = 0.0;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long m = e / Z;
    const int c = (int)(e - m * Z);
    const float mu = __ldg(st + m * 2 * Z + c), lv = __ldg(st + m * 2 * Z + Z + c);
    const float sd = expf(0.5f * lv);
    z[e] = fmaf(sd, __ldg(eps + e), mu);
    // torch.distributions.kl._kl_normal_normal with p = N(mu, sd), q = N(0, 1): 0.5 (var_ratio + t1 - 1 - log var_ratio)
    part += 0.5 * ((double)sd * sd + (double)mu * mu - 1.0 - (double)lv);
  }
  part = warp_sum_d(part);
  __shared__ double red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    double s = 0.0;
    for (int i = 0; i < 8; ++i) s += red[i];
    atomicAdd(kl_sum, s);
  }
}

__global__ void __launch_bounds__(256)
vae_reparam_kl_bwd_kernel(const float* __restrict__ st, const float* __restrict__ eps, const float* __restrict__ dz,
                          float kl_scale, long long pixels, int Z, float* __restrict__ dst) {
  pdl_wait();
  pdl_trigger();
  const long long total = pixels * Z;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < total; e += (long long)gridDim.x * blockDim.x) {
    const long long m = e / Z;
    const int c = (int)(e - m * Z);
    const float mu = __ldg(st + m * 2 * Z + c), lv = __ldg(st + m * 2 * Z + Z + c);
    const float sd = expf(0.5f * lv);
    const float g = dz ? __ldg(dz + e) : 0.f;
    dst[m * 2 * Z + c] = fmaf(kl_scale, mu, g);
    dst[m * 2 * Z + Z + c] = g * __ldg(eps + e) * 0.5f * sd + kl_scale * 0.5f * (sd * sd - 1.0f);
  }
}

__global__ void __launch_bounds__(256)
sqerr_sum_kernel(const float* __restrict__ a, const float* __restrict__ b, long long n, double* __restrict__ sum,
                 float grad_scale, int through_tanh, float* __restrict__ grad) {
  pdl_wait();
  pdl_trigger();
  double part = 0.0;
  for (long long e = (long long)blockIdx.x * blockDim.x + threadIdx.x; e < n; e += (long long)gridDim.x * blockDim.x) {
    const float av = __ldg(a + e), d = av - __ldg(b + e);
    part += (double)d * d;
    if (grad) grad[e] = grad_scale * 2.0f * d * (through_tanh ? 1.0f - av * av : 1.0f);
  }
  part = warp_sum_d(part);
  __shared__ double red[8];
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = part;
  __syncthreads();
  if (threadIdx.x == 0 && sum) {
    double s = 0.0;
    for (int i = 0; i < 8; ++i) s += red[i];
    atomicAdd(sum, s);
  }
}

// one warp per row: max, sum of exp, then the first class whose running probability mass exceeds u
__global__ void __launch_bounds__(256)
softmax_sample_kernel(const float* __restrict__ logits, long long rows, int classes, long long row_stride,
                      long long offset, const float* __restrict__ u, long long* __restrict__ out, long long out_stride) {
  pdl_wait();
  pdl_trigger();
  const int lane = threadIdx.x & 31;
  const long long row = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= rows) return;
  const float* p = logits + row * row_stride + offset;
  float mx = -INFINITY;
  for (int c = lane; c < classes; c += 32) mx = fmaxf(mx, __ldg(p + c));
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  float tot = 0.f;
  for (int c = lane; c < classes; c += 32) tot += expf(__ldg(p + c) - mx);
  tot = warp_sum(tot);
  const float target = __ldg(u + row) * tot;
  // classes in chunks of 32: inclusive scan inside the chunk, running base across chunks
  float base = 0.f;
  int pick = classes - 1;
  bool found = false;
  for (int c0 = 0; c0 < classes && !found; c0 += 32) {
    const int c = c0 + lane;
    float v = c < classes ? expf(__ldg(p + c) - mx) : 0.f;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float t = __shfl_up_sync(0xffffffffu, v, o);
      if (lane >= o) v += t;
    }
    const unsigned hit = __ballot_sync(0xffffffffu, c < classes && base + v > target);
    if (hit) {
      pick = c0 + __ffs(hit) - 1;
      found = true;
    }
    base += __shfl_sync(0xffffffffu, v, 31);
  }
  if (lane == 0) out[row * out_stride] = pick;
}

static int grid_for(long long n) {
  long long b = (n + 255) / 256;
  const long long cap = 8LL * num_sms();
  return (int)(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace nsg

using namespace nsg;

extern "C" int nsg_gated_act(const float* x, const float* add2, const float* cond, long long pixels, int channels,
                             long long pixels_per_item, float* out, void* stream) {
  NSG_REQUIRE(x && out && pixels >= 0 && channels > 0 && pixels_per_item > 0, "gated_act: bad arguments");
  if (pixels == 0) return NSG_OK;
  launch_pdl(gated_act_kernel, dim3(grid_for(pixels * channels)), dim3(256), 0, static_cast<cudaStream_t>(stream), x, add2,
             cond, pixels, channels, pixels_per_item, out);
  NSG_CHECK_LAUNCH("gated_act_kernel");
  return NSG_OK;
}

extern "C" int nsg_vae_reparam_kl(const float* stats, const float* eps, long long pixels, int z_channels, float* z,
                                  double* kl_sum, void* stream) {
  NSG_REQUIRE(stats && eps && z && kl_sum && pixels >= 0 && z_channels > 0, "vae_reparam_kl: bad arguments");
  if (pixels == 0) return NSG_OK;
  launch_pdl(vae_reparam_kl_kernel, dim3(grid_for(pixels * z_channels)), dim3(256), 0, static_cast<cudaStream_t>(stream),
             stats, eps, pixels, z_channels, z, kl_sum);
  NSG_CHECK_LAUNCH("vae_reparam_kl_kernel");
  return NSG_OK;
}

extern "C" int nsg_vae_reparam_kl_bwd(const float* stats, const float* eps, const float* dz, float kl_scale,
                                      long long pixels, int z_channels, float* dstats, void* stream) {
  NSG_REQUIRE(stats && eps && dstats && pixels >= 0 && z_channels > 0, "vae_reparam_kl_bwd: bad arguments");
  if (pixels == 0) return NSG_OK;
  launch_pdl(vae_reparam_kl_bwd_kernel, dim3(grid_for(pixels * z_channels)), dim3(256), 0,
             static_cast<cudaStream_t>(stream), stats, eps, dz, kl_scale, pixels, z_channels, dstats);
  NSG_CHECK_LAUNCH("vae_reparam_kl_bwd_kernel");
  return NSG_OK;
}

extern "C" int nsg_sqerr_sum(const float* a, const float* b, long long n, double* sum, float grad_scale, int through_tanh,
                             float* grad, void* stream) {
  NSG_REQUIRE(a && b && n >= 0 && (sum || grad), "sqerr_sum: bad arguments");
  if (n == 0) return NSG_OK;
  launch_pdl(sqerr_sum_kernel, dim3(grid_for(n)), dim3(256), 0, static_cast<cudaStream_t>(stream), a, b, n, sum, grad_scale,
             through_tanh, grad);
  NSG_CHECK_LAUNCH("sqerr_sum_kernel");
  return NSG_OK;
}

extern "C" int nsg_softmax_sample(const float* logits, long long rows, int classes, long long row_stride, long long offset,
                                  const float* u, long long* out, long long out_stride, void* stream) {
  NSG_REQUIRE(logits && u && out && rows >= 0 && classes > 0, "softmax_sample: bad arguments");
  if (rows == 0) return NSG_OK;
  launch_pdl(softmax_sample_kernel, dim3((unsigned)((rows + 7) / 8)), dim3(256), 0, static_cast<cudaStream_t>(stream),
             logits, rows, classes, row_stride, offset, u, out, out_stride);
  NSG_CHECK_LAUNCH("softmax_sample_kernel");
  return NSG_OK;
}
