// Streaming companions of the TMA convolution kernels for bf16 NHWC tensors with 64 channels: they produce the
// OPERAND-READY activations the tensor-core kernels read (models.py:148-183):
//   bn_act        v = [relu]( [res +] gamma*(h - mean)*invstd + beta ), BatchNorm2d statistics finalised in-kernel
//                 from the (sum, sum of squares) the convolution epilogue accumulated (train) or from the running
//                 statistics (eval); block 0 also publishes mean / invstd / scale / shift and updates the running
//                 statistics exactly as torch does (momentum, unbiased variance, num_batches_tracked)
//   bn_bwd        dh = gamma*invstd*(g - sum_g/M - xhat*sum_gx/M); dgamma += sum_gx; dbeta += sum_g;
//                 dbias_prev += column sums of dh (the gradient of the bias of the convolution that feeds the BN)
//   relu          out = max(x, 0)
//   vq_bwd_stats  dz = g_st + coef*(z - zq) with the BatchNorm-backward reductions of the following layer fused
// One thread owns 8 consecutive channels (16-byte accesses) of a grid-strided set of pixels, so per-channel
// coefficients live in registers and per-channel reductions need two shuffles, one shared-memory and one global atomic.
#include "common.cuh"
#define NSG_TU_ID 4
#include "peer.cuh"

namespace nsg {

constexpr int FE_C = 64;

// Synchronised BatchNorm without a collective and without an extra launch: CTA 0 swaps the layer's 128 partial sums
// with the other GPUs over peer memory (peer.cuh) while every CTA's tensor loads are already in flight, then opens
// `gate` (one zeroed counter per launch); the other CTAs wait at it before deriving their coefficients.  CTA 0 is
// dispatched first, so the waiters cannot keep it from becoming resident.
__device__ __forceinline__ void fe_global_stats(const PeerSync& peer, double* stats, unsigned int* gate) {
  if (peer.world <= 1) return;
  if (blockIdx.x == 0) {
    peer_exchange(peer, stats, (int)threadIdx.x, 1, (int)blockDim.x);
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(gate), "r"(1u) : "memory");
    }
  } else {
    if (threadIdx.x == 0) {
      unsigned int v = 0;
      const unsigned long long t0 = global_ns();
      for (;;) {
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(gate) : "memory");
        if (v != 0) break;
        if (global_ns() - t0 > 30 * kHangNs) hang_trap(__LINE__, 0u);
        __nanosleep(64);
      }
    }
    __syncthreads();
  }
}

__device__ __forceinline__ void fe_ld8(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ void fe_unpack(const uint4& r, float (&v)[8]) {
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}
__device__ __forceinline__ uint32_t fe_pack(float a, float b) {
  __nv_bfloat162 t = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ void fe_st8(__nv_bfloat16* p, const float (&v)[8]) {
  *reinterpret_cast<uint4*>(p) =
      make_uint4(fe_pack(v[0], v[1]), fe_pack(v[2], v[3]), fe_pack(v[4], v[5]), fe_pack(v[6], v[7]));
}
// sum over the 32 pixel lanes of a CTA for each of the thread's 8 channels; result added to dst[c0 + i] by one thread
template <typename OutT>
__device__ __forceinline__ void fe_channel_reduce(float (&s)[8], float* sred, OutT* dst, int c0) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    s[i] += __shfl_xor_sync(0xffffffffu, s[i], 8);
    s[i] += __shfl_xor_sync(0xffffffffu, s[i], 16);
  }
  const int tid = threadIdx.x;
  if (tid < FE_C) sred[tid] = 0.f;
  __syncthreads();
  if ((tid & 31) < 8) {
#pragma unroll
    for (int i = 0; i < 8; ++i) atomicAdd(&sred[c0 + i], s[i]);
  }
  __syncthreads();
  if (tid < FE_C) atomicAdd(dst + tid, (OutT)sred[tid]);
  __syncthreads();
}

// two statistics at once, without shared-memory atomics: lanes 0-7 of every warp hold the warp's sums of 8 channels
// each after two shuffles; they go to [statistic][warp][channel] and 128 threads add the 8 warps in a FIXED order
// (one barrier instead of six, and the CTA's contribution no longer depends on the order in which its warps arrive)
template <typename OutT>
__device__ __forceinline__ void fe_channel_reduce2(float (&a)[8], float (&b)[8], float* sred2 /* [2][8][FE_C] */,
                                                   OutT* dst /* [2 * FE_C] */, int c0) {
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    a[i] += __shfl_xor_sync(0xffffffffu, a[i], 8);
    a[i] += __shfl_xor_sync(0xffffffffu, a[i], 16);
    b[i] += __shfl_xor_sync(0xffffffffu, b[i], 8);
    b[i] += __shfl_xor_sync(0xffffffffu, b[i], 16);
  }
  const int tid = threadIdx.x, warp = tid >> 5;
  if ((tid & 31) < 8) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      sred2[(0 * 8 + warp) * FE_C + c0 + i] = a[i];
      sred2[(1 * 8 + warp) * FE_C + c0 + i] = b[i];
    }
  }
  __syncthreads();
  if (tid < 2 * FE_C) {
    const int which = tid >> 6, c = tid & (FE_C - 1);
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += sred2[(which * 8 + w) * FE_C + c];
    atomicAdd(dst + which * FE_C + c, (OutT)t);
  }
}

struct BnActArgs {
  const __nv_bfloat16* h;
  const __nv_bfloat16* res;
  __nv_bfloat16* out;
  const double* stats;           // train: [sum(64) | sumsq(64)]; null = eval (running statistics)
  double count;
  const float* gamma;
  const float* beta;
  float eps, momentum;
  float* running_mean;
  float* running_var;
  long long* nbt;
  float* mean;                   // published coefficients (nullable)
  float* invstd;
  float* scale;
  float* shift;
  int relu_out;
  long long pixels;
  PeerSync peer;                 // world > 1: the statistics are summed over the GPUs inside this launch
  unsigned int* gate;
  double inv_count;              // 1 / count from the host: a double-precision division is a ~100-instruction sequence, and
                                 // the whole CTA waits at a barrier for the 64 threads that derive the coefficients
  double unbias;                 // count / (count - 1), or 1
};

__global__ void __launch_bounds__(256) bn_act_kernel(const BnActArgs a) {
  __shared__ float sSc[FE_C], sSh[FE_C];
  const int tid = threadIdx.x;
  pdl_wait();
  pdl_trigger();
  const int c0 = (tid & 7) * 8;
  // The first 4 pixels of this thread are requested BEFORE the coefficients are derived: the statistics load, the
  // fp64 arithmetic behind mean / invstd and the CTA barrier then run under the latency of the tensor loads instead
  // of in front of it (at configs[1] every CTA has exactly one such batch: the kernel is one memory round trip).
  long long p0 = (long long)blockIdx.x * 128 + (tid >> 3);
  uint4 hv[4], rv[4];
  auto request = [&](long long q0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long p = q0 + 32 * u;
      if (p < a.pixels) {
        hv[u] = __ldg(reinterpret_cast<const uint4*>(a.h + p * FE_C + c0));
        if (a.res) rv[u] = __ldg(reinterpret_cast<const uint4*>(a.res + p * FE_C + c0));
      }
    }
  };
  request(p0);
  if (a.stats) fe_global_stats(a.peer, const_cast<double*>(a.stats), a.gate);
  if (tid < FE_C) {
    float mu, is;
    if (a.stats) {
      const double m = __ldcg(a.stats + tid) * a.inv_count;
      double var = __ldcg(a.stats + FE_C + tid) * a.inv_count - m * m;
      if (var < 0.0) var = 0.0;
      mu = (float)m;
      // (the moments and their difference in double -- cancellation --, the reciprocal square root in single precision:
      // a double sqrt + division is a few hundred dependent instructions in front of the CTA's barrier)
      is = 1.0f / sqrtf((float)(var + (double)a.eps));
      if (blockIdx.x == 0 && a.running_mean) {
        const double unbiased = var * a.unbias;
        a.running_mean[tid] = (1.f - a.momentum) * a.running_mean[tid] + a.momentum * mu;
        a.running_var[tid] = (1.f - a.momentum) * a.running_var[tid] + a.momentum * (float)unbiased;
        if (tid == 0 && a.nbt) *a.nbt += 1;
      }
    } else {
      mu = a.running_mean[tid];
      is = 1.0f / sqrtf(a.running_var[tid] + a.eps);
    }
    const float sc = a.gamma[tid] * is;
    const float sh = a.beta[tid] - mu * sc;
    sSc[tid] = sc;
    sSh[tid] = sh;
    if (blockIdx.x == 0 && a.mean) {
      a.mean[tid] = mu;
      a.invstd[tid] = is;
      a.scale[tid] = sc;
      a.shift[tid] = sh;
    }
  }
  __syncthreads();
  float sc[8], sh[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    sc[i] = sSc[c0 + i];
    sh[i] = sSh[c0 + i];
  }
  for (; p0 < a.pixels; p0 += (long long)gridDim.x * 128) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long p = p0 + 32 * u;
      if (p >= a.pixels) break;
      float v[8], r[8];
      fe_unpack(hv[u], v);
      if (a.res) fe_unpack(rv[u], r);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        float x = fmaf(sc[i], v[i], sh[i]);
        if (a.res) x += r[i];
        v[i] = a.relu_out ? fmaxf(x, 0.f) : x;
      }
      fe_st8(a.out + p * FE_C + c0, v);
    }
    request(p0 + (long long)gridDim.x * 128);          // next batch (none at configs[1])
  }
}

struct BnBwdArgs {
  const __nv_bfloat16* g;
  const __nv_bfloat16* h;
  __nv_bfloat16* dh;
  const double* stats;           // [sum g (64) | sum g*xhat (64)]
  double count;
  const float* mean;
  const float* invstd;
  const float* gamma;
  float* dgamma;
  float* dbeta;
  float* dbias_prev;             // nullable
  long long pixels;
  PeerSync peer;
  unsigned int* gate;
  double inv_count;              // 1 / count (see BnActArgs)
};

__global__ void __launch_bounds__(256) bn_bwd_kernel(const BnBwdArgs a) {
  __shared__ float sred[FE_C];
  __shared__ float sA[FE_C], sB[FE_C], sC[FE_C];     // dh = A*g + B*h + C per channel
  const int tid = threadIdx.x;
  pdl_wait();
  pdl_trigger();
  const int c0 = (tid & 7) * 8;
  // (first batch of tensor loads in flight before the coefficients are derived: see bn_act_kernel)
  long long p0 = (long long)blockIdx.x * 128 + (tid >> 3);
  uint4 gr[4], hr[4];
  auto request = [&](long long q0) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long p = q0 + 32 * u;
      if (p < a.pixels) {
        gr[u] = __ldg(reinterpret_cast<const uint4*>(a.g + p * FE_C + c0));
        hr[u] = __ldg(reinterpret_cast<const uint4*>(a.h + p * FE_C + c0));
      }
    }
  };
  request(p0);
  fe_global_stats(a.peer, const_cast<double*>(a.stats), a.gate);
  if (tid < FE_C) {
    const double sg = __ldcg(a.stats + tid) * a.inv_count, sgx = __ldcg(a.stats + FE_C + tid) * a.inv_count;
    const float mu = a.mean[tid], is = a.invstd[tid];
    const float k0 = a.gamma[tid] * is;
    const float kb = k0 * is * (float)sgx;
    sA[tid] = k0;
    sB[tid] = -kb;
    sC[tid] = kb * mu - k0 * (float)sg;
    if (blockIdx.x == 0) {
      if (a.dbeta) a.dbeta[tid] += (float)__ldcg(a.stats + tid);
      if (a.dgamma) a.dgamma[tid] += (float)__ldcg(a.stats + FE_C + tid);
    }
  }
  __syncthreads();
  float ca[8], cb[8], cc[8], col[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    ca[i] = sA[c0 + i];
    cb[i] = sB[c0 + i];
    cc[i] = sC[c0 + i];
    col[i] = 0.f;
  }
  for (; p0 < a.pixels; p0 += (long long)gridDim.x * 128) {
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long p = p0 + 32 * u;
      if (p >= a.pixels) break;
      float gv[8], hv[8];
      fe_unpack(gr[u], gv);
      fe_unpack(hr[u], hv);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        gv[i] = fmaf(ca[i], gv[i], fmaf(cb[i], hv[i], cc[i]));
        col[i] += gv[i];
      }
      fe_st8(a.dh + p * FE_C + c0, gv);
    }
    request(p0 + (long long)gridDim.x * 128);
  }
  if (a.dbias_prev) fe_channel_reduce<float>(col, sred, a.dbias_prev, c0);
}

__global__ void __launch_bounds__(256)
relu_kernel(const __nv_bfloat16* __restrict__ x, __nv_bfloat16* __restrict__ out, long long nvec) {
  pdl_wait();
  pdl_trigger();
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (long long)gridDim.x * 256) {
    float v[8];
    fe_ld8(x + i * 8, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = fmaxf(v[j], 0.f);
    fe_st8(out + i * 8, v);
  }
}

// dz = g_st + coef*(z - zq);  stats[c] += sum dz, stats[64 + c] += sum dz * (h - mean)*invstd
__global__ void __launch_bounds__(256)
vq_bwd_stats_kernel(const __nv_bfloat16* __restrict__ g_st, const __nv_bfloat16* __restrict__ z,
                    const __nv_bfloat16* __restrict__ zq, float coef, __nv_bfloat16* __restrict__ dz,
                    const __nv_bfloat16* __restrict__ h, const float* __restrict__ mean,
                    const float* __restrict__ invstd, double* __restrict__ stats, long long pixels) {
  pdl_wait();
  pdl_trigger();
  __shared__ float sred2[2 * 8 * FE_C];
  const int tid = threadIdx.x;
  const int c0 = (tid & 7) * 8;
  float mu[8], is[8], s1[8], s2[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    mu[i] = stats ? __ldg(mean + c0 + i) : 0.f;
    is[i] = stats ? __ldg(invstd + c0 + i) : 0.f;
    s1[i] = s2[i] = 0.f;
  }
  // 4 pixels per thread and iteration, every load of the iteration requested before the first use (one memory round
  // trip per 128 pixels of a CTA instead of four)
  for (long long p0 = (long long)blockIdx.x * 128 + (tid >> 3); p0 < pixels; p0 += (long long)gridDim.x * 128) {
    uint4 rz[4], rq[4], rg[4], rh[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long p = p0 + 32 * u;
      if (p < pixels) {
        const long long o = p * FE_C + c0;
        rz[u] = __ldg(reinterpret_cast<const uint4*>(z + o));
        if (zq) rq[u] = __ldg(reinterpret_cast<const uint4*>(zq + o));
        if (g_st) rg[u] = __ldg(reinterpret_cast<const uint4*>(g_st + o));
        if (stats) rh[u] = __ldg(reinterpret_cast<const uint4*>(h + o));
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long p = p0 + 32 * u;
      if (p >= pixels) break;
      const long long o = p * FE_C + c0;
      float gv[8], zv[8], qv[8], hv[8];
      fe_unpack(rz[u], zv);
      if (zq) fe_unpack(rq[u], qv);
      if (g_st) fe_unpack(rg[u], gv);
      if (stats) fe_unpack(rh[u], hv);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float d = (g_st ? gv[i] : 0.f) + coef * (zv[i] - (zq ? qv[i] : 0.f));
        gv[i] = d;
        if (stats) {
          s1[i] += d;
          s2[i] = fmaf(d, (hv[i] - mu[i]) * is[i], s2[i]);
        }
      }
      fe_st8(dz + o, gv);
    }
  }
  if (stats) fe_channel_reduce2<double>(s1, s2, sred2, stats, c0);
}

static inline int fe_grid(long long pixels) {
  long long blocks = (pixels + 127) / 128;
  const long long cap = (long long)num_sms() * 4;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : (int)blocks;
}

}  // namespace nsg

namespace nsg {
int fused_elem_hang_init(unsigned int* host_mapped_words) {
  cudaError_t e = cudaMemcpyToSymbol(tc::g_hang_words, &host_mapped_words, sizeof(host_mapped_words));
  return e == cudaSuccess ? NSG_OK : cuda_fail(e, "cudaMemcpyToSymbol(g_hang_words)");
}
}  // namespace nsg

using namespace nsg;

extern "C" {

int nsg_bn_act64_sync(const void* h, const void* res, void* out, double* stats, double count, const float* gamma,
                      const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                      long long* num_batches_tracked, float* mean, float* invstd, float* scale, float* shift,
                      int relu_out, long long pixels, const nsg_peer_sync* peer, unsigned int* gate, void* stream);

int nsg_bn_act64(const void* h, const void* res, void* out, const double* stats, double count, const float* gamma,
                 const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                 long long* num_batches_tracked, float* mean, float* invstd, float* scale, float* shift, int relu_out,
                 long long pixels, void* stream) {
  return nsg_bn_act64_sync(h, res, out, const_cast<double*>(stats), count, gamma, beta, eps, momentum, running_mean,
                           running_var, num_batches_tracked, mean, invstd, scale, shift, relu_out, pixels, nullptr,
                           nullptr, stream);
}

int nsg_bn_act64_sync(const void* h, const void* res, void* out, double* stats, double count, const float* gamma,
                      const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                      long long* num_batches_tracked, float* mean, float* invstd, float* scale, float* shift,
                      int relu_out, long long pixels, const nsg_peer_sync* peer, unsigned int* gate, void* stream) {
  NSG_REQUIRE(h && out && gamma && beta, "bn_act64: null tensor");
  NSG_REQUIRE(stats || (running_mean && running_var), "bn_act64: eval mode needs the running statistics");
  NSG_REQUIRE(!stats || count > 0, "bn_act64: count must be positive");
  if (pixels == 0) return NSG_OK;
  BnActArgs a{static_cast<const __nv_bfloat16*>(h), static_cast<const __nv_bfloat16*>(res),
              static_cast<__nv_bfloat16*>(out), stats, count, gamma, beta, eps, momentum,
              stats ? running_mean : running_mean, stats ? running_var : running_var, num_batches_tracked,
              mean, invstd, scale, shift, relu_out, pixels, PeerSync{}, gate, 0.0, 1.0};
  a.inv_count = count > 0 ? 1.0 / count : 0.0;
  a.unbias = count > 1.0 ? count / (count - 1.0) : 1.0;
  a.peer.world = 1;
  if (peer && peer->world > 1) {
    NSG_REQUIRE(stats && gate, "bn_act64_sync: the exchange needs the statistics and a zeroed gate counter");
    if (int rc = make_peer(peer, &a.peer)) return rc;
  }
  launch_pdl(bn_act_kernel, dim3(fe_grid(pixels)), dim3(256), 0, static_cast<cudaStream_t>(stream), a);
  NSG_CHECK_LAUNCH("bn_act_kernel");
  return NSG_OK;
}

int nsg_bn_bwd64_sync(const void* g, const void* h, void* dh, double* stats, double count, const float* mean,
                      const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* dbias_prev,
                      long long pixels, const nsg_peer_sync* peer, unsigned int* gate, void* stream);

int nsg_bn_bwd64(const void* g, const void* h, void* dh, const double* stats, double count, const float* mean,
                 const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* dbias_prev,
                 long long pixels, void* stream) {
  return nsg_bn_bwd64_sync(g, h, dh, const_cast<double*>(stats), count, mean, invstd, gamma, dgamma, dbeta, dbias_prev,
                           pixels, nullptr, nullptr, stream);
}

int nsg_bn_bwd64_sync(const void* g, const void* h, void* dh, double* stats, double count, const float* mean,
                      const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* dbias_prev,
                      long long pixels, const nsg_peer_sync* peer, unsigned int* gate, void* stream) {
  NSG_REQUIRE(g && h && dh && stats && mean && invstd && gamma, "bn_bwd64: null tensor");
  NSG_REQUIRE(count > 0, "bn_bwd64: count must be positive");
  if (pixels == 0) return NSG_OK;
  BnBwdArgs a{static_cast<const __nv_bfloat16*>(g), static_cast<const __nv_bfloat16*>(h),
              static_cast<__nv_bfloat16*>(dh), stats, count, mean, invstd, gamma, dgamma, dbeta, dbias_prev, pixels,
              PeerSync{}, gate, 1.0 / count};
  a.peer.world = 1;
  if (peer && peer->world > 1) {
    NSG_REQUIRE(gate, "bn_bwd64_sync: the exchange needs a zeroed gate counter");
    if (int rc = make_peer(peer, &a.peer)) return rc;
  }
  launch_pdl(bn_bwd_kernel, dim3(fe_grid(pixels)), dim3(256), 0, static_cast<cudaStream_t>(stream), a);
  NSG_CHECK_LAUNCH("bn_bwd_kernel");
  return NSG_OK;
}

int nsg_relu64(const void* x, void* out, long long n, void* stream) {
  NSG_REQUIRE(x && out && n % 8 == 0, "relu64: null tensor or element count not a multiple of 8");
  if (n == 0) return NSG_OK;
  long long blocks = (n / 8 + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  launch_pdl(relu_kernel, dim3((int)blocks), dim3(256), 0, static_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(x), static_cast<__nv_bfloat16*>(out), n / 8);
  NSG_CHECK_LAUNCH("relu_kernel");
  return NSG_OK;
}

int nsg_vq_bwd_stats64(const void* g_st, const void* z, const void* zq, float coef, void* dz, const void* h,
                       const float* mean, const float* invstd, double* stats, long long pixels, void* stream) {
  NSG_REQUIRE(z && dz, "vq_bwd_stats64: null tensor");
  NSG_REQUIRE(!stats || (h && mean && invstd), "vq_bwd_stats64: statistics need h, mean and invstd");
  if (pixels == 0) return NSG_OK;
  launch_pdl(vq_bwd_stats_kernel, dim3(fe_grid(pixels)), dim3(256), 0, static_cast<cudaStream_t>(stream), 
      static_cast<const __nv_bfloat16*>(g_st), static_cast<const __nv_bfloat16*>(z),
      static_cast<const __nv_bfloat16*>(zq), coef, static_cast<__nv_bfloat16*>(dz),
      static_cast<const __nv_bfloat16*>(h), mean, invstd, stats, pixels);
  NSG_CHECK_LAUNCH("vq_bwd_stats_kernel");
  return NSG_OK;
}

}  // extern "C"
