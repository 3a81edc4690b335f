// Vector quantisation on CUDA cores: exact-fp32 nearest-codeword search with fused gather and
// squared-error reduction, scatter-add (codebook gradient / EMA statistics), bottleneck gradient.
// Restates /root/reference/src/vector_quantization.py:6-63 (see include/nsg_b200.h).
#include "common.cuh"

namespace nsg {

constexpr int VR = 64;   // rows per CTA
constexpr int VC = 64;   // codewords per tile
constexpr int VD = 16;   // depth chunk
constexpr int VLD = VR + 4;

// argmin_k (|e_k|^2 - 2 z.e_k): the row-constant |z|^2 of the reference (vector_quantization.py:13)
// cannot change the argmin; dropping it only moves results inside the documented tie window.
template <typename T>
__global__ void __launch_bounds__(256)
vq_search_simt_kernel(const T* __restrict__ z, const float* __restrict__ cb, long long N, int K, int D,
                      long long* __restrict__ indices, T* __restrict__ codes, double* __restrict__ sqerr) {
  __shared__ __align__(16) float Zs[VD][VLD];
  __shared__ __align__(16) float Es[VD][VLD];
  __shared__ float esq[VC];
  __shared__ float sBest[VR][17];
  __shared__ int sIdx[VR][17];
  __shared__ int rowIdx[VR];
  __shared__ float part[8];

  const int tid = threadIdx.x;
  const long long r0 = (long long)blockIdx.x * VR;
  const int tx = tid & 15, ty = tid >> 4;   // tx: code quad, ty: row quad
  const int l_dk = tid & 15, l_r = tid >> 4;  // loader: depth lane, row base

  float best[4];
  int bidx[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    best[i] = INFINITY;
    bidx[i] = 0;
  }

  for (int k0 = 0; k0 < K; k0 += VC) {
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int e = 0; e < 4; ++e) acc[i][e] = 0.f;
    float my_esq = 0.f;
    for (int d0 = 0; d0 < D; d0 += VD) {
      const int d = d0 + l_dk;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int r = l_r + 16 * j;
        const long long row = r0 + r;
        Zs[l_dk][r] = (d < D && row < N) ? Act<T>::ld(z + row * D + d) : 0.f;
        const int code = k0 + r;
        Es[l_dk][r] = (d < D && code < K) ? __ldg(cb + (long long)code * D + d) : 0.f;
      }
      __syncthreads();
      if (tid < VC) {
#pragma unroll
        for (int k = 0; k < VD; ++k) my_esq = fmaf(Es[k][tid], Es[k][tid], my_esq);
      }
#pragma unroll
      for (int k = 0; k < VD; ++k) {
        const float4 a = *reinterpret_cast<const float4*>(&Zs[k][ty * 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Es[k][tx * 4]);
        const float av[4] = {a.x, a.y, a.z, a.w};
        const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int e = 0; e < 4; ++e) acc[i][e] = fmaf(av[i], bv[e], acc[i][e]);
      }
      __syncthreads();
    }
    if (tid < VC) esq[tid] = my_esq;
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const int code = k0 + tx * 4 + e;
      if (code < K) {
        const float es = esq[tx * 4 + e];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float dist = fmaf(-2.f, acc[i][e], es);
          if (dist < best[i]) {   // strict: the lowest index wins ties (torch.min, vector_quantization.py:19)
            best[i] = dist;
            bidx[i] = code;
          }
        }
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    sBest[ty * 4 + i][tx] = best[i];
    sIdx[ty * 4 + i][tx] = bidx[i];
  }
  __syncthreads();
  if (tid < VR) {
    float b = sBest[tid][0];
    int bi = sIdx[tid][0];
    for (int j = 1; j < 16; ++j) {
      const float v = sBest[tid][j];
      const int vi = sIdx[tid][j];
      if (v < b || (v == b && vi < bi)) {
        b = v;
        bi = vi;
      }
    }
    rowIdx[tid] = bi;
    if (r0 + tid < N) indices[r0 + tid] = (long long)bi;
  }
  __syncthreads();
  // gather + squared error: warp w handles rows w*8..w*8+7, lanes stride over D (coalesced)
  const int warp = tid >> 5, lane = tid & 31;
  float err = 0.f;
  if (codes || sqerr) {
    for (int i = 0; i < 8; ++i) {
      const int r = warp * 8 + i;
      const long long row = r0 + r;
      if (row >= N) break;
      const float* e = cb + (long long)rowIdx[r] * D;
      for (int d = lane; d < D; d += 32) {
        const float ev = __ldg(e + d);
        if (codes) Act<T>::st(codes + row * D + d, ev);
        const float df = Act<T>::ld(z + row * D + d) - ev;
        err = fmaf(df, df, err);
      }
    }
  }
  if (sqerr) {
    err = warp_sum(err);
    if (lane == 0) part[warp] = err;
    __syncthreads();
    if (tid == 0) {
      double s = 0.0;
      for (int i = 0; i < 8; ++i) s += (double)part[i];
      atomicAdd(sqerr, s);
    }
  }
}

// table[idx[n], :] += alpha * (rows[n,:] - sub[n,:]); counts[idx[n]] += 1
template <typename T, int VW>
__global__ void scatter_add_kernel(const T* __restrict__ rows, const T* __restrict__ sub,
                                   const long long* __restrict__ idx, long long N, int K, int D, float alpha,
                                   float* __restrict__ table, float* __restrict__ counts) {
  pdl_wait();
  pdl_trigger();
  const int dq = D / VW;
  const long long total = N * dq;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / dq;
    const int d = (int)(i % dq) * VW;
    const long long k = idx[n];
    if (k < 0 || k >= K) continue;
    if (VW == 4) {
      float4 v = Act<T>::ld4(rows + n * D + d);
      if (sub) {
        float4 s = Act<T>::ld4(sub + n * D + d);
        v.x -= s.x; v.y -= s.y; v.z -= s.z; v.w -= s.w;
      }
      v.x *= alpha; v.y *= alpha; v.z *= alpha; v.w *= alpha;
      atomicAdd(reinterpret_cast<float4*>(table + k * D + d), v);   // red.global.add.v4.f32
    } else {
      float v = Act<T>::ld(rows + n * D + d);
      if (sub) v -= Act<T>::ld(sub + n * D + d);
      atomicAdd(table + k * D + d, alpha * v);
    }
    if (counts && d == 0) atomicAdd(counts + k, 1.0f);
  }
}

template <typename T, int VW>
__global__ void vq_bwd_inputs_kernel(const T* __restrict__ g, const T* __restrict__ z, const T* __restrict__ zq,
                                     float coef, T* __restrict__ dz, long long total_vec) {
  pdl_wait();
  pdl_trigger();
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total_vec;
       i += (long long)gridDim.x * blockDim.x) {
    const long long o = i * VW;
    if (VW == 4) {
      float4 a = g ? Act<T>::ld4(g + o) : make_float4(0.f, 0.f, 0.f, 0.f);
      const float4 zz = Act<T>::ld4(z + o);
      const float4 q = zq ? Act<T>::ld4(zq + o) : make_float4(0.f, 0.f, 0.f, 0.f);
      a.x = fmaf(coef, zz.x - q.x, a.x);
      a.y = fmaf(coef, zz.y - q.y, a.y);
      a.z = fmaf(coef, zz.z - q.z, a.z);
      a.w = fmaf(coef, zz.w - q.w, a.w);
      Act<T>::st4(dz + o, a);
    } else {
      float a = g ? Act<T>::ld(g + o) : 0.f;
      Act<T>::st(dz + o, fmaf(coef, Act<T>::ld(z + o) - (zq ? Act<T>::ld(zq + o) : 0.f), a));
    }
  }
}

// out[n, :] = table[idx[n], :]   (nn.Embedding lookup / index_select, models.py:137,194)
template <typename T>
__global__ void gather_rows_kernel(const float* __restrict__ table, const long long* __restrict__ idx, long long N,
                                   int K, int D, T* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  const long long total = N * D;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total;
       i += (long long)gridDim.x * blockDim.x) {
    const long long n = i / D;
    const int d = (int)(i % D);
    const long long k = idx[n];
    Act<T>::st(out + i, (k >= 0 && k < K) ? __ldg(table + k * D + d) : 0.f);
  }
}

// D % 4 == 0: one thread moves 4 channels (a 16-byte table read, an 8- or 16-byte store); the lanes of a warp cover
// consecutive channel groups of consecutive rows, so stores are full lines and the index is read once per row group
template <typename T>
__global__ void __launch_bounds__(256)
gather_rows_vec_kernel(const float* __restrict__ table, const long long* __restrict__ idx, long long N, int K, int dq,
                       T* __restrict__ out) {
  pdl_wait();
  pdl_trigger();
  const long long total = N * dq;
  const long long stride = (long long)gridDim.x * 256;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += stride) {
    const long long n = i / dq;
    const int c4 = (int)(i - n * dq);
    const long long k = __ldg(idx + n);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k >= 0 && k < K) v = __ldg(reinterpret_cast<const float4*>(table + (k * dq + c4) * 4));
    Act<T>::st4(out + i * 4, v);
  }
}

int vq_forward_simt(const void* z, const float* codebook, long long n, int k, int d, int dtype,
                    long long* indices, void* codes, double* sqerr, cudaStream_t stream) {
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    const unsigned grid = (unsigned)((n + VR - 1) / VR);
    vq_search_simt_kernel<T><<<grid, 256, 0, stream>>>((const T*)z, codebook, n, k, d, indices, (T*)codes, sqerr);
    NSG_CHECK_LAUNCH("vq_search_simt_kernel");
    return NSG_OK;
  });
}

int vq_forward_tc(const void* z, const float* codebook, long long n, int k, int d, int dtype, long long* indices,
                  void* codes, double* sqerr, void* ws, long long ws_bytes, cudaStream_t stream);
long long vq_tc_workspace_bytes(int k, int d);

}  // namespace nsg

using namespace nsg;

namespace nsg {
int scatter_add_agg(const void* rows, const void* sub, const long long* indices, long long n, int k, int d, float alpha,
                    float* table, float* counts, int dtype, cudaStream_t stream);
}
// measured on B200 (profiles/r1_final_vq_kernels.txt, N = 40960 / 2^20, K = 512, D = 64, bf16): the per-element
// vector-atomic kernel is the faster one when the assignment is uniform (12 vs 21 us; 246 vs 493 us) but collapses when
// a few codewords take most rows, which is exactly what an untrained encoder produces (146 us; 3461 us), while the
// block-aggregated kernel stays at 19 us / 444 us.  The bounded one is the default; mode 1 selects the other.
static int g_scatter_mode = 0;
namespace nsg {
extern int g_vq_resident;
}

extern "C" {

int nsg_vq_resident(int enable) {
  const int old = g_vq_resident;
  if (enable == 0 || enable == 1) g_vq_resident = enable;
  return old;
}

int nsg_scatter_mode(int mode) {
  const int old = g_scatter_mode;
  if (mode == 0 || mode == 1) g_scatter_mode = mode;
  return old;
}

long long nsg_vq_workspace_bytes(int k, int d, int mode) {
  if (mode == 0) return 0;
  return vq_tc_workspace_bytes(k, d);
}

int nsg_vq_forward(const void* z, const float* codebook, long long n, int k, int d, int dtype, int mode,
                   long long* indices, void* codes, double* sqerr, void* ws, long long ws_bytes, void* stream) {
  NSG_REQUIRE(z && codebook && indices, "vq_forward: null pointer");
  NSG_REQUIRE(n >= 0 && k >= 1 && d >= 1, "vq_forward: bad sizes n=%lld k=%d d=%d", n, k, d);
  if (n == 0) return NSG_OK;
  if (mode == 0) return vq_forward_simt(z, codebook, n, k, d, dtype, indices, codes, sqerr, (cudaStream_t)stream);
  if (mode == 1)
    return vq_forward_tc(z, codebook, n, k, d, dtype, indices, codes, sqerr, ws, ws_bytes, (cudaStream_t)stream);
  set_error("vq_forward: unknown mode %d", mode);
  return NSG_E_INVALID;
}

int nsg_scatter_add(const void* rows, const void* sub, const long long* indices, long long n, int k, int d,
                    float alpha, float* table, float* counts, int dtype, void* stream) {
  NSG_REQUIRE(rows && indices && table && n >= 0 && k >= 1 && d >= 1, "scatter_add: bad arguments");
  if (n == 0) return NSG_OK;
  if (g_scatter_mode == 0) {      // block-aggregated kernel (scatter_ema.cu) when the shape is served
    const int rc = scatter_add_agg(rows, sub, indices, n, k, d, alpha, table, counts, dtype, (cudaStream_t)stream);
    if (rc != NSG_E_UNSUPPORTED) return rc;
  }
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    const int threads = 256;
    if (d % 4 == 0) {
      long long work = n * (d / 4);
      long long blocks = (work + threads - 1) / threads;
      if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
      launch_pdl(scatter_add_kernel<T, 4>, dim3((int)blocks), dim3(threads), 0, (cudaStream_t)stream, 
          (const T*)rows, (const T*)sub, indices, n, k, d, alpha, table, counts);
    } else {
      long long work = n * d;
      long long blocks = (work + threads - 1) / threads;
      if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
      launch_pdl(scatter_add_kernel<T, 1>, dim3((int)blocks), dim3(threads), 0, (cudaStream_t)stream, 
          (const T*)rows, (const T*)sub, indices, n, k, d, alpha, table, counts);
    }
    NSG_CHECK_LAUNCH("scatter_add_kernel");
    return NSG_OK;
  });
}

int nsg_gather_rows(const float* table, const long long* indices, long long n, int k, int d, void* out, int dtype,
                    void* stream) {
  NSG_REQUIRE(table && indices && out && n >= 0 && k >= 1 && d >= 1, "gather_rows: bad arguments");
  if (n == 0) return NSG_OK;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    if (d % 4 == 0 && (reinterpret_cast<uintptr_t>(table) & 15) == 0 && (reinterpret_cast<uintptr_t>(out) & 15) == 0) {
      long long blocks = (n * (d / 4) + 255) / 256;
      if (blocks > (long long)num_sms() * 32) blocks = (long long)num_sms() * 32;
      launch_pdl(gather_rows_vec_kernel<T>, dim3((int)blocks), dim3(256), 0, (cudaStream_t)stream, table, indices, n, k, d / 4, (T*)out);
      NSG_CHECK_LAUNCH("gather_rows_vec_kernel");
      return NSG_OK;
    }
    long long blocks = (n * d + 255) / 256;
    if (blocks > (long long)num_sms() * 16) blocks = (long long)num_sms() * 16;
    launch_pdl(gather_rows_kernel<T>, dim3((int)blocks), dim3(256), 0, (cudaStream_t)stream, table, indices, n, k, d, (T*)out);
    NSG_CHECK_LAUNCH("gather_rows_kernel");
    return NSG_OK;
  });
}

int nsg_vq_backward_inputs(const void* g_st, const void* z, const void* zq, float coef, void* dz, long long n,
                           int dtype, void* stream) {
  NSG_REQUIRE(z && dz && n >= 0, "vq_backward_inputs: bad arguments");
  if (n == 0) return NSG_OK;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    const int threads = 256;
    if (n % 4 == 0) {
      long long blocks = (n / 4 + threads - 1) / threads;
      if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
      launch_pdl(vq_bwd_inputs_kernel<T, 4>, dim3((int)blocks), dim3(threads), 0, (cudaStream_t)stream, 
          (const T*)g_st, (const T*)z, (const T*)zq, coef, (T*)dz, n / 4);
    } else {
      long long blocks = (n + threads - 1) / threads;
      if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
      launch_pdl(vq_bwd_inputs_kernel<T, 1>, dim3((int)blocks), dim3(threads), 0, (cudaStream_t)stream, 
          (const T*)g_st, (const T*)z, (const T*)zq, coef, (T*)dz, n);
    }
    NSG_CHECK_LAUNCH("vq_bwd_inputs_kernel");
    return NSG_OK;
  });
}

}  // extern "C"
