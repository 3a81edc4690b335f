// Codebook scatter-add with block-level aggregation, and the EMA codebook update built on it.
//
//   table[idx[n], :] += alpha * (rows[n, :] - sub[n, :]);  counts[idx[n]] += 1
//
// is the reference's index_add_ (/root/reference/src/vector_quantization.py:58-61; autograd of the index_select at
// src/models.py:137-138 = the codebook gradient) and, with counts, the statistics pass of the EMA codebook
// extension.  N rows fold into K << N table rows, so the cost is the number of atomics that reach L2,
// not the bytes: one vector atomic per (row, 4 channels) -- the first-cut kernel in vq_simt.cu -- runs at ~3 % of HBM.
//
// Here the grid is P row partitions x G codeword groups.  CTA (p, g) owns the codewords k with k % G == g in a
// shared-memory table, scans its partition's indices in coalesced chunks (staged once, compacted to a match list),
// and lane groups fold the matching rows: each group walks a CONTIGUOUS slice of the list, aggregates runs of
// equal codeword in registers (so a skewed assignment costs fewer, not more, atomics) and merges run totals into the
// shared table.  At the end every touched table row leaves through one 16-byte vector atomic per 4 channels:
// P * K * D/4 global atomics instead of N * D/4.  P and G are chosen so that P*G fills the machine twice, a
// partition holds >= 4096 rows, and the redundant index scan (each index is read by G CTAs) stays below the row bytes.
#include "common.cuh"

namespace nsg {

constexpr int SA_THREADS = 256;
constexpr int SA_CHUNK = 2048;            // indices staged per pass
constexpr int SA_MAXQ = 2;                // float4 groups per lane (D <= 256)
constexpr int SA_TABLE_BYTES = 96 * 1024;

struct ScatterPlan {
  int P, G, slots, lpr;                   // partitions, groups, table rows per CTA, lanes per row
  long long rows_per_part;
  size_t smem;
};

template <typename T>
__global__ void __launch_bounds__(SA_THREADS)
scatter_agg_kernel(const T* __restrict__ rows, const T* __restrict__ sub, const long long* __restrict__ idx,
                   long long N, int K, int D, float alpha, float* __restrict__ table, float* __restrict__ counts,
                   int G, int slots, int lpr, long long rows_per_part) {
  pdl_wait();
  pdl_trigger();
  extern __shared__ __align__(16) uint8_t sa_smem[];
  float* tab = reinterpret_cast<float*>(sa_smem);                         // [slots][D]
  int* cnt = reinterpret_cast<int*>(tab + (size_t)slots * D);             // [slots]
  int* list_row = cnt + slots;                                            // [SA_CHUNK] row offset inside the chunk
  int* list_slot = list_row + SA_CHUNK;                                   // [SA_CHUNK]
  __shared__ int list_n;

  const int tid = threadIdx.x;
  const int g = blockIdx.x % G;
  const long long p = blockIdx.x / G;
  const long long r_begin = p * rows_per_part;
  const long long r_end = r_begin + rows_per_part < N ? r_begin + rows_per_part : N;
  const int dq = D >> 2;

  for (int i = tid; i < slots * D; i += SA_THREADS) tab[i] = 0.f;
  for (int i = tid; i < slots; i += SA_THREADS) cnt[i] = 0;

  const int ngroups = SA_THREADS / lpr;
  const int grp = tid / lpr, lane = tid % lpr;

  for (long long c0 = r_begin; c0 < r_end; c0 += SA_CHUNK) {
    if (tid == 0) list_n = 0;
    __syncthreads();
    const int cn = (int)(r_end - c0 < SA_CHUNK ? r_end - c0 : SA_CHUNK);
    // stage + compact: coalesced index loads, one shared counter bump per match (matches are 1/G of the indices)
    for (int j = tid; j < cn; j += SA_THREADS) {
      const long long k = __ldg(idx + c0 + j);
      if (k >= 0 && k < K && (int)(k % G) == g) {
        const int pos = atomicAdd(&list_n, 1);
        list_row[pos] = j;
        list_slot[pos] = (int)(k / G);
      }
    }
    __syncthreads();
    const int n = list_n;
    const int per = (n + ngroups - 1) / ngroups;
    const int e0 = grp * per, e1 = e0 + per < n ? e0 + per : n;
    int cur = -1, run = 0;
    float4 acc[SA_MAXQ];
#pragma unroll
    for (int q = 0; q < SA_MAXQ; ++q) acc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
    auto flush = [&]() {
      if (cur < 0) return;
#pragma unroll
      for (int q = 0; q < SA_MAXQ; ++q) {
        const int c4 = lane + q * lpr;
        if (c4 < dq) {
          float* dst = tab + (size_t)cur * D + c4 * 4;
          atomicAdd(dst + 0, acc[q].x);
          atomicAdd(dst + 1, acc[q].y);
          atomicAdd(dst + 2, acc[q].z);
          atomicAdd(dst + 3, acc[q].w);
          acc[q] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      if (lane == 0) atomicAdd(&cnt[cur], run);
      run = 0;
    };
    // four entries' loads in flight per lane group
    for (int e = e0; e < e1; e += 4) {
      float4 v[4][SA_MAXQ];
      int s[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const bool on = e + u < e1;
        s[u] = on ? list_slot[e + u] : -1;
        const long long rbase = on ? (c0 + list_row[e + u]) * (long long)D : 0;
#pragma unroll
        for (int q = 0; q < SA_MAXQ; ++q) {
          const int c4 = lane + q * lpr;
          v[u][q] = make_float4(0.f, 0.f, 0.f, 0.f);
          if (on && c4 < dq) {
            v[u][q] = Act<T>::ld4(rows + rbase + c4 * 4);
            if (sub) {
              const float4 t = Act<T>::ld4(sub + rbase + c4 * 4);
              v[u][q].x -= t.x; v[u][q].y -= t.y; v[u][q].z -= t.z; v[u][q].w -= t.w;
            }
          }
        }
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (s[u] < 0) continue;
        if (s[u] != cur) {
          flush();
          cur = s[u];
        }
        ++run;
#pragma unroll
        for (int q = 0; q < SA_MAXQ; ++q) {
          acc[q].x = fmaf(alpha, v[u][q].x, acc[q].x);
          acc[q].y = fmaf(alpha, v[u][q].y, acc[q].y);
          acc[q].z = fmaf(alpha, v[u][q].z, acc[q].z);
          acc[q].w = fmaf(alpha, v[u][q].w, acc[q].w);
        }
      }
    }
    flush();
    __syncthreads();
  }
  // touched table rows -> global, one vector atomic per 4 channels
  for (int i = tid; i < slots * dq; i += SA_THREADS) {
    const int s = i / dq, c4 = i % dq;
    const int k = s * G + g;
    if (k < K && cnt[s] > 0) {
      const float4 v = *reinterpret_cast<const float4*>(tab + (size_t)s * D + c4 * 4);
      atomicAdd(reinterpret_cast<float4*>(table + (size_t)k * D + c4 * 4), v);
      if (counts && c4 == 0) atomicAdd(counts + k, (float)cnt[s]);
    }
  }
}

static bool plan_scatter(long long n, int k, int d, int row_bytes, ScatterPlan* pl) {
  if (d % 4 != 0 || d > 4 * 32 * SA_MAXQ) return false;
  int lpr = 1;
  while (lpr < 32 && lpr * 2 <= d / 4) lpr *= 2;
  // whole table in one CTA's shared memory (K = 512, D = 64: 128 KB): no codeword groups at all -- every CTA owns a
  // slice of the rows, each index is read once, the work is balanced whatever the assignment looks like, and only
  // the touched table rows leave through global atomics (a handful under the heavy skew of an untrained encoder)
  const size_t whole = (size_t)k * d * 4 + (size_t)k * 4 + 2 * SA_CHUNK * 4;
  if (whole <= 200 * 1024) {
    long long P1 = (n + 63) / 64;
    if (P1 > num_sms()) P1 = num_sms();
    pl->P = (int)P1;
    pl->G = 1;
    pl->slots = k;
    pl->lpr = lpr;
    pl->rows_per_part = (n + P1 - 1) / P1;
    pl->smem = whole;
    return true;
  }
  const int gmin = (int)(((long long)k * d * 4 + SA_TABLE_BYTES - 1) / SA_TABLE_BYTES);
  const int target = 2 * num_sms();
  long long P = (n + 4095) / 4096;
  if (P > target) P = target;
  if (P < 1) P = 1;
  int G = (int)((target + P - 1) / P);
  const int gmax = row_bytes / 8 > 1 ? row_bytes / 8 : 1;     // redundant index scan <= row bytes
  if (G > gmax) G = gmax;
  if (G < gmin) G = gmin;
  if (G > k) G = k;
  if (G < gmin) return false;                                  // table does not fit even at one codeword per CTA
  pl->P = (int)P;
  pl->G = G;
  pl->slots = (k + G - 1) / G;
  pl->lpr = lpr;
  pl->rows_per_part = (n + P - 1) / P;
  pl->smem = (size_t)pl->slots * d * 4 + (size_t)pl->slots * 4 + 2 * SA_CHUNK * 4;
  return pl->smem <= 200 * 1024;
}

// returns NSG_E_UNSUPPORTED when the shape is left to the per-element kernel
int scatter_add_agg(const void* rows, const void* sub, const long long* indices, long long n, int k, int d, float alpha,
                    float* table, float* counts, int dtype, cudaStream_t stream) {
  ScatterPlan pl;
  const int esz = dtype == NSG_BF16 ? 2 : 4;
  if (n < 4096 || !plan_scatter(n, k, d, d * esz * (sub ? 2 : 1), &pl)) return NSG_E_UNSUPPORTED;
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    static size_t configured = 0;
    if (pl.smem > 48 * 1024 && pl.smem > configured) {
      cudaError_t e = cudaFuncSetAttribute(scatter_agg_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e == cudaSuccess)
        e = cudaFuncSetAttribute(scatter_agg_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(scatter_agg_kernel)");
      configured = 200 * 1024;
    }
    launch_pdl(scatter_agg_kernel<T>, dim3(pl.P * pl.G), dim3(SA_THREADS), pl.smem, stream, 
        (const T*)rows, (const T*)sub, indices, n, k, d, alpha, table, counts, pl.G, pl.slots, pl.lpr, pl.rows_per_part);
    NSG_CHECK_LAUNCH("scatter_agg_kernel");
    return NSG_OK;
  });
}

// ---------------------------------------------------------------------------------------------- EMA codebook update
// N <- g N + (1-g) counts;  m <- g m + (1-g) sums;  e = m / ((N + eps) / (sum N + K eps) * sum N)
// (van den Oord et al. 2017, appendix A.1; extension, not in the reference).  One CTA reduces sum N (K <= 65536),
// then the grid streams the K x D table with 16-byte accesses.
__global__ void __launch_bounds__(256)
ema_count_kernel(const float* __restrict__ counts, float* __restrict__ ema_count, int K, float decay,
                 double* __restrict__ total) {
  pdl_wait();
  pdl_trigger();
  __shared__ double part[8];
  double s = 0.0;
  for (int k = threadIdx.x; k < K; k += 256) {
    const float nk = decay * ema_count[k] + (1.f - decay) * counts[k];
    ema_count[k] = nk;
    s += (double)nk;
  }
  for (int o = 16; o >= 1; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    for (int i = 0; i < 8; ++i) t += part[i];
    *total = t;
  }
}

__global__ void __launch_bounds__(256)
ema_table_kernel(const float* __restrict__ sums, const float* __restrict__ ema_count, float* __restrict__ ema_sum,
                 float* __restrict__ codebook, int K, int D, float decay, float eps, const double* __restrict__ total) {
  pdl_wait();
  pdl_trigger();
  const double tot = *total;
  const long long nvec = (long long)K * D / 4;
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (long long)gridDim.x * 256) {
    const int k = (int)(i * 4 / D);
    const float4 s = __ldg(reinterpret_cast<const float4*>(sums) + i);
    float4 m = reinterpret_cast<const float4*>(ema_sum)[i];
    m.x = decay * m.x + (1.f - decay) * s.x;
    m.y = decay * m.y + (1.f - decay) * s.y;
    m.z = decay * m.z + (1.f - decay) * s.z;
    m.w = decay * m.w + (1.f - decay) * s.w;
    reinterpret_cast<float4*>(ema_sum)[i] = m;
    const double ns = ((double)ema_count[k] + (double)eps) / (tot + (double)K * (double)eps) * tot;
    const float inv = (float)(1.0 / ns);
    reinterpret_cast<float4*>(codebook)[i] = make_float4(m.x * inv, m.y * inv, m.z * inv, m.w * inv);
  }
}

}  // namespace nsg

using namespace nsg;

extern "C" int nsg_ema_update(const float* counts, const float* sums, float* ema_count, float* ema_sum, float* codebook,
                              int k, int d, float decay, float eps, double* scratch, void* stream) {
  NSG_REQUIRE(counts && sums && ema_count && ema_sum && codebook && scratch, "ema_update: null pointer");
  NSG_REQUIRE(k >= 1 && d >= 4 && d % 4 == 0, "ema_update: needs d %% 4 == 0 (got k=%d d=%d)", k, d);
  NSG_REQUIRE(decay >= 0.f && decay <= 1.f, "ema_update: decay %f outside [0, 1]", (double)decay);
  cudaStream_t st = (cudaStream_t)stream;
  launch_pdl(ema_count_kernel, dim3(1), dim3(256), 0, st, counts, ema_count, k, decay, scratch);
  NSG_CHECK_LAUNCH("ema_count_kernel");
  long long blocks = ((long long)k * d / 4 + 255) / 256;
  if (blocks > (long long)num_sms() * 8) blocks = (long long)num_sms() * 8;
  launch_pdl(ema_table_kernel, dim3((int)blocks), dim3(256), 0, st, sums, ema_count, ema_sum, codebook, k, d, decay, eps, scratch);
  NSG_CHECK_LAUNCH("ema_table_kernel");
  return NSG_OK;
}
