// Nearest-codeword search on the 5th-generation tensor cores (tcgen05 + TMEM), D = 64.
//
// scores S = z . e^T are one bf16 GEMM whose operands are split hi/lo so that the fp32 reference's
// indices survive: S ~= z_hi.e_hi + z_hi.e_lo + z_lo.e_hi, three tcgen05.mma passes into
// the same TMEM accumulator.  Per CTA: 128 rows of z (M = 128), the codebook streamed in chunks of 256
// codewords (N = 256) through two 64 KB shared-memory buffers filled by 1-D bulk async copies of a
// pre-swizzled image, two 256-column TMEM accumulators ping-ponged between the MMA thread and 8 epilogue
// warps.  The epilogue keeps the two best codewords per row from the tensor-core distances
// (|e|^2 - 2S), then re-scores those two EXACTLY in fp32 (sum (z-e)^2), which also yields the fused
// squared-error (train.py:131,133) and the code gather (vector_quantization.py:40-42).
//
// warp roles: 0-7 operand staging + epilogue (one TMEM row per thread, two column halves), 8 = TMEM
// allocator + bulk-copy producer + MMA issuer (one elected lane).
#include "common.cuh"
#include "tc_common.cuh"

namespace nsg {

using namespace tc;

constexpr int VT_M = 128, VT_N = 256, VT_D = 64;
constexpr uint32_t VT_IMG_HALF = VT_N * 128;          // one 256-row bf16 K-major image (hi or lo)
constexpr uint32_t VT_CHUNK_BYTES = 2 * VT_IMG_HALF;  // hi + lo
constexpr int VT_MAX_CHUNKS = 32;                     // K <= 8192

// codebook fp32 (K,64) -> per chunk: swizzled bf16 hi image, swizzled bf16 lo image; |e|^2 (inf padded)
__global__ void __launch_bounds__(256)
vq_prep_kernel(const float* __restrict__ cb, int K, uint8_t* __restrict__ image, float* __restrict__ esq_pad) {
  const int gid = blockIdx.x * 256 + threadIdx.x;
  const int c = gid & 7, r = (gid >> 3) & 255, ch = gid >> 11;
  const int code = ch * VT_N + r;
  float v[8];
  if (code < K) {
    const float4 a = __ldg(reinterpret_cast<const float4*>(cb + (long long)code * VT_D + c * 8));
    const float4 b = __ldg(reinterpret_cast<const float4*>(cb + (long long)code * VT_D + c * 8 + 4));
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  } else {
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = 0.f;
  }
  float hi[8], lo[8], sq = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    split_bf16(v[i], hi[i], lo[i]);
    sq = fmaf(v[i], v[i], sq);
  }
  uint8_t* dst = image + (size_t)ch * VT_CHUNK_BYTES + sw128_offset(r, c);
  *reinterpret_cast<uint4*>(dst) =
      make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]), pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
  *reinterpret_cast<uint4*>(dst + VT_IMG_HALF) =
      make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
  sq += __shfl_xor_sync(0xffffffffu, sq, 4);
  sq += __shfl_xor_sync(0xffffffffu, sq, 2);
  sq += __shfl_xor_sync(0xffffffffu, sq, 1);
  if (c == 0) esq_pad[code] = code < K ? sq : INFINITY;
}

struct Top2 {
  float d1, d2;
  int i1, i2;
  __device__ __forceinline__ void init() {
    d1 = d2 = INFINITY;
    i1 = i2 = -1;
  }
  // candidates arrive in increasing index order: strict '<' keeps the lowest index among equals
  __device__ __forceinline__ void push_ordered(float d, int k) {
    const bool lt1 = d < d1, lt2 = d < d2;
    d2 = lt1 ? d1 : (lt2 ? d : d2);
    i2 = lt1 ? i1 : (lt2 ? k : i2);
    d1 = lt1 ? d : d1;
    i1 = lt1 ? k : i1;
  }
  // arbitrary order: lexicographic (distance, index)
  __device__ __forceinline__ void push_any(float d, int k) {
    if (k < 0) return;
    const bool lt1 = d < d1 || (d == d1 && (i1 < 0 || k < i1));
    const bool lt2 = d < d2 || (d == d2 && (i2 < 0 || k < i2));
    if (lt1) {
      d2 = d1; i2 = i1; d1 = d; i1 = k;
    } else if (lt2) {
      d2 = d; i2 = k;
    }
  }
};

template <typename T>
__device__ __forceinline__ void load8(const T* p, float (&v)[8]);
template <>
__device__ __forceinline__ void load8<float>(const float* p, float (&v)[8]) {
  const float4 a = __ldg(reinterpret_cast<const float4*>(p));
  const float4 b = __ldg(reinterpret_cast<const float4*>(p + 4));
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float (&v)[8]) {
  const uint4 r = __ldg(reinterpret_cast<const uint4*>(p));
  const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    v[2 * i] = __uint_as_float(w[i] << 16);
    v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}

template <typename T>
__global__ void __launch_bounds__(288, 1)
vq_search_tc_kernel(const T* __restrict__ z, const float* __restrict__ cb, const uint8_t* __restrict__ image,
                    const float* __restrict__ esq_pad, long long N, int K, int nchunks,
                    long long* __restrict__ indices, T* __restrict__ codes, double* __restrict__ sqerr) {
  constexpr bool kSplitZ = sizeof(T) == 4;   // bf16 activations have no low part
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* A_hi = sm;
  uint8_t* A_lo = sm + 16384;
  uint8_t* Bbuf = sm + 32768;                                   // 2 x 64 KB
  float* esq_s = reinterpret_cast<float*>(sm + 32768 + 2 * VT_CHUNK_BYTES);
  uint8_t* tail = reinterpret_cast<uint8_t*>(esq_s + nchunks * VT_N);
  float* sCD = reinterpret_cast<float*>(tail);                  // [2][128][2]
  int* sCI = reinterpret_cast<int*>(tail + 2048);               // [2][128][2]
  int* sWin = reinterpret_cast<int*>(tail + 4096);              // [128]
  float* sErr = reinterpret_cast<float*>(tail + 4608);          // [128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail + 5120);    // b_full[2], b_empty[2], acc_full[2], acc_empty[2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 5184);
  float* part = reinterpret_cast<float*>(tail + 5200);          // [4]
  uint64_t* b_full = bars;
  uint64_t* b_empty = bars + 2;
  uint64_t* acc_full = bars + 4;
  uint64_t* acc_empty = bars + 6;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long row0 = (long long)blockIdx.x * VT_M;

  if (tid == 0) {
    mbar_init(&b_full[0], 1); mbar_init(&b_full[1], 1);
    mbar_init(&b_empty[0], 1); mbar_init(&b_empty[1], 1);
    mbar_init(&acc_full[0], 1); mbar_init(&acc_full[1], 1);
    mbar_init(&acc_empty[0], 8); mbar_init(&acc_empty[1], 8);
    fence_mbar_init();
  }
  if (warp == 8) tmem_alloc<512>(tmem_slot);
  if (tid < 256) {
    // stage the z tile: thread = (row, half); 4 x 16-byte chunks of bf16 hi (and lo)
    const int row = tid >> 1, half = tid & 1;
    const long long grow = row0 + row;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int c = half * 4 + j;
      float v[8];
      if (grow < N) {
        load8<T>(z + grow * VT_D + c * 8, v);
      } else {
#pragma unroll
        for (int i = 0; i < 8; ++i) v[i] = 0.f;
      }
      float hi[8], lo[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) split_bf16(v[i], hi[i], lo[i]);
      const uint32_t off = sw128_offset(row, c);
      *reinterpret_cast<uint4*>(A_hi + off) = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]),
                                                         pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
      if (kSplitZ)
        *reinterpret_cast<uint4*>(A_lo + off) = make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]),
                                                           pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
    }
    for (int i = tid; i < nchunks * VT_N; i += 256) esq_s[i] = __ldg(esq_pad + i);
    fence_proxy_async_smem();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 8) {
    if (lane == 0) {
      // ---------------- producer + MMA issuer
      const uint32_t idesc = idesc_bf16(VT_M, VT_N);
      const uint64_t a_hi = sw128_desc(smem_u32(A_hi)), a_lo = sw128_desc(smem_u32(A_lo));
      auto load_chunk = [&](int c) {
        const int b = c & 1;
        mbar_arrive_expect_tx(&b_full[b], VT_CHUNK_BYTES);
        bulk_g2s(Bbuf + (size_t)b * VT_CHUNK_BYTES, image + (size_t)c * VT_CHUNK_BYTES, VT_CHUNK_BYTES, &b_full[b]);
      };
      load_chunk(0);
      if (nchunks > 1) load_chunk(1);
      for (int c = 0; c < nchunks; ++c) {
        const int b = c & 1;
        const uint32_t ph = (c >> 1) & 1;
        mbar_wait(&b_full[b], ph);
        if (c >= 2) mbar_wait(&acc_empty[b], ph ^ 1);
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)b * VT_N;
        const uint64_t b_hi = sw128_desc(smem_u32(Bbuf + (size_t)b * VT_CHUNK_BYTES));
        const uint64_t b_lo = sw128_desc(smem_u32(Bbuf + (size_t)b * VT_CHUNK_BYTES + VT_IMG_HALF));
#pragma unroll
        for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_hi + 2 * ks, b_hi + 2 * ks, idesc, ks > 0);
#pragma unroll
        for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_hi + 2 * ks, b_lo + 2 * ks, idesc, 1);
        if (kSplitZ) {
#pragma unroll
          for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_lo + 2 * ks, b_hi + 2 * ks, idesc, 1);
        }
        umma_commit(&acc_full[b]);
        umma_commit(&b_empty[b]);
        if (c >= 1 && c + 1 < nchunks) {      // refill the buffer chunk c-1 used, now that its MMAs retired
          const int pb = (c - 1) & 1;
          mbar_wait(&b_empty[pb], ((c - 1) >> 1) & 1);
          load_chunk(c + 1);
        }
      }
    }
  } else {
    // ---------------- epilogue: thread = TMEM row q*32+lane, column half g of every chunk
    const int q = warp & 3, g = warp >> 2;
    const int row = q * 32 + lane;
    Top2 t2;
    t2.init();
    for (int c = 0; c < nchunks; ++c) {
      const int b = c & 1;
      mbar_wait(&acc_full[b], (c >> 1) & 1);
      tc_fence_after();
#pragma unroll 1
      for (int j = 0; j < 4; ++j) {
        const int col0 = g * 128 + j * 32;
        uint32_t r[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(b * VT_N + col0), r);
        tmem_ld_wait();
        const float4* es = reinterpret_cast<const float4*>(esq_s + c * VT_N + col0);
        const int kbase = c * VT_N + col0;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float4 e = es[i];
          t2.push_ordered(fmaf(-2.f, __uint_as_float(r[4 * i + 0]), e.x), kbase + 4 * i + 0);
          t2.push_ordered(fmaf(-2.f, __uint_as_float(r[4 * i + 1]), e.y), kbase + 4 * i + 1);
          t2.push_ordered(fmaf(-2.f, __uint_as_float(r[4 * i + 2]), e.z), kbase + 4 * i + 2);
          t2.push_ordered(fmaf(-2.f, __uint_as_float(r[4 * i + 3]), e.w), kbase + 4 * i + 3);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&acc_empty[b]);
    }
    // padded codewords carry +inf and never win, but -1 guards K < 2
    if (t2.i1 >= K) t2.i1 = -1;
    if (t2.i2 >= K) t2.i2 = -1;
    sCD[(g * VT_M + row) * 2 + 0] = t2.d1; sCI[(g * VT_M + row) * 2 + 0] = t2.i1;
    sCD[(g * VT_M + row) * 2 + 1] = t2.d2; sCI[(g * VT_M + row) * 2 + 1] = t2.i2;
    asm volatile("bar.sync 1, 256;" ::: "memory");

    // ---------------- exact fp32 re-score of the two best candidates: 8 lanes per row
    const int sub = tid & 7;
#pragma unroll 1
    for (int pass = 0; pass < 4; ++pass) {
      const int rr = pass * 32 + (tid >> 3);
      const long long grow = row0 + rr;
      Top2 m;
      m.init();
#pragma unroll
      for (int gg = 0; gg < 2; ++gg)
#pragma unroll
        for (int s = 0; s < 2; ++s) m.push_any(sCD[(gg * VT_M + rr) * 2 + s], sCI[(gg * VT_M + rr) * 2 + s]);
      float p1 = 0.f, p2 = 0.f;
      if (grow < N && m.i1 >= 0) {
        float zv[8], ev[8];
        load8<T>(z + grow * VT_D + sub * 8, zv);
        load8<float>(cb + (long long)m.i1 * VT_D + sub * 8, ev);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float df = zv[i] - ev[i];
          p1 = fmaf(df, df, p1);
        }
        if (m.i2 >= 0) {
          load8<float>(cb + (long long)m.i2 * VT_D + sub * 8, ev);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float df = zv[i] - ev[i];
            p2 = fmaf(df, df, p2);
          }
        }
      }
#pragma unroll
      for (int o = 4; o > 0; o >>= 1) {
        p1 += __shfl_xor_sync(0xffffffffu, p1, o);
        p2 += __shfl_xor_sync(0xffffffffu, p2, o);
      }
      if (sub == 0) {
        int win = m.i1 < 0 ? 0 : m.i1;
        float dw = p1;
        if (m.i2 >= 0 && (p2 < p1 || (p2 == p1 && m.i2 < m.i1))) {
          win = m.i2;
          dw = p2;
        }
        sWin[rr] = win;
        sErr[rr] = grow < N ? dw : 0.f;
      }
    }
    asm volatile("bar.sync 1, 256;" ::: "memory");
    if (tid < VT_M) {
      const long long grow = row0 + tid;
      if (grow < N) indices[grow] = (long long)sWin[tid];
      if (sqerr) {
        const float e = warp_sum(sErr[tid]);
        if (lane == 0) part[warp] = e;
      }
    }
    if (codes) {
      for (int i = tid; i < VT_M * (VT_D / 4); i += 256) {
        const int rr = i >> 4, c4 = i & 15;
        const long long grow = row0 + rr;
        if (grow < N) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(cb + (long long)sWin[rr] * VT_D + c4 * 4));
          Act<T>::st4(codes + grow * VT_D + c4 * 4, v);
        }
      }
    }
    if (sqerr) {
      asm volatile("bar.sync 1, 256;" ::: "memory");
      if (tid == 0) atomicAdd(sqerr, (double)part[0] + (double)part[1] + (double)part[2] + (double)part[3]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    __syncwarp();
    tmem_dealloc<512>(tmem_base);
  }
}

static size_t vq_tc_smem_bytes(int nchunks) {
  return 1024 + 32768 + 2 * (size_t)VT_CHUNK_BYTES + (size_t)nchunks * VT_N * 4 + 5376;
}

long long vq_tc_workspace_bytes(int k, int d) {
  if (d != VT_D || k < 1 || k > VT_MAX_CHUNKS * VT_N) return -1;
  const long long nchunks = (k + VT_N - 1) / VT_N;
  return nchunks * (long long)VT_CHUNK_BYTES + nchunks * VT_N * 4;
}

int vq_forward_tc(const void* z, const float* codebook, long long n, int k, int d, int dtype, long long* indices,
                  void* codes, double* sqerr, void* ws, long long ws_bytes, cudaStream_t stream) {
  const long long need = vq_tc_workspace_bytes(k, d);
  if (need < 0) {
    set_error("vq_forward(mode=1): tensor-core search supports D == 64 and K <= %d (got K=%d, D=%d)",
              VT_MAX_CHUNKS * VT_N, k, d);
    return NSG_E_UNSUPPORTED;
  }
  NSG_REQUIRE(ws && ws_bytes >= need, "vq_forward(mode=1): workspace of %lld bytes required (got %lld)", need, ws_bytes);
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, "vq_forward(mode=1): workspace must be 16-byte aligned");
  const int nchunks = (k + VT_N - 1) / VT_N;
  uint8_t* image = static_cast<uint8_t*>(ws);
  float* esq_pad = reinterpret_cast<float*>(image + (size_t)nchunks * VT_CHUNK_BYTES);
  vq_prep_kernel<<<nchunks * 8, 256, 0, stream>>>(codebook, k, image, esq_pad);
  NSG_CHECK_LAUNCH("vq_prep_kernel");
  const size_t smem = vq_tc_smem_bytes(nchunks);
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    static bool configured = false;
    if (!configured) {
      cudaError_t e = cudaFuncSetAttribute(vq_search_tc_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)vq_tc_smem_bytes(VT_MAX_CHUNKS));
      if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(vq_search_tc_kernel)");
      configured = true;
    }
    const unsigned grid = (unsigned)((n + VT_M - 1) / VT_M);
    vq_search_tc_kernel<T><<<grid, 288, smem, stream>>>((const T*)z, codebook, image, esq_pad, n, k, nchunks,
                                                        indices, (T*)codes, sqerr);
    NSG_CHECK_LAUNCH("vq_search_tc_kernel");
    return NSG_OK;
  });
}

}  // namespace nsg
