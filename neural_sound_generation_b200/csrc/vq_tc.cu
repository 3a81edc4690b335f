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
#define NSG_TU_ID 3
#include "tc_common.cuh"
#include "tma.cuh"
#include <cstring>

namespace nsg {

using namespace tc;

constexpr int VT_M = 128, VT_N = 256, VT_D = 64;
constexpr uint32_t VT_IMG_HALF = VT_N * 128;          // one 256-row bf16 K-major image (hi or lo)
constexpr uint32_t VT_CHUNK_BYTES = 2 * VT_IMG_HALF;  // hi + lo
constexpr int VT_MAX_CHUNKS = 32;                     // K <= 8192

// codebook fp32 (K, D = 64 * nd) -> per (chunk of 256 codewords, block of 64 dims): swizzled bf16 hi image, swizzled bf16
// lo image (unit u = chunk * nd + dblock at image + u * VT_CHUNK_BYTES); |e|^2 over all D (inf padded)
__global__ void __launch_bounds__(256)
vq_prep_kernel(const float* __restrict__ cb, int K, int nd, uint8_t* __restrict__ image, float* __restrict__ esq_pad) {
  pdl_wait();
  pdl_trigger();
  const int gid = blockIdx.x * 256 + threadIdx.x;
  const int c = gid & 7, r = (gid >> 3) & 255, ch = gid >> 11;
  const int code = ch * VT_N + r;
  const int D = nd * VT_D;
  float sq = 0.f;
  for (int db = 0; db < nd; ++db) {
    float v[8];
    if (code < K) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(cb + (long long)code * D + db * VT_D + c * 8));
      const float4 b = __ldg(reinterpret_cast<const float4*>(cb + (long long)code * D + db * VT_D + c * 8 + 4));
      v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) v[i] = 0.f;
    }
    float hi[8], lo[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      split_bf16(v[i], hi[i], lo[i]);
      sq = fmaf(v[i], v[i], sq);
    }
    uint8_t* dst = image + (size_t)(ch * nd + db) * VT_CHUNK_BYTES + sw128_offset(r, c);
    *reinterpret_cast<uint4*>(dst) =
        make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]), pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
    *reinterpret_cast<uint4*>(dst + VT_IMG_HALF) =
        make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]), pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
  }
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

// Streaming search, D = 64 * ND (ND = 1, 2, 4): the codebook streams through the two 64 KB buffers in units of
// (256 codewords, 64 dims) and the ND units of a chunk accumulate into the same 256-column TMEM accumulator, so the
// pipeline and the epilogue are those of D = 64; only the z tile (ND K-blocks of 16 KB per image) and the exact
// re-score / gather (D dims per row) grow.
template <typename T, int ND>
__global__ void __launch_bounds__(288, 1)
vq_search_tc_kernel(const T* __restrict__ z, const float* __restrict__ cb, const uint8_t* __restrict__ image,
                    const float* __restrict__ esq_pad, long long N, int K, int nchunks,
                    long long* __restrict__ indices, T* __restrict__ codes, double* __restrict__ sqerr) {
  constexpr bool kSplitZ = sizeof(T) == 4;   // bf16 activations have no low part
  constexpr int D = ND * VT_D;
  constexpr uint32_t A_IMG = ND * 16384;     // one image of the z tile: ND K-blocks of 128 rows x 64 bf16
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  uint8_t* A_hi = sm;
  uint8_t* A_lo = sm + A_IMG;                                   // (fp32 z only)
  uint8_t* Bbuf = sm + (kSplitZ ? 2 : 1) * A_IMG;               // 2 x 64 KB
  float* esq_s = reinterpret_cast<float*>(Bbuf + 2 * VT_CHUNK_BYTES);
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
    // stage the z tile: thread = (row, half); per 64-dim block 4 x 16-byte chunks of bf16 hi (and lo)
    const int row = tid >> 1, half = tid & 1;
    const long long grow = row0 + row;
#pragma unroll
    for (int db = 0; db < ND; ++db) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int c = half * 4 + j;
        float v[8];
        if (grow < N) {
          load8<T>(z + grow * D + db * VT_D + c * 8, v);
        } else {
#pragma unroll
          for (int i = 0; i < 8; ++i) v[i] = 0.f;
        }
        float hi[8], lo[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) split_bf16(v[i], hi[i], lo[i]);
        const uint32_t off = (uint32_t)db * 16384 + sw128_offset(row, c);
        *reinterpret_cast<uint4*>(A_hi + off) = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]),
                                                           pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
        if (kSplitZ)
          *reinterpret_cast<uint4*>(A_lo + off) = make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]),
                                                             pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
      }
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
      // ---------------- producer + MMA issuer: unit u = chunk * ND + dblock streams through buffer u & 1
      const uint32_t idesc = idesc_bf16(VT_M, VT_N);
      const int nunits = nchunks * ND;
      auto load_unit = [&](int u) {
        const int b = u & 1;
        mbar_arrive_expect_tx(&b_full[b], VT_CHUNK_BYTES);
        bulk_g2s(Bbuf + (size_t)b * VT_CHUNK_BYTES, image + (size_t)u * VT_CHUNK_BYTES, VT_CHUNK_BYTES, &b_full[b]);
      };
      load_unit(0);
      if (nunits > 1) load_unit(1);
      for (int u = 0; u < nunits; ++u) {
        const int b = u & 1;
        const int c = u / ND, db = u - c * ND;
        const int ab = c & 1;                               // accumulator of this chunk
        mbar_wait(&b_full[b], (uint32_t)((u >> 1) & 1));
        if (db == 0 && c >= 2) mbar_wait(&acc_empty[ab], (uint32_t)(((c >> 1) & 1) ^ 1));
        tc_fence_after();
        const uint32_t d_tmem = tmem_base + (uint32_t)ab * VT_N;
        const uint64_t a_hi = sw128_desc(smem_u32(A_hi + (size_t)db * 16384));
        const uint64_t a_lo = sw128_desc(smem_u32(A_lo + (size_t)db * 16384));
        const uint64_t b_hi = sw128_desc(smem_u32(Bbuf + (size_t)b * VT_CHUNK_BYTES));
        const uint64_t b_lo = sw128_desc(smem_u32(Bbuf + (size_t)b * VT_CHUNK_BYTES + VT_IMG_HALF));
#pragma unroll
        for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_hi + 2 * ks, b_hi + 2 * ks, idesc, (db > 0 || ks > 0) ? 1u : 0u);
#pragma unroll
        for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_hi + 2 * ks, b_lo + 2 * ks, idesc, 1);
        if (kSplitZ) {
#pragma unroll
          for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_lo + 2 * ks, b_hi + 2 * ks, idesc, 1);
        }
        if (db == ND - 1) umma_commit(&acc_full[ab]);
        umma_commit(&b_empty[b]);
        if (u >= 1 && u + 1 < nunits) {       // refill the buffer unit u-1 used, now that its MMAs retired
          const int pb = (u - 1) & 1;
          mbar_wait(&b_empty[pb], (uint32_t)(((u - 1) >> 1) & 1));
          load_unit(u + 1);
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

    // ---------------- exact fp32 re-score of the two best candidates: 8 lanes per row, D / 8 dims per lane
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
#pragma unroll
        for (int db = 0; db < ND; ++db) {
          float zv[8], ev[8];
          load8<T>(z + grow * D + db * VT_D + sub * 8, zv);
          load8<float>(cb + (long long)m.i1 * D + db * VT_D + sub * 8, ev);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float df = zv[i] - ev[i];
            p1 = fmaf(df, df, p1);
          }
          if (m.i2 >= 0) {
            load8<float>(cb + (long long)m.i2 * D + db * VT_D + sub * 8, ev);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const float df = zv[i] - ev[i];
              p2 = fmaf(df, df, p2);
            }
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
      for (int i = tid; i < VT_M * (D / 4); i += 256) {
        const int rr = i / (D / 4), c4 = i % (D / 4);
        const long long grow = row0 + rr;
        if (grow < N) {
          const float4 v = __ldg(reinterpret_cast<const float4*>(cb + (long long)sWin[rr] * D + c4 * 4));
          Act<T>::st4(codes + grow * D + c4 * 4, v);
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

// ---------------------------------------------------------------------------------------------- resident-codebook search
// K <= 512 (the reference's z_dim): the whole split codebook image (2 chunks x 64 KB) stays in shared memory for the
// CTA's lifetime and ONE CTA per SM walks many 128-row tiles, so the 128 KB of codebook is read from L2 once per SM
// instead of once per tile.  Two epilogue groups (4 warps each) own alternate tiles end to end -- scan the
// 256-column accumulator of each chunk for the two best codewords, re-score them exactly in fp32, write index /
// code / squared error -- and each owns one 256-column TMEM accumulator, so the MMA thread always has the other
// group's tile to work on while one group is in its latency-bound tail.
//   bf16 z (production): z tiles arrive UMMA-ready through a SWIZZLE_128B tensor map into a 4-deep ring (no thread
//       touches them before the tensor core does); the exact re-score reads the same tile back from shared memory.
//   fp32 z (parity mode): each group splits its own tile into hi / lo bf16 images (one 32 KB buffer per group).
constexpr int VR_THREADS = 320;     // warps 0-3 group 0, warps 4-7 group 1, warp 8 = loader + MMA issuer, warp 9 = z producer
constexpr int VR_STAGES = 4;        // bf16 z ring (4 x 16 KB = the two 32 KB buffers of the fp32 path)
constexpr int VR_CAND = 8;          // candidate list per row (4 per 256-codeword chunk)
constexpr int VR_OVERFLOW = 255;    // candidate count marker: list too short, scan the whole codebook exactly

template <typename T>
__global__ void __launch_bounds__(VR_THREADS, 1)
vq_search_res_kernel(const __grid_constant__ CUtensorMap tmz, const T* __restrict__ z, const float* __restrict__ cb,
                     const uint8_t* __restrict__ image, const float* __restrict__ esq_pad, long long N, int K,
                     int nchunks, int ntiles, long long* __restrict__ indices, T* __restrict__ codes,
                     double* __restrict__ sqerr) {
  constexpr bool kSplitZ = sizeof(T) == 4;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* sm = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);    // (pointer arithmetic on the __shared__ array: the compiler keeps the address space -- LDS / STS instead of generic loads)
  uint8_t* Bres = sm;                                           // nchunks x (hi 32 KB | lo 32 KB)
  uint8_t* Abuf = sm + (size_t)nchunks * VT_CHUNK_BYTES;        // fp32: [2 groups][hi 16 KB | lo 16 KB]; bf16: [4][16 KB]
  float* esq_s = reinterpret_cast<float*>(Abuf + 2 * 32768);    // [nchunks * 256]
  uint8_t* tail = reinterpret_cast<uint8_t*>(esq_s + nchunks * VT_N);
  int* sCI = reinterpret_cast<int*>(tail);                      // [2][128][VR_CAND] candidate lists
  int* sWin = reinterpret_cast<int*>(tail + 8192);              // [2][128]
  int* sCnt = reinterpret_cast<int*>(tail + 9216);              // [2][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(tail + 10240);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tail + 10368);
  __shared__ float sEmax[VR_THREADS / 32];
  uint64_t* b_full = bars;                                      // [1]
  uint64_t* a_full = bars + 1;                                  // [4]
  uint64_t* a_empty = bars + 5;                                 // [4]
  uint64_t* acc_full = bars + 9;                                // [2]
  uint64_t* acc_empty = bars + 11;                              // [2]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int ntiles_mine = ((int)blockIdx.x < ntiles) ? (ntiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

  if (tid == 0) {
    mbar_init(b_full, 1);
    for (int s = 0; s < VR_STAGES; ++s) {
      mbar_init(&a_full[s], kSplitZ ? 4 : 1);
      mbar_init(&a_empty[s], 4);
    }
    for (int g = 0; g < 2; ++g) {
      mbar_init(&acc_full[g], 1);
      mbar_init(&acc_empty[g], 4);
    }
    fence_mbar_init();
    if (!kSplitZ) tma::prefetch_desc(&tmz);
  }
  if (warp == 8) tmem_alloc<512>(tmem_slot);
  pdl_wait();                // barriers / TMEM above are independent of earlier kernels (common.cuh, PDL)
  pdl_trigger();
  float emx = 0.f;                 // max |e|^2 over the real codewords (the padding rows carry +inf)
  for (int i = tid; i < nchunks * VT_N; i += VR_THREADS) {
    const float v = __ldg(esq_pad + i);
    esq_s[i] = v;
    if (i < K) emx = fmaxf(emx, v);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) emx = fmaxf(emx, __shfl_xor_sync(0xffffffffu, emx, o));
  if (lane == 0) sEmax[warp] = emx;
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  float emax = 0.f;
#pragma unroll
  for (int w = 0; w < VR_THREADS / 32; ++w) emax = fmaxf(emax, sEmax[w]);
  emax = sqrtf(emax);

  if (warp == 9) {
    if (!kSplitZ && lane == 0) {
      // ---------------- z producer (bf16): one 128-row box per tile, rows past N are zero-filled by the TMA unit
      for (int i = 0; i < ntiles_mine; ++i) {
        const int s = i % VR_STAGES;
        mbar_wait(&a_empty[s], (uint32_t)(((i / VR_STAGES) & 1) ^ 1));
        mbar_arrive_expect_tx(&a_full[s], 16384);
        const long long row0 = ((long long)blockIdx.x + (long long)i * gridDim.x) * VT_M;
        tma::load_4d(Abuf + (size_t)s * 16384, &tmz, 0, (int)row0, 0, 0, &a_full[s]);
      }
    }
  } else if (warp == 8) {
    if (lane == 0 && ntiles_mine > 0) {
      // ---------------- codebook loader + MMA issuer
      mbar_arrive_expect_tx(b_full, (uint32_t)nchunks * VT_CHUNK_BYTES);
      for (int c = 0; c < nchunks; ++c)
        for (int h = 0; h < 2; ++h)
          bulk_g2s(Bres + (size_t)c * VT_CHUNK_BYTES + (size_t)h * VT_IMG_HALF,
                   image + (size_t)c * VT_CHUNK_BYTES + (size_t)h * VT_IMG_HALF, VT_IMG_HALF, b_full);
      const uint32_t idesc = idesc_bf16(VT_M, VT_N);
      mbar_wait(b_full, 0);
      // the two groups advance independently: whichever group's next (tile, chunk) has its operands in place and
      // its accumulator drained is served first, so one group's latency-bound tail never stalls the other's scan
      int used[2] = {0, 0};                 // accumulator uses so far, per group
      int tile_g[2] = {0, 1}, chunk_g[2] = {0, 0};
      int remaining = ntiles_mine * nchunks;
      while (remaining > 0) {
        bool progressed = false;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const int i = tile_g[g];
          if (i >= ntiles_mine) continue;
          const int c = chunk_g[g];
          // fp32: buffer g, filled once per tile of group g; bf16: ring slot i % 4
          const int slot = kSplitZ ? g : i % VR_STAGES;
          const uint32_t aph = kSplitZ ? (uint32_t)((i >> 1) & 1) : (uint32_t)((i / VR_STAGES) & 1);
          if (c == 0 && !mbar_try_wait(&a_full[slot], aph)) continue;
          if (used[g] > 0 && !mbar_try_wait(&acc_empty[g], (uint32_t)((used[g] - 1) & 1))) continue;
          ++used[g];
          --remaining;
          progressed = true;
          if (++chunk_g[g] == nchunks) {
            chunk_g[g] = 0;
            tile_g[g] += 2;
          }
          tc_fence_after();
          const uint32_t d_tmem = tmem_base + (uint32_t)g * VT_N;
          uint8_t* abase = kSplitZ ? Abuf + (size_t)slot * 32768 : Abuf + (size_t)slot * 16384;
          const uint64_t a_hi = sw128_desc(smem_u32(abase));
          const uint64_t a_lo = sw128_desc(smem_u32(abase + 16384));
          const uint64_t b_hi = sw128_desc(smem_u32(Bres + (size_t)c * VT_CHUNK_BYTES));
          const uint64_t b_lo = sw128_desc(smem_u32(Bres + (size_t)c * VT_CHUNK_BYTES + VT_IMG_HALF));
#pragma unroll
          for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_hi + 2 * ks, b_hi + 2 * ks, idesc, ks > 0);
#pragma unroll
          for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_hi + 2 * ks, b_lo + 2 * ks, idesc, 1);
          if (kSplitZ) {
#pragma unroll
            for (int ks = 0; ks < VT_D / 16; ++ks) umma_bf16(d_tmem, a_lo + 2 * ks, b_hi + 2 * ks, idesc, 1);
          }
          umma_commit(&acc_full[g]);
        }
        // nothing to issue: both groups are scanning.  This thread shares its scheduler with two scanning warps and a
        // polling warp is always eligible (ncu: 6 % of the kernel's instructions were this loop) -- back off
        if (!progressed) __nanosleep(64);
      }
    }
  } else {
    // ---------------- epilogue group g: tiles g, g+2, ... of this CTA
    // Exact-minimum search in two sweeps over each 256-column accumulator (it stays in TMEM between them):
    //   sweep 1: the chunk minimum of d = |e|^2 - 2 z.e -- one FFMA and half a 3-input FMNMX per (row, codeword);
    //   sweep 2: every codeword within `margin` of that minimum becomes a CANDIDATE (FFMA + FSETP per element; the
    //            rare hits are appended in increasing index order).
    // margin = 2^-14 |z| max|e| is > 2x the worst error of the split-bf16 GEMM (the dropped operand residuals are
    // < 2^-17 |z||e| each, doubled by the factor 2), so the fp32 argmin is always among the candidates; they are then
    // re-scored EXACTLY in fp32 in increasing index order with a strict '<' -- the lowest index wins among equal
    // distances, however many duplicates there are (vector_quantization.py:19).  A row with more candidates than the
    // list holds (degenerate codebooks: many identical codewords) falls back to an exact scan of all K codewords.
    // The previous form tracked the two best codewords in one sweep at ~10 instructions per (row, codeword).
    const int g = warp >> 2, q = warp & 3;
    const int gt = tid & 127;                         // thread within the group = TMEM lane = tile row
    int* myCand = sCI + g * (128 * VR_CAND);
    int* myCnt = sCnt + g * 128;
    int* myWin = sWin + g * 128;
    const int barid = 1 + g;
    float err_acc = 0.f;
    int nfull = 0;                                    // acc_full completions consumed by this group

    auto stage = [&](int i) {                         // fp32 only: split this group's tile into hi / lo images
      uint8_t* A_hi = Abuf + (size_t)g * 32768;
      uint8_t* A_lo = A_hi + 16384;
      const long long row0 = ((long long)blockIdx.x + (long long)i * gridDim.x) * VT_M;
#pragma unroll
      for (int rnd = 0; rnd < 2; ++rnd) {
        const int row = rnd * 64 + (gt >> 1), half = gt & 1;
        const long long grow = row0 + row;
        float v[4][8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (grow < N) {
            load8<T>(z + grow * VT_D + (half * 4 + j) * 8, v[j]);
          } else {
#pragma unroll
            for (int e = 0; e < 8; ++e) v[j][e] = 0.f;
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float hi[8], lo[8];
#pragma unroll
          for (int e = 0; e < 8; ++e) split_bf16(v[j][e], hi[e], lo[e]);
          const uint32_t off = sw128_offset(row, half * 4 + j);
          *reinterpret_cast<uint4*>(A_hi + off) = make_uint4(pack_bf16(hi[0], hi[1]), pack_bf16(hi[2], hi[3]),
                                                             pack_bf16(hi[4], hi[5]), pack_bf16(hi[6], hi[7]));
          *reinterpret_cast<uint4*>(A_lo + off) = make_uint4(pack_bf16(lo[0], lo[1]), pack_bf16(lo[2], lo[3]),
                                                             pack_bf16(lo[4], lo[5]), pack_bf16(lo[6], lo[7]));
        }
      }
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) mbar_arrive(&a_full[g]);
    };

    if (kSplitZ && g < ntiles_mine) stage(g);
    for (int i = g; i < ntiles_mine; i += 2) {
      const long long row0 = ((long long)blockIdx.x + (long long)i * gridDim.x) * VT_M;
      const int slot = i % VR_STAGES;
      const uint8_t* ztile = Abuf + (size_t)slot * 16384;           // bf16 path: this tile's rows, as TMA wrote them
      float cmin[2] = {INFINITY, INFINITY};
      int cn[2] = {0, 0};
      int cand[2][VR_CAND / 2];
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int j = 0; j < VR_CAND / 2; ++j) cand[c][j] = -1;
      float margin = 0.f;
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (c < nchunks) {
          mbar_wait_backoff(&acc_full[g], (uint32_t)(nfull & 1));
          ++nfull;
          tc_fence_after();
          const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(g * VT_N);
          uint32_t ra[32], rb[32];
          tmem_ld32(tbase, ra);
          if (c == 0) {
            // |z| of this row (the accumulator being complete implies its z tile has landed / been staged)
            float zz = 0.f;
            if (kSplitZ) {
              const long long grow = row0 + gt;
              if (grow < N) {
#pragma unroll
                for (int j = 0; j < 8; ++j) {
                  float v[8];
                  load8<T>(z + grow * VT_D + j * 8, v);
#pragma unroll
                  for (int e = 0; e < 8; ++e) zz = fmaf(v[e], v[e], zz);
                }
              }
            } else {
              mbar_wait(&a_full[slot], (uint32_t)((i / VR_STAGES) & 1));     // (complete long ago: acquires the TMA write)
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const uint4 raw = *reinterpret_cast<const uint4*>(ztile + sw128_offset(gt, j));
                const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  const float lo = __uint_as_float(w[e] << 16), hi = __uint_as_float(w[e] & 0xffff0000u);
                  zz = fmaf(lo, lo, fmaf(hi, hi, zz));
                }
              }
            }
            margin = sqrtf(zz) * emax * 6.103515625e-05f + 1e-30f;      // 2^-14 |z| max|e|
          }
          // ---- sweep 1: chunk minimum
          float m0 = INFINITY, m1 = INFINITY, m2 = INFINITY, m3 = INFINITY;
#pragma unroll 1
          for (int j = 0; j < 8; j += 2) {
            tmem_ld_wait();
            tmem_ld_fence(ra);
            tmem_ld32(tbase + (uint32_t)((j + 1) * 32), rb);
            {
              const float4* es = reinterpret_cast<const float4*>(esq_s + c * VT_N + j * 32);
#pragma unroll
              for (int e4 = 0; e4 < 8; e4 += 2) {
                const float4 e = es[e4], f = es[e4 + 1];
                m0 = fminf(fminf(m0, fmaf(-2.f, __uint_as_float(ra[4 * e4 + 0]), e.x)), fmaf(-2.f, __uint_as_float(ra[4 * e4 + 1]), e.y));
                m1 = fminf(fminf(m1, fmaf(-2.f, __uint_as_float(ra[4 * e4 + 2]), e.z)), fmaf(-2.f, __uint_as_float(ra[4 * e4 + 3]), e.w));
                m2 = fminf(fminf(m2, fmaf(-2.f, __uint_as_float(ra[4 * e4 + 4]), f.x)), fmaf(-2.f, __uint_as_float(ra[4 * e4 + 5]), f.y));
                m3 = fminf(fminf(m3, fmaf(-2.f, __uint_as_float(ra[4 * e4 + 6]), f.z)), fmaf(-2.f, __uint_as_float(ra[4 * e4 + 7]), f.w));
              }
            }
            tmem_ld_wait();
            tmem_ld_fence(rb);
            tmem_ld32(tbase + (uint32_t)(((j + 2) & 7) * 32), ra);      // (wraps to slice 0: sweep 2 starts there)
            {
              const float4* es = reinterpret_cast<const float4*>(esq_s + c * VT_N + (j + 1) * 32);
#pragma unroll
              for (int e4 = 0; e4 < 8; e4 += 2) {
                const float4 e = es[e4], f = es[e4 + 1];
                m0 = fminf(fminf(m0, fmaf(-2.f, __uint_as_float(rb[4 * e4 + 0]), e.x)), fmaf(-2.f, __uint_as_float(rb[4 * e4 + 1]), e.y));
                m1 = fminf(fminf(m1, fmaf(-2.f, __uint_as_float(rb[4 * e4 + 2]), e.z)), fmaf(-2.f, __uint_as_float(rb[4 * e4 + 3]), e.w));
                m2 = fminf(fminf(m2, fmaf(-2.f, __uint_as_float(rb[4 * e4 + 4]), f.x)), fmaf(-2.f, __uint_as_float(rb[4 * e4 + 5]), f.y));
                m3 = fminf(fminf(m3, fmaf(-2.f, __uint_as_float(rb[4 * e4 + 6]), f.z)), fmaf(-2.f, __uint_as_float(rb[4 * e4 + 7]), f.w));
              }
            }
          }
          const float mc = fminf(fminf(m0, m1), fminf(m2, m3));
          cmin[c] = mc;
          const float thr = mc + margin;
          // ---- sweep 2: candidates within the margin, in increasing index order
          int n = 0;
          auto push = [&](int k) {
            if (n == 0) cand[c][0] = k;
            else if (n == 1) cand[c][1] = k;
            else if (n == 2) cand[c][2] = k;
            else if (n == 3) cand[c][3] = k;
            ++n;
          };
          // (branch-free per element: the hits of a 32-column slice are collected in a bit mask -- FFMA + FSETP +
          // predicated OR -- and the rare set bits are walked once per slice; a branch per 4 columns made the warp run
          // the divergent push path for every group in which ANY of its 32 rows had a candidate: as many instructions
          // again as the scan itself, at 16-21 active threads per warp)
          auto scan32 = [&](const uint32_t (&rr)[32], int jj) {
            const float4* es = reinterpret_cast<const float4*>(esq_s + c * VT_N + jj * 32);
            const int kbase = c * VT_N + jj * 32;
            uint32_t hit = 0;
#pragma unroll
            for (int e4 = 0; e4 < 8; ++e4) {
              const float4 e = es[e4];
              const float d0 = fmaf(-2.f, __uint_as_float(rr[4 * e4 + 0]), e.x);
              const float d1 = fmaf(-2.f, __uint_as_float(rr[4 * e4 + 1]), e.y);
              const float d2 = fmaf(-2.f, __uint_as_float(rr[4 * e4 + 2]), e.z);
              const float d3 = fmaf(-2.f, __uint_as_float(rr[4 * e4 + 3]), e.w);
              if (d0 <= thr) hit |= 1u << (4 * e4 + 0);
              if (d1 <= thr) hit |= 1u << (4 * e4 + 1);
              if (d2 <= thr) hit |= 1u << (4 * e4 + 2);
              if (d3 <= thr) hit |= 1u << (4 * e4 + 3);
            }
            while (hit) {                                  // increasing index order
              push(kbase + __ffs((int)hit) - 1);
              hit &= hit - 1;
            }
          };
#pragma unroll 1
          for (int j = 0; j < 8; j += 2) {
            tmem_ld_wait();
            tmem_ld_fence(ra);
            tmem_ld32(tbase + (uint32_t)((j + 1) * 32), rb);
            scan32(ra, j);
            tmem_ld_wait();
            tmem_ld_fence(rb);
            if (j + 2 < 8) tmem_ld32(tbase + (uint32_t)((j + 2) * 32), ra);
            scan32(rb, j + 1);
          }
          cn[c] = n;
          tc_fence_before();
          __syncwarp();
          if (lane == 0) mbar_arrive(&acc_empty[g]);
        }
      }
      // fp32: every MMA that read this group's A buffer has retired (its last accumulator was just consumed): hand
      // the MMA thread the next tile before the latency-bound tail of this one
      if (kSplitZ && i + 2 < ntiles_mine) stage(i + 2);
      {
        // this row's candidate list: the chunks whose minimum is within the margin of the row minimum
        const float gthr = fminf(cmin[0], cmin[1]) + margin;
        int total = 0;
        bool ovf = false;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (cmin[c] <= gthr) {
            if (cn[c] > VR_CAND / 2) ovf = true;
#pragma unroll
            for (int j = 0; j < VR_CAND / 2; ++j)
              if (j < cn[c] && cand[c][j] < K) myCand[gt * VR_CAND + total++] = cand[c][j];
          }
        }
        myCnt[gt] = ovf ? VR_OVERFLOW : total;
      }
      if (!kSplitZ) mbar_wait(&a_full[slot], (uint32_t)((i / VR_STAGES) & 1));   // (complete long ago: ordering only)
      asm volatile("bar.sync %0, 128;" ::"r"(barid) : "memory");

      // exact fp32 re-score: 8 lanes per row, 16 rows per pass, 4 passes' first candidates in flight
      const int sub = gt & 7;
#pragma unroll 1
      for (int pb = 0; pb < 8; pb += 4) {
        float zv[4][8], e1[4][8];
        int cnt[4], c1[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int rr = (pb + u) * 16 + (gt >> 3);
          const long long grow = row0 + rr;
          cnt[u] = grow < N ? myCnt[rr] : 0;
          c1[u] = (cnt[u] > 0 && cnt[u] != VR_OVERFLOW) ? myCand[rr * VR_CAND] : -1;
          if (cnt[u] > 0) {
            if (kSplitZ) {
              load8<T>(z + grow * VT_D + sub * 8, zv[u]);
            } else {
              const uint4 raw = *reinterpret_cast<const uint4*>(ztile + sw128_offset(rr, sub));
              const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                zv[u][2 * e] = __uint_as_float(w[e] << 16);
                zv[u][2 * e + 1] = __uint_as_float(w[e] & 0xffff0000u);
              }
            }
            if (c1[u] >= 0) load8<float>(cb + (long long)c1[u] * VT_D + sub * 8, e1[u]);
          }
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int rr = (pb + u) * 16 + (gt >> 3);
          float best = INFINITY;
          int win = 0;
          {
            float p1 = 0.f;
            if (c1[u] >= 0) {
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float df = zv[u][e] - e1[u][e];
                p1 = fmaf(df, df, p1);
              }
            }
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) p1 += __shfl_xor_sync(0xffffffffu, p1, o);
            if (c1[u] >= 0) {
              best = p1;
              win = c1[u];
            }
          }
          // further candidates (rare): warp-uniform trip count, rows without one idle
          const int extra = (cnt[u] != VR_OVERFLOW) ? cnt[u] : 0;
          const int jmax = __reduce_max_sync(0xffffffffu, extra);
          for (int j = 1; j < jmax; ++j) {
            const int k = j < extra ? myCand[rr * VR_CAND + j] : -1;
            float pj = 0.f;
            if (k >= 0) {
              float ev[8];
              load8<float>(cb + (long long)k * VT_D + sub * 8, ev);
#pragma unroll
              for (int e = 0; e < 8; ++e) {
                const float df = zv[u][e] - ev[e];
                pj = fmaf(df, df, pj);
              }
            }
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) pj += __shfl_xor_sync(0xffffffffu, pj, o);
            if (k >= 0 && pj < best) {          // increasing index order + strict '<': lowest index among equals
              best = pj;
              win = k;
            }
          }
          // candidate list overflow (many identical codewords): exact scan of the whole codebook for this row
          if (__any_sync(0xffffffffu, cnt[u] == VR_OVERFLOW)) {
            const bool mine = cnt[u] == VR_OVERFLOW;
            for (int k = 0; k < K; ++k) {
              float pj = 0.f;
              if (mine) {
                float ev[8];
                load8<float>(cb + (long long)k * VT_D + sub * 8, ev);
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                  const float df = zv[u][e] - ev[e];
                  pj = fmaf(df, df, pj);
                }
              }
#pragma unroll
              for (int o = 4; o > 0; o >>= 1) pj += __shfl_xor_sync(0xffffffffu, pj, o);
              if (mine && pj < best) {
                best = pj;
                win = k;
              }
            }
          }
          if (sub == 0) {
            myWin[rr] = win;
            const long long grow = row0 + rr;
            if (grow < N) {
              indices[grow] = (long long)win;
              err_acc += best;
            }
          }
        }
      }
      if (!kSplitZ) {                                   // the z ring slot is free once every warp has re-scored
        __syncwarp();
        if (lane == 0) mbar_arrive(&a_empty[slot]);
      }
      asm volatile("bar.sync %0, 128;" ::"r"(barid) : "memory");
      if (codes) {
#pragma unroll 8      // (8 codebook rows in flight per thread: two L2 round trips per tile instead of four)
        for (int it = gt; it < VT_M * (VT_D / 4); it += 128) {
          const int rr = it >> 4, c4 = it & 15;
          const long long grow = row0 + rr;
          if (grow < N) {
            const float4 v = __ldg(reinterpret_cast<const float4*>(cb + (long long)myWin[rr] * VT_D + c4 * 4));
            Act<T>::st4(codes + grow * VT_D + c4 * 4, v);
          }
        }
      }
      asm volatile("bar.sync %0, 128;" ::"r"(barid) : "memory");      // myCand / myCnt / myWin are reused by the next tile
    }
    if (sqerr) {
      const float e = warp_sum(err_acc);
      if (lane == 0 && e != 0.f) atomicAdd(sqerr, (double)e);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 8) {
    __syncwarp();
    tmem_dealloc<512>(tmem_base);
  }
}

static size_t vq_res_smem_bytes(int nchunks) {
  return 1024 + (size_t)nchunks * VT_CHUNK_BYTES + 2 * 32768 + (size_t)nchunks * VT_N * 4 + 10752;
}

// dynamic shared memory of the streaming kernel: z tile images (hi, + lo for fp32 z), two codebook units, |e|^2, tail
static size_t vq_tc_smem_bytes(int nchunks, int nd, bool split_z) {
  return 1024 + (size_t)(split_z ? 2 : 1) * nd * 16384 + 2 * (size_t)VT_CHUNK_BYTES + (size_t)nchunks * VT_N * 4 + 5376;
}
constexpr size_t VT_SMEM_MAX = 227 * 1024;

int g_vq_resident = 1;     // development switch (nsg_vq_resident): 0 = always the streaming kernel

// D = 64, 128 or 256 and as many codewords as the kernel's shared memory holds |e|^2 for beside a bf16 z tile
// (K <= 8192 at D = 64 / 128, 7168 at D = 256).  fp32 z needs a second z image: where that no longer fits (D = 256, or
// D = 128 with K > 7168) vq_forward_tc runs the exact CUDA-core search instead -- same indices, no tensor cores.
long long vq_tc_workspace_bytes(int k, int d) {
  if ((d != 64 && d != 128 && d != 256) || k < 1 || k > VT_MAX_CHUNKS * VT_N) return -1;
  const long long nchunks = (k + VT_N - 1) / VT_N;
  const int nd = d / VT_D;
  if (vq_tc_smem_bytes((int)nchunks, nd, false) > VT_SMEM_MAX) return -1;
  return nchunks * nd * (long long)VT_CHUNK_BYTES + nchunks * VT_N * 4;
}

int vq_forward_simt(const void* z, const float* codebook, long long n, int k, int d, int dtype, long long* indices,
                    void* codes, double* sqerr, cudaStream_t stream);

template <typename T, int ND>
static int launch_vq_tc(unsigned grid, size_t smem, cudaStream_t stream, const void* z, const float* codebook,
                        const uint8_t* image, const float* esq_pad, long long n, int k, int nchunks, long long* indices,
                        void* codes, double* sqerr) {
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(vq_search_tc_kernel<T, ND>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         (int)VT_SMEM_MAX);
    if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(vq_search_tc_kernel)");
    configured = true;
  }
  vq_search_tc_kernel<T, ND><<<grid, 288, smem, stream>>>((const T*)z, codebook, image, esq_pad, n, k, nchunks, indices,
                                                          (T*)codes, sqerr);
  NSG_CHECK_LAUNCH("vq_search_tc_kernel");
  return NSG_OK;
}

int vq_forward_tc(const void* z, const float* codebook, long long n, int k, int d, int dtype, long long* indices,
                  void* codes, double* sqerr, void* ws, long long ws_bytes, cudaStream_t stream) {
  const long long need = vq_tc_workspace_bytes(k, d);
  if (need < 0) {
    set_error("vq_forward(mode=1): tensor-core search supports D in {64, 128, 256} and K <= %d (7168 at D = 256) "
              "(got K=%d, D=%d)", VT_MAX_CHUNKS * VT_N, k, d);
    return NSG_E_UNSUPPORTED;
  }
  NSG_REQUIRE(ws && ws_bytes >= need, "vq_forward(mode=1): workspace of %lld bytes required (got %lld)", need, ws_bytes);
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(ws) & 15) == 0, "vq_forward(mode=1): workspace must be 16-byte aligned");
  const int nchunks = (k + VT_N - 1) / VT_N;
  const int nd = d / VT_D;
  const size_t smem = vq_tc_smem_bytes(nchunks, nd, dtype == NSG_F32);
  if (smem > VT_SMEM_MAX)      // fp32 z whose hi + lo images do not fit: the exact CUDA-core search (see above)
    return vq_forward_simt(z, codebook, n, k, d, dtype, indices, codes, sqerr, stream);
  uint8_t* image = static_cast<uint8_t*>(ws);
  float* esq_pad = reinterpret_cast<float*>(image + (size_t)nchunks * nd * VT_CHUNK_BYTES);
  launch_pdl(vq_prep_kernel, dim3(nchunks * 8), dim3(256), 0, stream, codebook, k, nd, image, esq_pad);
  NSG_CHECK_LAUNCH("vq_prep_kernel");
  return NSG_DISPATCH_DTYPE(dtype, [&]() -> int {
    const unsigned grid = (unsigned)((n + VT_M - 1) / VT_M);
    if (nd == 1 && nchunks <= 2 && g_vq_resident) {      // D = 64, K <= 512: persistent CTAs, codebook resident in shared memory
      static bool configured_res = false;
      if (!configured_res) {
        cudaError_t e = cudaFuncSetAttribute(vq_search_res_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             (int)vq_res_smem_bytes(2));
        if (e != cudaSuccess) return cuda_fail(e, "cudaFuncSetAttribute(vq_search_res_kernel)");
        configured_res = true;
      }
      const unsigned pgrid = grid < (unsigned)num_sms() ? grid : (unsigned)num_sms();
      CUtensorMap tmz;
      memset(&tmz, 0, sizeof(tmz));
      if (sizeof(T) == 2) {       // z (n, 64) bf16 as a 64-channel "image" one pixel high: 128-row SWIZZLE_128B boxes
        NSG_REQUIRE((reinterpret_cast<uintptr_t>(z) & 15) == 0, "vq_forward(mode=1): z must be 16-byte aligned");
        if (int rc = tma::encode_nhwc64(&tmz, z, n, 1, 1, VT_D, (long long)n * VT_D, (long long)n * VT_D, VT_M, 1))
          return rc;
      }
      launch_pdl(vq_search_res_kernel<T>, dim3(pgrid), dim3(VR_THREADS), vq_res_smem_bytes(nchunks), stream, 
          tmz, (const T*)z, codebook, image, esq_pad, n, k, nchunks, (int)grid, indices, (T*)codes, sqerr);
      NSG_CHECK_LAUNCH("vq_search_res_kernel");
      return NSG_OK;
    }
    if (nd == 1) return launch_vq_tc<T, 1>(grid, smem, stream, z, codebook, image, esq_pad, n, k, nchunks, indices, codes, sqerr);
    if (nd == 2) return launch_vq_tc<T, 2>(grid, smem, stream, z, codebook, image, esq_pad, n, k, nchunks, indices, codes, sqerr);
    return launch_vq_tc<T, 4>(grid, smem, stream, z, codebook, image, esq_pad, n, k, nchunks, indices, codes, sqerr);
  });
}

}  // namespace nsg

namespace nsg {
int vq_tc_hang_init(unsigned int* host_mapped_words) {
  cudaError_t e = cudaMemcpyToSymbol(tc::g_hang_words, &host_mapped_words, sizeof(host_mapped_words));
  return e == cudaSuccess ? NSG_OK : cuda_fail(e, "cudaMemcpyToSymbol(g_hang_words)");
}
}  // namespace nsg
