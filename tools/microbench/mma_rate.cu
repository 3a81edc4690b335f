// Microbenchmark: cycles per tcgen05.mma (kind::f16, bf16 operands in shared memory, cta_group::1) as a function of N,
// of the number of independent accumulators the instructions rotate over, and of how the issuing thread is selected.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../neural_sound_generation_b200/csrc mma_rate.cu -o mma_rate
#include <cstdio>
#include "common.cuh"
#include "tc_common.cuh"
using namespace nsg;
using namespace nsg::tc;

template <int N, int CHAINS, int MODE>
__global__ void __launch_bounds__(320, 1) rate_kernel(unsigned long long* out, int iters) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint64_t bar2;      // MODE 2/3: eight more warps wait on this one while the instructions are issued
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (16384 + N * 128) / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(sm)[i] = 0;
  if (tid == 0) {
    mbar_init(&bar, 1);
    mbar_init(&bar2, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  const uint32_t idesc = idesc_bf16(128, N);
  const uint64_t a = sw128_desc(smem_u32(sm)), b = sw128_desc(smem_u32(sm + 16384));
  if (warp == 0) {
    if (MODE == 0 || MODE >= 2) {    // one lane in a divergent branch (what the conv kernels do)
      if (lane == 0) {
        const unsigned long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
#pragma unroll
          for (int c = 0; c < CHAINS; ++c) umma_bf16(tm + c * (512 / CHAINS), a + 2 * (c & 3), b + 2 * (c & 3), idesc, 1);
        }
        umma_commit(&bar);
        const unsigned long long t1 = clock64();
        mbar_wait(&bar, 0);
        const unsigned long long t2 = clock64();
        out[0] = t1 - t0;
        out[1] = t2 - t0;
        if (MODE >= 2) mbar_arrive(&bar2);
      }
    } else {                         // whole warp runs the loop, elect.sync picks the issuing lane per instruction
      const unsigned long long t0 = clock64();
      for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int c = 0; c < CHAINS; ++c) {
          asm volatile(
              "{\n\t.reg .pred p;\n\t"
              "elect.sync _|p, 0xffffffff;\n\t"
              "@p tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, 1;\n\t}" ::"r"(tm + c * (512 / CHAINS)),
              "l"(a + 2 * (c & 3)), "l"(b + 2 * (c & 3)), "r"(idesc)
              : "memory");
        }
      }
      asm volatile(
          "{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\t"
          "@p tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(smem_u32(&bar))
          : "memory");
      const unsigned long long t1 = clock64();
      mbar_wait(&bar, 0);
      const unsigned long long t2 = clock64();
      if (lane == 0) {
        out[0] = t1 - t0;
        out[1] = t2 - t0;
      }
    }
  }
  if (MODE >= 2 && warp >= 2) {
    if (MODE == 2) {
      mbar_wait(&bar2, 0);                    // every lane polls (what the conv epilogue warps do)
    } else {
      while (!mbar_try_wait(&bar2, 0)) __nanosleep(200);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

// The conv kernel's issue pattern: 9 taps x 4 K steps per tile, per-tap descriptors from (VAR 0) a shared-memory table
// read by the issuing lane, (VAR 1) arithmetic progressions, (VAR 2) table values fetched one tap ahead,
// (VAR 3) the whole warp runs the loop convergently, lane-held table values are broadcast by shuffle, elect.sync issues.
template <int VAR>
__global__ void __launch_bounds__(128, 1) conv_like_kernel(unsigned long long* out, int tiles, int rnd) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  __shared__ uint32_t sTapA[16], sTapB[16];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < (3 * 24576 + 9 * 8192) / 4; i += 128) {
    uint32_t h = (uint32_t)i * 2654435761u + blockIdx.x * 97u;
    h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
    // two bf16 values in [-2, 2): random mantissas, exponents around 1
    reinterpret_cast<uint32_t*>(sm)[i] = rnd ? ((h & 0x807F807Fu) | 0x3F803F80u) : 0u;
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  const uint32_t sW = smem_u32(sm + 3 * 24576);
  if (tid < 9) {
    sTapA[tid] = ((tid % 3) * 4096u) >> 4;
    sTapB[tid] = (((sW + tid * 8192u) >> 4) & 0x3FFFu) | (1u << 16);
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc<512>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  const uint32_t idesc = idesc_bf16(128, 64);
  constexpr uint64_t HI = (64ull << 32) | (1ull << 46) | (2ull << 61);
  const uint32_t a_base = ((smem_u32(sm) >> 4) & 0x3FFFu) | (1u << 16);
  if (warp == 0) {
    if (VAR <= 2) {
      if (lane == 0) {
        const unsigned long long t0 = clock64();
        for (int i = 0; i < tiles; ++i) {
          const uint32_t d = tm + (i & 1) * 64;
          int t = 0;
          uint32_t na = sTapA[0], nb = sTapB[0];
          for (int g = 0; g < 3; ++g) {
            const uint32_t box = a_base + g * (24576u >> 4);
            for (int j = 0; j < 3; ++j, ++t) {
              uint32_t ao, bo;
              if (VAR == 0) {
                ao = sTapA[t]; bo = sTapB[t];
              } else if (VAR == 1) {
                ao = j * (4096u >> 4); bo = (((sW >> 4) & 0x3FFFu) | (1u << 16)) + (g + 3 * j) * (8192u >> 4);
              } else {
                ao = na; bo = nb;
                na = sTapA[(t + 1) & 15]; nb = sTapB[(t + 1) & 15];
              }
              const uint64_t a = HI | (uint64_t)(box + ao), b = HI | (uint64_t)bo;
              umma_bf16(d, a, b, idesc, t != 0);
              umma_bf16(d, a + 2, b + 2, idesc, 1);
              umma_bf16(d, a + 4, b + 4, idesc, 1);
              umma_bf16(d, a + 6, b + 6, idesc, 1);
            }
          }
        }
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        if (blockIdx.x == 0) out[0] = clock64() - t0;
      }
    } else {
      const uint32_t myA = sTapA[lane & 15], myB = sTapB[lane & 15];
      const unsigned long long t0 = clock64();
      for (int i = 0; i < tiles; ++i) {
        const uint32_t d = tm + (i & 1) * 64;
        int t = 0;
        for (int g = 0; g < 3; ++g) {
          const uint32_t box = a_base + g * (24576u >> 4);
          for (int j = 0; j < 3; ++j, ++t) {
            const uint32_t ao = __shfl_sync(0xffffffffu, myA, t), bo = __shfl_sync(0xffffffffu, myB, t);
            const uint64_t a = HI | (uint64_t)(box + ao), b = HI | (uint64_t)bo;
#pragma unroll
            for (int ks = 0; ks < 4; ++ks)
              asm volatile(
                  "{\n\t.reg .pred p, q;\n\t"
                  "setp.ne.b32 q, %4, 0;\n\t"
                  "elect.sync _|p, 0xffffffff;\n\t"
                  "@p tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, q;\n\t}" ::"r"(d),
                  "l"(a + 2 * ks), "l"(b + 2 * ks), "r"(idesc), "r"((uint32_t)((t | ks) != 0))
                  : "memory");
          }
        }
      }
      if (lane == 0) {
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        out[0] = clock64() - t0;
      }
      __syncwarp();
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<512>(tm);
}

template <int VAR>
static void run_conv_like(const char* what, int grid = 1, int rnd = 0) {
  unsigned long long* d;
  cudaMalloc(&d, 16);
  const int tiles = 32;
  const size_t smem = 1024 + 3 * 24576 + 9 * 8192;
  cudaFuncSetAttribute(conv_like_kernel<VAR>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int rep = 0; rep < 2; ++rep) conv_like_kernel<VAR><<<grid, 128, smem>>>(d, tiles, rnd);
  unsigned long long h[2] = {0, 0};
  cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  printf("conv-like issue loop: %-44s grid %3d %s data %7.1f cyc/mma%s\n", what, grid, rnd ? "random" : "zero  ", h[0] / (36.0 * tiles), e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}

template <int N, int CHAINS, int MODE>
static void run(const char* what) {
  unsigned long long* d;
  cudaMalloc(&d, 16);
  const int iters = 256;
  const size_t smem = 1024 + 16384 + N * 128;
  cudaFuncSetAttribute(rate_kernel<N, CHAINS, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  for (int rep = 0; rep < 2; ++rep) rate_kernel<N, CHAINS, MODE><<<1, MODE >= 2 ? 320 : 128, smem>>>(d, iters);
  unsigned long long h[2] = {0, 0};
  cudaError_t e = cudaMemcpy(h, d, 16, cudaMemcpyDeviceToHost);
  const double n = (double)iters * CHAINS;
  printf("%-34s N=%3d chains=%d  issue %7.1f cyc/mma  complete %7.1f cyc/mma  (floor %d)%s\n", what, N, CHAINS, h[0] / n,
         h[1] / n, 128 * N / 256, e == cudaSuccess ? "" : cudaGetErrorString(e));
  cudaFree(d);
}

int main() {
  run_conv_like<0>("per-tap table in shared memory");
  run_conv_like<1>("arithmetic descriptors");
  run_conv_like<0>("per-tap table in shared memory", 1, 1);
  run_conv_like<0>("per-tap table in shared memory", 148, 0);
  run_conv_like<0>("per-tap table in shared memory", 148, 1);
  run_conv_like<2>("table, fetched one tap ahead");
  run_conv_like<3>("convergent warp, shuffle broadcast, elect.sync");
  run<64, 1, 0>("single lane, divergent");
  run<64, 4, 0>("single lane, divergent");
  run<64, 8, 0>("single lane, divergent");
  run<64, 1, 1>("elect.sync, convergent warp");
  run<64, 4, 1>("elect.sync, convergent warp");
  run<64, 8, 1>("elect.sync, convergent warp");
  run<64, 1, 2>("single lane + 8 polling warps");
  run<64, 1, 3>("single lane + 8 warps, nanosleep");
  run<128, 1, 0>("single lane, divergent");
  run<128, 4, 0>("single lane, divergent");
  run<128, 4, 1>("elect.sync, convergent warp");
  run<256, 1, 0>("single lane, divergent");
  run<256, 2, 0>("single lane, divergent");
  run<256, 2, 1>("elect.sync, convergent warp");
  return 0;
}
