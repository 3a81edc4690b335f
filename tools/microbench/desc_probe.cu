// Probe: which tcgen05 shared-memory descriptors let ONE activation box (with halo pixels) serve every tap of a 3 x 3
// convolution?  A tile is 8 pixels along the box's fast pixel axis x 16 lines; tap (dh, dw) reads the sub-block that
// starts (1 + dh) lines and (1 + dw) pixels into the box.  Vertical shifts move the start by whole lines; horizontal
// shifts by single pixels, i.e. off the 1024-byte swizzle atom -- the question is what the hardware then needs.
// Variants (A operand; B is the verified K-major SWIZZLE_128B weight image):
//   sw128 / P pixels per line: rows of 128 B, chunk16 ^= (byte address >> 7) & 7 as TMA writes them, SBO = P * 128,
//                              matrix-base-offset field = 0 or (start >> 7) & 7
//   none  / P pixels per line: no swizzle, [channel group of 8][line][pixel][16 B], SBO = P * 16, LBO = lines * P * 16
// Prints, per variant and tap, the largest |D - reference|.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -I../../neural_sound_generation_b200/csrc desc_probe.cu -o desc_probe
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>
#include "common.cuh"
#include "tc_common.cuh"
using namespace nsg;
using namespace nsg::tc;

constexpr int LINES = 18, C = 64;

struct Variant {
  int swizzle;      // 1 = SWIZZLE_128B, 0 = none (interleaved core matrices)
  int P;            // pixels per box line
  int base_mode;    // sw128: 0 = base-offset field 0, 1 = (start >> 7) & 7
};

__device__ __forceinline__ uint16_t bf16_bits(float v) {
  __nv_bfloat16 b = __float2bfloat16_rn(v);
  return *reinterpret_cast<uint16_t*>(&b);
}

// box value at (line, px, ch): small integers, exactly representable in bf16
__host__ __device__ inline float box_val(int line, int px, int ch) { return (float)(((line * 31 + px * 7 + ch * 3) % 17) - 8); }
__host__ __device__ inline float w_val(int co, int ci) { return (float)(((co * 5 + ci * 11) % 13) - 6) * 0.125f; }

__global__ void __launch_bounds__(128, 1) probe_kernel(Variant v, int dh, int dw, float* out /* [128][64] */) {
  extern __shared__ uint8_t raw[];
  uint8_t* sm = raw + ((1024u - (smem_u32(raw) & 1023u)) & 1023u);
  uint8_t* sB = sm;                   // 64 x 64 weights, K-major SWIZZLE_128B (8 KB)
  uint8_t* sA = sm + 8192;            // the box
  __shared__ uint64_t bar;
  __shared__ uint32_t slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  const int P = v.P;
  for (int i = tid; i < 64 * 8; i += 128) {          // weights: row co, chunk c (8 input channels)
    const int co = i >> 3, c = i & 7;
    uint32_t w[4];
    for (int k = 0; k < 4; ++k)
      w[k] = (uint32_t)bf16_bits(w_val(co, c * 8 + 2 * k)) | ((uint32_t)bf16_bits(w_val(co, c * 8 + 2 * k + 1)) << 16);
    *reinterpret_cast<uint4*>(sB + sw128_offset(co, c)) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  for (int i = tid; i < LINES * P * 8; i += 128) {   // the box: (line, px, chunk of 8 channels)
    const int c = i & 7, px = (i >> 3) % P, line = (i >> 3) / P;
    uint32_t w[4];
    for (int k = 0; k < 4; ++k)
      w[k] = (uint32_t)bf16_bits(box_val(line, px, c * 8 + 2 * k)) | ((uint32_t)bf16_bits(box_val(line, px, c * 8 + 2 * k + 1)) << 16);
    uint32_t off;
    if (v.swizzle) {
      const uint32_t row = (uint32_t)(line * P + px);            // 128-byte row index from the 1024-aligned base
      off = row * 128u + (((uint32_t)c ^ (row & 7u)) << 4);
    } else {
      off = ((uint32_t)(c * LINES + line) * P + px) * 16u;
    }
    *reinterpret_cast<uint4*>(sA + off) = make_uint4(w[0], w[1], w[2], w[3]);
  }
  if (tid == 0) {
    mbar_init(&bar, 1);
    fence_mbar_init();
  }
  fence_proxy_async_smem();
  if (warp == 0) tmem_alloc<64>(&slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tm = slot;
  if (tid == 0) {
    const uint32_t idesc = idesc_bf16(128, 64);
    const uint64_t b = sw128_desc(smem_u32(sB));
    for (int ks = 0; ks < 4; ++ks) {
      uint64_t a;
      if (v.swizzle) {
        const uint32_t start = smem_u32(sA) + (uint32_t)((1 + dh) * P + (1 + dw)) * 128u + (uint32_t)ks * 32u;
        const uint32_t sbo = (uint32_t)P * 128u;
        const uint64_t bo = v.base_mode ? (uint64_t)((start >> 7) & 7u) : 0ull;
        a = (uint64_t)((start >> 4) & 0x3FFFu) | (1ull << 16) | ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46) |
            (bo << 49) | (2ull << 61);
      } else {
        const uint32_t lbo = (uint32_t)LINES * P * 16u, sbo = (uint32_t)P * 16u;
        const uint32_t start = smem_u32(sA) + (uint32_t)((1 + dh) * P + (1 + dw)) * 16u + (uint32_t)ks * 2u * lbo;
        a = (uint64_t)((start >> 4) & 0x3FFFu) | ((uint64_t)((lbo >> 4) & 0x3FFFu) << 16) |
            ((uint64_t)((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
      }
      umma_bf16(tm, a, b + 2 * ks, idesc, ks > 0);
    }
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  tc_fence_after();
  uint32_t acc[32];
  for (int half = 0; half < 2; ++half) {
    tmem_ld32(tm + ((uint32_t)(warp * 32) << 16) + (uint32_t)half * 32, acc);
    tmem_ld_wait();
    for (int j = 0; j < 32; ++j) out[(size_t)tid * 64 + half * 32 + j] = __uint_as_float(acc[j]);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc<64>(tm);
}

int main() {
  float* d_out;
  cudaMalloc(&d_out, 128 * 64 * sizeof(float));
  std::vector<float> h(128 * 64);
  const Variant vs[] = {{1, 8, 0}, {1, 16, 0}, {1, 16, 1}, {1, 10, 0}, {1, 10, 1}, {1, 12, 0}, {1, 12, 1}, {0, 10, 0}, {0, 16, 0}};
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
  for (const Variant& v : vs) {
    printf("%s P=%2d base_offset=%s :", v.swizzle ? "sw128" : "none ", v.P, v.swizzle ? (v.base_mode ? "phase" : "0    ") : "-    ");
    for (int dh = -1; dh <= 1; ++dh)
      for (int dw = -1; dw <= 1; ++dw) {
        if (v.P == 8 && dw != -1) continue;           // no halo pixels: only the aligned start exists (dw = -1 -> offset 0)
        cudaMemset(d_out, 0, 128 * 64 * sizeof(float));
        probe_kernel<<<1, 128, 100 * 1024>>>(v, dh, dw, d_out);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) {
          printf(" [CUDA error %s]\n", cudaGetErrorString(e));
          return 1;
        }
        cudaMemcpy(h.data(), d_out, h.size() * sizeof(float), cudaMemcpyDeviceToHost);
        double worst = 0;
        for (int r = 0; r < 128; ++r) {
          const int line = r >> 3, px = r & 7;
          for (int co = 0; co < 64; ++co) {
            double ref = 0;
            for (int ci = 0; ci < 64; ++ci) ref += (double)box_val(line + 1 + dh, px + 1 + dw, ci) * w_val(co, ci);
            worst = fmax(worst, fabs(ref - h[r * 64 + co]));
          }
        }
        printf(" (%+d,%+d)%s", dh, dw, worst < 1e-3 ? "ok" : "BAD");
      }
    printf("\n");
  }
  return 0;
}
