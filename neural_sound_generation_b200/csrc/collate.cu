// Device half of the mel batch assembly (the reference's collate_fn, /root/reference/src/dataloader.py:410-415):
// the host copies each utterance's KEPT frames (after the random crop of dataloader.py:356-362), time-major and back to
// back, into one pinned staging buffer; this kernel zero-pads them to the longest item and transposes
// (T_i, n_mels) -> (n_mels, T) into the (B, n_mels, T) fp32 batch train_vqvae reads (train.py:115) and, in the same
// pass, into the bf16 copy the tensor-core encoder reads (a (B, n_mels, T, 1) NHWC tensor has the same layout), so the
// step needs no separate conversion launch.  Pure data movement: bit-exact against numpy.
#include "common.cuh"

namespace nsg {

// grid (ceil(T / 32), ceil(n_mels / 32), B), block (32, 8): 32 x 32 tiles through shared memory, both sides coalesced
__global__ void __launch_bounds__(256)
mel_collate_kernel(const float* __restrict__ frames, const int* __restrict__ offs, int n_mels, int t_out,
                   float* __restrict__ out_f32, __nv_bfloat16* __restrict__ out_bf16) {
  __shared__ float tile[32][33];
  pdl_wait();
  pdl_trigger();
  const int b = blockIdx.z;
  const int f0 = offs[b], len = offs[b + 1] - f0;          // this item's first frame in `frames` and its frame count
  const int t0 = blockIdx.x * 32, m0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {              // rows of the tile = frames, columns = mel bins (contiguous)
    const int t = t0 + i, m = m0 + threadIdx.x;
    tile[i][threadIdx.x] = (t < len && m < n_mels) ? __ldg(frames + (long long)(f0 + t) * n_mels + m) : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {              // rows = mel bins, columns = frames (contiguous in the output)
    const int m = m0 + i, t = t0 + threadIdx.x;
    if (m < n_mels && t < t_out) {
      const float v = tile[threadIdx.x][i];
      const long long o = ((long long)b * n_mels + m) * t_out + t;
      if (out_f32) out_f32[o] = v;
      if (out_bf16) out_bf16[o] = __float2bfloat16_rn(v);
    }
  }
}

}  // namespace nsg

using namespace nsg;

extern "C" int nsg_mel_collate(const float* frames, const int* offsets, int batch, int n_mels, int t_out, float* out_f32,
                               void* out_bf16, void* stream) {
  NSG_REQUIRE(frames && offsets && batch >= 0 && n_mels > 0 && t_out >= 0, "mel_collate: bad arguments");
  NSG_REQUIRE(out_f32 || out_bf16, "mel_collate: no output");
  if (batch == 0 || t_out == 0) return NSG_OK;
  NSG_REQUIRE(batch <= 65535, "mel_collate: batch %d exceeds the grid limit", batch);
  launch_pdl(mel_collate_kernel, dim3((t_out + 31) / 32, (n_mels + 31) / 32, batch), dim3(32, 8), 0,
             static_cast<cudaStream_t>(stream), frames, offsets, n_mels, t_out, out_f32,
             static_cast<__nv_bfloat16*>(out_bf16));
  NSG_CHECK_LAUNCH("mel_collate_kernel");
  return NSG_OK;
}
