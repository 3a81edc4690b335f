// mel <-> audio on the device: the step after decode in the reference's main.py:164-197.
//   /root/reference/src/audio_tacotron.py:70-78    melspectrogram      (pre-emphasis, STFT, mel basis, dB, normalise)
//   /root/reference/src/audio_tacotron.py:99-116   inv_mel_spectrogram (denormalise, pinv(mel basis), ** power, Griffin-Lim,
//                                                                       inverse pre-emphasis)
//   /root/reference/src/audio_tacotron.py:142-161  _griffin_lim / _stft / _istft = librosa.stft / librosa.istft
// The reference does all of this on the host in numpy / librosa (60 x (STFT + ISTFT) of the whole utterance, a few
// seconds per sample).  Here:
//   * the FFT is a shared-memory Stockham radix-4 (+ one radix-2 stage when log2 n is odd) of n_fft complex points by
//     one CTA; frames are REAL, so a CTA transforms two frames at once (frame a in the real, frame b in the imaginary
//     part; X_a = (Z[k] + conj Z[n-k]) / 2, X_b = (Z[k] - conj Z[n-k]) / 2i) and inverts two Hermitian spectra at once;
//   * Griffin-Lim is ONE persistent launch for all its iterations: CTA c owns the frame pairs c, c + grid, ...; an
//     iteration is "rebuild my frames' samples from the previous iteration's windowed inverse transforms (overlap-add
//     gather + window sum-of-squares normalisation + librosa's centred reflect padding), FFT, keep the phase, put the
//     target magnitude back, inverse FFT, window, store" and ONE grid barrier; the signal never exists as a separate
//     tensor between iterations and everything stays in L2 (3.5 MB for a 10 s utterance);
//   * pre-emphasis is folded into the STFT's frame load; its inverse (a one-pole IIR) is a block-wide scan of affine
//     maps with the carry handed from chunk to chunk, exact for any coefficient.
// Arithmetic: fp32 (the reference: float32 mel algebra, complex128 Griffin-Lim); parity bars in tests/test_gpu_audio.py.
#include <mutex>
#include "common.cuh"

namespace nsg {

constexpr int AU_THREADS = 256;
constexpr int AU_MAX_FFT = 4096;

__device__ __forceinline__ float2 cmul(float2 a, float2 b) {
  return make_float2(fmaf(a.x, b.x, -a.y * b.y), fmaf(a.x, b.y, a.y * b.x));
}

// shared-memory layout of one CTA: two n-point complex buffers, n twiddles exp(-2 pi i m / n), the n-point window
struct AuSmem {
  float2* a;
  float2* b;
  float2* tw;
  float* win;
};
__device__ __forceinline__ AuSmem au_carve(uint8_t* raw, int n) {
  AuSmem s;
  s.a = reinterpret_cast<float2*>(raw);
  s.b = s.a + n;
  s.tw = s.b + n;
  s.win = reinterpret_cast<float*>(s.tw + n);
  return s;
}
static size_t au_smem_bytes(int n) { return (size_t)n * (3 * sizeof(float2) + sizeof(float)); }

// twiddles and the periodic Hann window (scipy.signal.get_window('hann', n, fftbins=True) = 0.5 - 0.5 cos(2 pi j / n))
__device__ __forceinline__ void au_tables(const AuSmem& s, int n) {
  for (int m = threadIdx.x; m < n; m += blockDim.x) {
    float sn, cs;
    sincospif(2.0f * (float)m / (float)n, &sn, &cs);
    s.tw[m] = make_float2(cs, -sn);
    s.win[m] = 0.5f - 0.5f * cs;
  }
  __syncthreads();
}

// forward DFT of x[0..n) (n a power of two >= 4); the result is in the returned buffer (x or y), natural order.
// Stockham autosort: every stage reads x[i + q t] and writes y[j + q p], no bit reversal.
__device__ __forceinline__ float2* au_fft(float2* x, float2* y, const float2* tw, int n) {
  int p = 1;
  while (p * 4 <= n) {
    const int t = n >> 2, step = n / (4 * p);
    for (int i = threadIdx.x; i < t; i += blockDim.x) {
      const int k = i & (p - 1), j = ((i - k) << 2) + k, m = k * step;
      const float2 u0 = x[i], u1 = cmul(tw[m], x[i + t]), u2 = cmul(tw[2 * m], x[i + 2 * t]),
                   u3 = cmul(tw[3 * m], x[i + 3 * t]);
      const float2 v0 = make_float2(u0.x + u2.x, u0.y + u2.y), v1 = make_float2(u0.x - u2.x, u0.y - u2.y);
      const float2 v2 = make_float2(u1.x + u3.x, u1.y + u3.y);
      const float2 v3 = make_float2(u1.y - u3.y, -(u1.x - u3.x));            // (u1 - u3) * -i
      y[j] = make_float2(v0.x + v2.x, v0.y + v2.y);
      y[j + p] = make_float2(v1.x + v3.x, v1.y + v3.y);
      y[j + 2 * p] = make_float2(v0.x - v2.x, v0.y - v2.y);
      y[j + 3 * p] = make_float2(v1.x - v3.x, v1.y - v3.y);
    }
    __syncthreads();
    float2* sw = x; x = y; y = sw;
    p <<= 2;
  }
  if (p < n) {
    const int t = n >> 1, step = n / (2 * p);
    for (int i = threadIdx.x; i < t; i += blockDim.x) {
      const int k = i & (p - 1), j = ((i - k) << 1) + k;
      const float2 u0 = x[i], u1 = cmul(tw[k * step], x[i + t]);
      y[j] = make_float2(u0.x + u1.x, u0.y + u1.y);
      y[j + p] = make_float2(u0.x - u1.x, u0.y - u1.y);
    }
    __syncthreads();
    float2* sw = x; x = y; y = sw;
  }
  return x;
}

// the two real frames' spectra out of the packed transform: k in [0, n/2]
__device__ __forceinline__ void au_split(const float2* Z, int n, int k, float2* xa, float2* xb) {
  const float2 z = Z[k], w = Z[(n - k) & (n - 1)];
  *xa = make_float2(0.5f * (z.x + w.x), 0.5f * (z.y - w.y));
  *xb = make_float2(0.5f * (z.y + w.y), -0.5f * (z.x - w.x));
}

// Hermitian spectra Ya, Yb (bins 0..n/2; the imaginary parts of bins 0 and n/2 are ignored, as numpy's irfft does) ->
// conj(Ya + i Yb) on all n bins, the input of the forward transform that yields the inverse one:
// ifft(Z) = conj(fft(conj Z)) / n, so frame a = Re(fft) / n and frame b = -Im(fft) / n.
__device__ __forceinline__ void au_pack_inverse(float2* dst, int n, int k, float2 ya, float2 yb) {
  if (k == 0 || k == n / 2) {
    dst[k] = make_float2(ya.x, -yb.x);
  } else {
    dst[k] = make_float2(ya.x - yb.y, -(ya.y + yb.x));           // conj(Ya + i Yb)
    dst[n - k] = make_float2(ya.x + yb.y, -(yb.x - ya.y));       // conj(conj Ya + i conj Yb)
  }
}

// sample `m` of librosa's centred, reflect-padded signal (np.pad(y, n_fft // 2, mode='reflect')) -> index into y
__device__ __forceinline__ long long au_reflect(long long m, int n_fft, long long len) {
  long long i = m - n_fft / 2;
  if (i < 0) i = -i;
  if (i >= len) i = 2 * (len - 1) - i;
  return i;
}

// sample i (0 <= i < hop (T - 1)) of librosa.istft's output, gathered from the windowed inverse transforms
// F [T][n_fft]: overlap-add of the frames that cover it (ascending frame order), divided by the window's
// sum-of-squares envelope where that exceeds tiny(float32), with the n_fft / 2 centring trim
__device__ __forceinline__ float au_ola(const float* F, const float* win, long long i, int n_fft, int hop, int T) {
  const long long m = i + n_fft / 2;
  long long g_lo = m < n_fft ? 0 : (m - n_fft) / hop + 1, g_hi = m / hop;
  if (g_hi > T - 1) g_hi = T - 1;
  float acc = 0.f, wss = 0.f;
  for (long long g = g_lo; g <= g_hi; ++g) {
    const int j = (int)(m - g * hop);
    acc += __ldcg(F + g * n_fft + j);
    wss = fmaf(win[j], win[j], wss);
  }
  return wss > 1.17549435e-38f ? acc / wss : acc;
}

// a counter in global memory as a grid barrier: legal because the grid is at most one CTA per SM and every CTA is
// resident (launch below); bounded (2 s) and trapping rather than hanging.  `epoch` = barriers passed so far.
__device__ __forceinline__ void au_grid_barrier(unsigned int* ctr, unsigned int epoch) {
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int expected = (epoch + 1) * gridDim.x;
    __threadfence();
    atomicAdd(ctr, 1u);
    unsigned int v = 0;
    unsigned long long t0, t1;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(ctr) : "memory");
      if (v >= expected) break;
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
      if (t1 - t0 > 2000000000ull) __trap();
      __nanosleep(32);
    }
    __threadfence();
  }
  __syncthreads();
}

// ---------------------------------------------------------------------------------------------- STFT
// one CTA per frame pair.  y (len) -> spec [T][n_fft/2 + 1] complex64, T = 1 + len / hop.  preemph != 0: the frames are
// taken from lfilter([1, -k], [1], y) (audio_tacotron.py:22-25) without materialising it.  mel_out != null: instead of
// the spectrum the kernel emits the normalised mel spectrogram [T][n_mels] (audio_tacotron.py:70-78).
struct MelArgs {
  const float* basis_t;      // [nb][n_mels]: librosa.filters.mel, transposed
  int n_mels;
  float min_level_db, ref_level_db, max_abs;
  float* mel;                // [T][n_mels]
};
__global__ void __launch_bounds__(AU_THREADS)
stft_kernel(const float* __restrict__ y, long long len, int n, int hop, int T, float preemph, float2* __restrict__ spec,
            MelArgs M) {
  extern __shared__ __align__(16) uint8_t au_raw[];
  const AuSmem s = au_carve(au_raw, n);
  au_tables(s, n);
  pdl_wait();
  pdl_trigger();
  const int nb = n / 2 + 1;
  for (int pair = blockIdx.x; 2 * pair < T; pair += gridDim.x) {
    const int f0 = 2 * pair, f1 = f0 + 1;
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
      float v[2];
#pragma unroll
      for (int q = 0; q < 2; ++q) {
        const int f = q ? f1 : f0;
        v[q] = 0.f;
        if (f < T) {
          const long long i = au_reflect((long long)f * hop + j, n, len);
          float x = __ldg(y + i);
          if (preemph != 0.f && i > 0) x = fmaf(-preemph, __ldg(y + i - 1), x);
          v[q] = x * s.win[j];
        }
      }
      s.a[j] = make_float2(v[0], v[1]);
    }
    __syncthreads();
    float2* Z = au_fft(s.a, s.b, s.tw, n);
    float* mag = reinterpret_cast<float*>(Z == s.a ? s.b : s.a);      // the free buffer: |X_a| then |X_b|, nb each
    for (int k = threadIdx.x; k < nb; k += blockDim.x) {
      float2 xa, xb;
      au_split(Z, n, k, &xa, &xb);
      if (M.mel) {
        mag[k] = sqrtf(xa.x * xa.x + xa.y * xa.y);
        mag[nb + k] = sqrtf(xb.x * xb.x + xb.y * xb.y);
      } else {
        spec[(long long)f0 * nb + k] = xa;
        if (f1 < T) spec[(long long)f1 * nb + k] = xb;
      }
    }
    __syncthreads();
    if (M.mel) {
      // S = 20 log10(max(min_level, basis . |X|)) - ref_level_db, normalised and clipped to [0, max_abs]
      const float min_level = exp10f(M.min_level_db * 0.05f);
      for (int e = threadIdx.x; e < 2 * M.n_mels; e += blockDim.x) {
        const int q = e / M.n_mels, m = e - q * M.n_mels, f = q ? f1 : f0;
        if (f >= T) continue;
        float acc = 0.f;
        for (int k = 0; k < nb; ++k) acc = fmaf(__ldg(M.basis_t + (long long)k * M.n_mels + m), mag[q * nb + k], acc);
        const float db = 20.f * log10f(fmaxf(min_level, acc)) - M.ref_level_db;
        const float nv = M.max_abs * ((db - M.min_level_db) / (-M.min_level_db));
        M.mel[(long long)f * M.n_mels + m] = fminf(fmaxf(nv, 0.f), M.max_abs);
      }
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------------------- ISTFT (two launches)
// frames: spec [T][nb] -> F [T][n_fft] = irfft(frame) * window
__global__ void __launch_bounds__(AU_THREADS)
istft_frames_kernel(const float2* __restrict__ spec, int n, int T, float* __restrict__ F) {
  extern __shared__ __align__(16) uint8_t au_raw[];
  const AuSmem s = au_carve(au_raw, n);
  au_tables(s, n);
  pdl_wait();
  pdl_trigger();
  const int nb = n / 2 + 1;
  const float inv_n = 1.0f / (float)n;
  for (int pair = blockIdx.x; 2 * pair < T; pair += gridDim.x) {
    const int f0 = 2 * pair, f1 = f0 + 1;
    for (int k = threadIdx.x; k < nb; k += blockDim.x) {
      const float2 ya = __ldg(spec + (long long)f0 * nb + k);
      const float2 yb = f1 < T ? __ldg(spec + (long long)f1 * nb + k) : make_float2(0.f, 0.f);
      au_pack_inverse(s.a, n, k, ya, yb);
    }
    __syncthreads();
    const float2* W = au_fft(s.a, s.b, s.tw, n);
    for (int j = threadIdx.x; j < n; j += blockDim.x) {
      const float2 w = W[j];
      F[(long long)f0 * n + j] = w.x * inv_n * s.win[j];
      if (f1 < T) F[(long long)f1 * n + j] = -w.y * inv_n * s.win[j];
    }
    __syncthreads();
  }
}
// overlap-add: F -> y (hop (T - 1) samples)
__global__ void __launch_bounds__(AU_THREADS)
istft_ola_kernel(const float* __restrict__ F, int n, int hop, int T, float* __restrict__ y) {
  extern __shared__ __align__(16) uint8_t au_raw[];
  float* win = reinterpret_cast<float*>(au_raw);
  for (int m = threadIdx.x; m < n; m += blockDim.x) win[m] = 0.5f - 0.5f * cospif(2.0f * (float)m / (float)n);
  __syncthreads();
  pdl_wait();
  pdl_trigger();
  const long long len = (long long)hop * (T - 1);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += (long long)gridDim.x * blockDim.x)
    y[i] = au_ola(F, win, i, n, hop, T);
}

// ---------------------------------------------------------------------------------------------- Griffin-Lim
// audio_tacotron.py:142-152, all iterations in one launch.  mag [T][nb] = the target magnitudes (S ** power), ang0
// [T][nb] = np.random.rand(*S.shape) of :146 (fractions of a turn).  F0 / F1 [T][n]: windowed inverse transforms of
// the even / odd steps.  Step 0 is  y = istft(S * exp(2 pi i ang0));  step s >= 1 is
// angles = exp(1j * angle(stft(y)));  y = istft(S * angles).   After the last step the signal is gathered into y.
__global__ void __launch_bounds__(AU_THREADS)
griffin_lim_kernel(const float* __restrict__ mag, const float* __restrict__ ang0, int n, int hop, int T, int iters,
                   float* F0, float* F1, float* __restrict__ y, unsigned int* barrier) {
  extern __shared__ __align__(16) uint8_t au_raw[];
  const AuSmem s = au_carve(au_raw, n);
  au_tables(s, n);
  pdl_wait();
  pdl_trigger();
  const int nb = n / 2 + 1;
  const float inv_n = 1.0f / (float)n;
  const long long len = (long long)hop * (T - 1);
  for (int step = 0; step <= iters; ++step) {
    const float* Fprev = (step & 1) ? F0 : F1;
    float* Fnext = (step & 1) ? F1 : F0;
    for (int pair = blockIdx.x; 2 * pair < T; pair += gridDim.x) {
      const int f0 = 2 * pair, f1 = f0 + 1;
      float2* src = s.a;
      if (step > 0) {
        // frames of the current signal, which exists only as the previous step's windowed inverse transforms
        for (int j = threadIdx.x; j < n; j += blockDim.x) {
          const float va = au_ola(Fprev, s.win, au_reflect((long long)f0 * hop + j, n, len), n, hop, T) * s.win[j];
          const float vb = f1 < T ? au_ola(Fprev, s.win, au_reflect((long long)f1 * hop + j, n, len), n, hop, T) * s.win[j]
                                  : 0.f;
          s.a[j] = make_float2(va, vb);
        }
        __syncthreads();
        float2* Z = au_fft(s.a, s.b, s.tw, n);
        float2* dst = Z == s.a ? s.b : s.a;
        for (int k = threadIdx.x; k < nb; k += blockDim.x) {
          float2 xa, xb;
          au_split(Z, n, k, &xa, &xb);
          // exp(1j * np.angle(X)): X / |X|, and 1 where X == 0 (np.angle(0) = 0)
          const float ma = xa.x * xa.x + xa.y * xa.y, mb = xb.x * xb.x + xb.y * xb.y;
          const float ra = ma > 0.f ? 1.0f / sqrtf(ma) : 0.f, rb = mb > 0.f ? 1.0f / sqrtf(mb) : 0.f;
          const float2 ua = ma > 0.f ? make_float2(xa.x * ra, xa.y * ra) : make_float2(1.f, 0.f);
          const float2 ub = mb > 0.f ? make_float2(xb.x * rb, xb.y * rb) : make_float2(1.f, 0.f);
          const float sa = __ldg(mag + (long long)f0 * nb + k);
          const float sb = f1 < T ? __ldg(mag + (long long)f1 * nb + k) : 0.f;
          au_pack_inverse(dst, n, k, make_float2(sa * ua.x, sa * ua.y), make_float2(sb * ub.x, sb * ub.y));
        }
        src = dst;
      } else {
        for (int k = threadIdx.x; k < nb; k += blockDim.x) {
          float sn, cs;
          sincospif(2.0f * __ldg(ang0 + (long long)f0 * nb + k), &sn, &cs);
          const float sa = __ldg(mag + (long long)f0 * nb + k);
          float2 yb = make_float2(0.f, 0.f);
          if (f1 < T) {
            float sn1, cs1;
            sincospif(2.0f * __ldg(ang0 + (long long)f1 * nb + k), &sn1, &cs1);
            const float sb = __ldg(mag + (long long)f1 * nb + k);
            yb = make_float2(sb * cs1, sb * sn1);
          }
          au_pack_inverse(s.a, n, k, make_float2(sa * cs, sa * sn), yb);
        }
      }
      __syncthreads();
      const float2* W = au_fft(src, src == s.a ? s.b : s.a, s.tw, n);
      for (int j = threadIdx.x; j < n; j += blockDim.x) {
        const float2 w = W[j];
        __stcg(Fnext + (long long)f0 * n + j, w.x * inv_n * s.win[j]);
        if (f1 < T) __stcg(Fnext + (long long)f1 * n + j, -w.y * inv_n * s.win[j]);
      }
      __syncthreads();
    }
    au_grid_barrier(barrier, (unsigned int)step);
  }
  const float* Flast = (iters & 1) ? F1 : F0;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < len; i += (long long)gridDim.x * blockDim.x)
    y[i] = au_ola(Flast, s.win, i, n, hop, T);
}

// ---------------------------------------------------------------------------------------------- mel -> linear magnitudes
// audio_tacotron.py:101-107 + the `** hparams.power` of :115: one CTA per frame.
//   amp[m] = 10 ** ((clip(mel[m][t], 0, max_abs) * -min_level_db / max_abs + min_level_db + ref_level_db) * 0.05)
//   mag[t][k] = max(1e-10, sum_m inv_t[m][k] * amp[m]) ** power            (inv_t = pinv(mel basis), transposed)
__global__ void __launch_bounds__(AU_THREADS)
mel_to_linear_kernel(const float* __restrict__ mel, int n_mels, int T, const float* __restrict__ inv_t, int nb,
                     float min_level_db, float ref_level_db, float max_abs, float power, float* __restrict__ mag) {
  extern __shared__ __align__(16) uint8_t au_raw[];
  float* amp = reinterpret_cast<float*>(au_raw);
  pdl_wait();
  pdl_trigger();
  const int t = blockIdx.x;
  for (int m = threadIdx.x; m < n_mels; m += blockDim.x) {
    const float d = fminf(fmaxf(__ldg(mel + (long long)m * T + t), 0.f), max_abs) * -min_level_db / max_abs + min_level_db;
    amp[m] = exp10f((d + ref_level_db) * 0.05f);
  }
  __syncthreads();
  for (int k = threadIdx.x; k < nb; k += blockDim.x) {
    float acc = 0.f;
    for (int m = 0; m < n_mels; ++m) acc = fmaf(__ldg(inv_t + (long long)m * nb + k), amp[m], acc);
    mag[(long long)t * nb + k] = powf(fmaxf(1e-10f, acc), power);
  }
}

// ---------------------------------------------------------------------------------------------- inverse pre-emphasis
// lfilter([1], [1, -k], x): y[i] = x[i] + k y[i-1] (audio_tacotron.py:27-30).  One CTA walks the signal in chunks of
// 1024 threads x 4 samples; inside a chunk the recurrence is a scan of affine maps c -> a c + b (a = k^4 per thread),
// the carry leaves through the last thread: exact for any k, no truncated history.
__global__ void __launch_bounds__(1024)
inv_preemphasis_kernel(const float* __restrict__ x, long long len, float k, float* __restrict__ y) {
  __shared__ float sA[32], sB[32];
  __shared__ float s_carry;
  pdl_wait();
  pdl_trigger();
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_carry = 0.f;
  __syncthreads();
  const float k2 = k * k, k4 = k2 * k2;
  for (long long base = 0; base < len; base += 4096) {
    const long long i0 = base + 4 * tid;
    float v[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) v[q] = i0 + q < len ? __ldg(x + i0 + q) : 0.f;
    // this thread's map with a zero carry-in
    float a = k4, b = fmaf(k, fmaf(k, fmaf(k, v[0], v[1]), v[2]), v[3]);
    // inclusive scan over the warp: (a1, b1) then (a2, b2) = (a1 a2, a2 b1 + b2)
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const float pa = __shfl_up_sync(0xffffffffu, a, o), pb = __shfl_up_sync(0xffffffffu, b, o);
      if (lane >= o) {
        b = fmaf(a, pb, b);
        a *= pa;
      }
    }
    if (lane == 31) {
      sA[warp] = a;
      sB[warp] = b;
    }
    __syncthreads();
    if (warp == 0) {
      float wa = sA[lane], wb = sB[lane];
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const float pa = __shfl_up_sync(0xffffffffu, wa, o), pb = __shfl_up_sync(0xffffffffu, wb, o);
        if (lane >= o) {
          wb = fmaf(wa, pb, wb);
          wa *= pa;
        }
      }
      sA[lane] = wa;
      sB[lane] = wb;
    }
    __syncthreads();
    const float carry = s_carry;
    // carry into this thread = (maps of all earlier threads of the chunk) applied to the chunk's carry
    float ea = __shfl_up_sync(0xffffffffu, a, 1), eb = __shfl_up_sync(0xffffffffu, b, 1);
    if (lane == 0) {
      ea = 1.f;
      eb = 0.f;
    }
    if (warp > 0) {                      // earlier warps first, then the earlier lanes of this warp
      const float wa = sA[warp - 1], wb = sB[warp - 1];
      eb = fmaf(ea, wb, eb);
      ea *= wa;
    }
    float c = fmaf(ea, carry, eb);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      c = fmaf(k, c, v[q]);
      if (i0 + q < len) y[i0 + q] = c;
    }
    __syncthreads();
    if (tid == 1023) s_carry = c;
    __syncthreads();
  }
}

}  // namespace nsg

using namespace nsg;

static int au_check_fft(int n_fft, int hop, const char* what) {
  NSG_REQUIRE(n_fft >= 16 && n_fft <= AU_MAX_FFT && (n_fft & (n_fft - 1)) == 0,
              "%s: n_fft %d must be a power of two in [16, %d]", what, n_fft, AU_MAX_FFT);
  NSG_REQUIRE(hop >= 1 && hop <= n_fft, "%s: hop %d must be in [1, n_fft]", what, hop);
  return NSG_OK;
}
template <typename K>
static int au_configure(K kernel, int n_fft, const char* what) {
  const size_t smem = au_smem_bytes(n_fft);
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return cuda_fail(e, what);
  }
  return NSG_OK;
}
static int au_grid(int n_frames) {
  const int pairs = (n_frames + 1) / 2;
  return pairs < num_sms() ? pairs : num_sms();
}

extern "C" int nsg_stft(const float* y, long long n_samples, int n_fft, int hop, float preemph, float* spec,
                        void* stream) {
  if (int rc = au_check_fft(n_fft, hop, "stft")) return rc;
  NSG_REQUIRE(y && spec, "stft: null pointer");
  NSG_REQUIRE(n_samples > n_fft / 2, "stft: reflect padding needs more than n_fft / 2 = %d samples (got %lld)", n_fft / 2,
              n_samples);
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(spec) & 7) == 0, "stft: spec must be 8-byte aligned");
  const int T = 1 + (int)(n_samples / hop);
  if (int rc = au_configure(stft_kernel, n_fft, "cudaFuncSetAttribute(stft_kernel)")) return rc;
  MelArgs M{};
  launch_pdl(stft_kernel, dim3(au_grid(T)), dim3(AU_THREADS), au_smem_bytes(n_fft), static_cast<cudaStream_t>(stream), y,
             n_samples, n_fft, hop, T, preemph, reinterpret_cast<float2*>(spec), M);
  NSG_CHECK_LAUNCH("stft_kernel");
  return NSG_OK;
}

extern "C" int nsg_melspectrogram(const float* wav, long long n_samples, int n_fft, int hop, float preemph,
                                  const float* basis_t, int n_mels, float min_level_db, float ref_level_db, float max_abs,
                                  float* mel, void* stream) {
  if (int rc = au_check_fft(n_fft, hop, "melspectrogram")) return rc;
  NSG_REQUIRE(wav && basis_t && mel, "melspectrogram: null pointer");
  NSG_REQUIRE(n_samples > n_fft / 2, "melspectrogram: needs more than n_fft / 2 = %d samples (got %lld)", n_fft / 2,
              n_samples);
  NSG_REQUIRE(n_mels >= 1 && n_mels <= 1024 && min_level_db < 0.f && max_abs > 0.f, "melspectrogram: bad mel arguments");
  const int T = 1 + (int)(n_samples / hop);
  if (int rc = au_configure(stft_kernel, n_fft, "cudaFuncSetAttribute(stft_kernel)")) return rc;
  MelArgs M{basis_t, n_mels, min_level_db, ref_level_db, max_abs, mel};
  launch_pdl(stft_kernel, dim3(au_grid(T)), dim3(AU_THREADS), au_smem_bytes(n_fft), static_cast<cudaStream_t>(stream), wav,
             n_samples, n_fft, hop, T, preemph, static_cast<float2*>(nullptr), M);
  NSG_CHECK_LAUNCH("stft_kernel(mel)");
  return NSG_OK;
}

extern "C" long long nsg_istft_workspace_bytes(int n_frames, int n_fft) {
  return (long long)n_frames * n_fft * (long long)sizeof(float);
}

extern "C" int nsg_istft(const float* spec, int n_frames, int n_fft, int hop, void* workspace, float* y, void* stream) {
  if (int rc = au_check_fft(n_fft, hop, "istft")) return rc;
  NSG_REQUIRE(spec && workspace && y, "istft: null pointer");
  NSG_REQUIRE(n_frames >= 2, "istft: at least two frames (got %d)", n_frames);
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(spec) & 7) == 0, "istft: spec must be 8-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (int rc = au_configure(istft_frames_kernel, n_fft, "cudaFuncSetAttribute(istft_frames_kernel)")) return rc;
  float* F = static_cast<float*>(workspace);
  launch_pdl(istft_frames_kernel, dim3(au_grid(n_frames)), dim3(AU_THREADS), au_smem_bytes(n_fft), s,
             reinterpret_cast<const float2*>(spec), n_fft, n_frames, F);
  NSG_CHECK_LAUNCH("istft_frames_kernel");
  const long long len = (long long)hop * (n_frames - 1);
  const int blocks = (int)((len + AU_THREADS - 1) / AU_THREADS < 4 * num_sms() ? (len + AU_THREADS - 1) / AU_THREADS
                                                                                 : 4 * num_sms());
  launch_pdl(istft_ola_kernel, dim3(blocks), dim3(AU_THREADS), (size_t)n_fft * sizeof(float), s,
             static_cast<const float*>(F), n_fft, hop, n_frames, y);
  NSG_CHECK_LAUNCH("istft_ola_kernel");
  return NSG_OK;
}

// ---------------------------------------------------------------------------------------------- ISTFT without allocation
// nsg_istft with its frame scratch (the windowed inverse transforms, n_frames * n_fft floats) in a file-scope device
// array instead of a caller-owned workspace: for callers that may not call cudaMalloc (code under CUDA graph capture,
// serving loops, harnesses that forbid it).  No counterpart in the reference.
// Every GPU that loads the library holds the array (4 MiB) until exit; every call on a device shares it, so
// calls that overlap on the GPU (different streams) race.  The array keeps the last call's contents; nothing in it has to
// start at zero: istft_frames_kernel overwrites every element that istft_ola_kernel reads in the same call.
constexpr long long AU_STATIC_SCRATCH_FLOATS = 1ll << 20;       // n_frames * n_fft <= 2^20 (e.g. 1024 frames of n_fft 1024)
static __device__ __align__(256) float g_istft_scratch[AU_STATIC_SCRATCH_FLOATS];

// In host code `g_istft_scratch` is a host-side stand-in, not a device address: the address a kernel may use comes
// from cudaGetSymbolAddress, once per device, with that device current.
static int au_static_scratch(float** out) {
  constexpr int kMaxDevices = 64;
  static float* ptr[kMaxDevices] = {};
  static std::mutex mu;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDevice(istft_static)");
  NSG_REQUIRE(dev >= 0 && dev < kMaxDevices, "istft_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(mu);
  if (!ptr[dev]) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_istft_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_istft_scratch)");
    NSG_REQUIRE(p != nullptr, "istft_static: cudaGetSymbolAddress returned a null address");
    ptr[dev] = p;
  }
  *out = ptr[dev];
  return NSG_OK;
}

extern "C" long long nsg_istft_static_max_floats(void) { return AU_STATIC_SCRATCH_FLOATS; }

extern "C" int nsg_istft_static(const float* spec, int n_frames, int n_fft, int hop, float* y, void* stream) {
  if (int rc = au_check_fft(n_fft, hop, "istft_static")) return rc;
  NSG_REQUIRE(spec && y, "istft_static: null pointer");
  NSG_REQUIRE(n_frames >= 2, "istft_static: at least two frames (got %d)", n_frames);
  NSG_REQUIRE((long long)n_frames * n_fft <= AU_STATIC_SCRATCH_FLOATS,
              "istft_static: n_frames * n_fft = %lld exceeds the built-in scratch of %lld floats; use nsg_istft with a "
              "caller-owned workspace", (long long)n_frames * n_fft, AU_STATIC_SCRATCH_FLOATS);
  float* F = nullptr;
  if (int rc = au_static_scratch(&F)) return rc;
  return nsg_istft(spec, n_frames, n_fft, hop, F, y, stream);
}

extern "C" long long nsg_griffin_lim_workspace_bytes(int n_frames, int n_fft) {
  return 2ll * n_frames * n_fft * (long long)sizeof(float) + 256;
}

extern "C" int nsg_griffin_lim(const float* mag, const float* angles0, int n_frames, int n_fft, int hop, int iters,
                               void* workspace, float* y, void* stream) {
  if (int rc = au_check_fft(n_fft, hop, "griffin_lim")) return rc;
  NSG_REQUIRE(mag && angles0 && workspace && y, "griffin_lim: null pointer");
  NSG_REQUIRE(n_frames >= 2 && iters >= 0, "griffin_lim: frames %d / iterations %d", n_frames, iters);
  NSG_REQUIRE((long long)hop * (n_frames - 1) > n_fft / 2,
              "griffin_lim: the signal (%lld samples) must be longer than n_fft / 2 (reflect padding)",
              (long long)hop * (n_frames - 1));
  NSG_REQUIRE((reinterpret_cast<uintptr_t>(workspace) & 255) == 0, "griffin_lim: workspace must be 256-byte aligned");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  if (int rc = au_configure(griffin_lim_kernel, n_fft, "cudaFuncSetAttribute(griffin_lim_kernel)")) return rc;
  // workspace: [barrier counter, 256 bytes][F0][F1]
  unsigned int* bar = static_cast<unsigned int*>(workspace);
  float* F0 = reinterpret_cast<float*>(static_cast<char*>(workspace) + 256);
  float* F1 = F0 + (size_t)n_frames * n_fft;
  cudaError_t e = cudaMemsetAsync(bar, 0, 256, s);
  if (e != cudaSuccess) return cuda_fail(e, "cudaMemsetAsync(griffin_lim barrier)");
  // grid barrier: every CTA must be resident -- at most one CTA per SM, greatest launch priority (common.cuh)
  launch_pdl_ex(true, griffin_lim_kernel, dim3(au_grid(n_frames)), dim3(AU_THREADS), au_smem_bytes(n_fft), s, mag, angles0,
                n_fft, hop, n_frames, iters, F0, F1, y, bar);
  NSG_CHECK_LAUNCH("griffin_lim_kernel");
  return NSG_OK;
}

extern "C" int nsg_mel_to_linear(const float* mel, int n_mels, int n_frames, const float* inv_basis_t, int n_bins,
                                 float min_level_db, float ref_level_db, float max_abs, float power, float* mag,
                                 void* stream) {
  NSG_REQUIRE(mel && inv_basis_t && mag, "mel_to_linear: null pointer");
  NSG_REQUIRE(n_mels >= 1 && n_mels <= 4096 && n_bins >= 1 && n_frames >= 0 && min_level_db < 0.f && max_abs > 0.f,
              "mel_to_linear: bad arguments");
  if (n_frames == 0) return NSG_OK;
  launch_pdl(mel_to_linear_kernel, dim3(n_frames), dim3(AU_THREADS), (size_t)n_mels * sizeof(float),
             static_cast<cudaStream_t>(stream), mel, n_mels, n_frames, inv_basis_t, n_bins, min_level_db, ref_level_db,
             max_abs, power, mag);
  NSG_CHECK_LAUNCH("mel_to_linear_kernel");
  return NSG_OK;
}

extern "C" int nsg_inv_preemphasis(const float* x, long long n_samples, float k, float* y, void* stream) {
  NSG_REQUIRE(x && y && n_samples >= 0, "inv_preemphasis: bad arguments");
  if (n_samples == 0) return NSG_OK;
  launch_pdl(inv_preemphasis_kernel, dim3(1), dim3(1024), 0, static_cast<cudaStream_t>(stream), x, n_samples, k, y);
  NSG_CHECK_LAUNCH("inv_preemphasis_kernel");
  return NSG_OK;
}
