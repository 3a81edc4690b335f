/* nsg_b200 -- C ABI of the B200-native VQ-VAE hot path.
 *
 * Drop-in boundary for dendisuhubdy/neural_sound_generation.  The reference has no FFI layer of
 * its own (its hot path is Python calling ATen); each entry point below names the reference call
 * site it replaces (paths relative to /root/reference).  INTEGRATION.md shows the ctypes binding a
 * maintainer adds on the reference side.
 *
 * Conventions
 *  - extern "C", plain pointers and sizes only.  Every pointer is a DEVICE pointer owned by the
 *    caller (the PyTorch caching allocator in the shipped host layer); the library allocates nothing.
 *  - every call takes the cudaStream_t to launch on (as void*), never synchronises, never touches
 *    the host heap after the first call: all entry points are CUDA-graph capturable.
 *  - return value: 0 = OK, negative = error (NSG_E_*); text via nsg_last_error() (thread local).
 *  - there is NO CPU path: a missing device, a non-sm_100 device or a null pointer is an error.
 *  - activations are NHWC ("pixel-major": (N, H, W, C) with C contiguous).  `dtype` selects the
 *    activation storage type: NSG_F32 or NSG_BF16.  Parameters, statistics, losses and optimiser
 *    state are always fp32 (sums in fp64).
 */
#ifndef NSG_B200_H_
#define NSG_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define NSG_OK 0
#define NSG_E_INVALID (-1)     /* bad argument / unsupported shape */
#define NSG_E_CUDA (-2)        /* CUDA runtime error, see nsg_last_error() */
#define NSG_E_UNSUPPORTED (-3) /* valid request this build cannot serve (e.g. tensor-core path shape) */

#define NSG_F32 0
#define NSG_BF16 1

/* operand transform applied while an activation tensor is being read ("prologue") */
#define NSG_PRO_NONE 0
#define NSG_PRO_RELU 1        /* max(x, 0)                           ResBlock leading ReLU, models.py:149 */
#define NSG_PRO_AFFINE_RELU 2 /* max(scale[c]*x + shift[c], 0)       BatchNorm2d + ReLU, models.py:151-152,166-167,180-181 */
#define NSG_PRO_AFFINE 3      /* scale[c]*x + shift[c]               BatchNorm2d alone */

/* gradient mask applied in a dgrad epilogue */
#define NSG_MASK_NONE 0
#define NSG_MASK_POS 1    /* keep where mask_src > 0                         (ReLU backward) */
#define NSG_MASK_AFFINE 2 /* keep where mscale[c]*mask_src + mshift[c] > 0   (BN+ReLU backward) */

/* per-channel statistics accumulated by an epilogue into double stats[2*C] (caller zeroes them) */
#define NSG_STATS_NONE 0
#define NSG_STATS_MOMENTS 1 /* sum v, sum v^2            -> BatchNorm forward statistics           */
#define NSG_STATS_BNBWD 2   /* sum g, sum g*xhat         -> BatchNorm backward reductions           */

const char* nsg_version(void);
const char* nsg_last_error(void);
/* 1 when the current device is sm_100 (B200); fills name (<= 127 chars) when not null.  Also the library's one-time
 * set-up (the pinned report words below): call it once per process and device before the first launch. */
int nsg_device_ok(char* name, int name_len);
/* Every wait inside the tensor-core kernels (mbarrier, grid barrier, peer flags) is bounded: after ~2 s without
 * progress the waiting thread records where it was stuck in pinned host memory and traps, so a lost pipeline stage
 * surfaces as a CUDA launch failure with a location instead of a hung device.  out: 8 words -- out[0] != 0 when a
 * report is present; out[1] translation unit (1 conv_tc.cu, 2 conv_tma.cu, 3 vq_tc.cu, 4 fused_elem.cu), out[2] source line, out[3]
 * block (x | y << 20), out[4] thread, out[5] barrier parity / counter value, out[6] grid size.  clear != 0 resets. */
int nsg_hang_report(unsigned int* out, int clear);

/* ---------------------------------------------------------------- convolution (models.py:150,153,165,168,179,182)
 * One implicit-GEMM-over-taps entry point serves nn.Conv2d forward, nn.ConvTranspose2d forward and both
 * data gradients (ATen convolution_backward in the reference path):
 *     out[n, oh, ow, co] = bias[co] + sum_{kh,kw,ci} pro(in[n, ih, iw, ci]) * w[kh*k+kw + ci*w_cin_stride + co*w_cout_stride]
 *   gather 0 ("conv"):        ih = oh*stride - pad + kh                      (Conv2d fwd, ConvTranspose2d dgrad)
 *   gather 1 ("transposed"):  ih = (oh + pad - kh)/stride when divisible     (ConvTranspose2d fwd, Conv2d dgrad)
 * out-of-range taps contribute 0 (zero padding is applied AFTER the prologue, as in the reference where the
 * activation precedes the padded convolution).  `w` is the reference's parameter tensor untouched (OIHW for
 * Conv2d, (Cin,Cout,KH,KW) for ConvTranspose2d); the two strides say which of its axes the GEMM contracts.
 * cin / cout are the channel counts of `in` / `out`.
 */
typedef struct {
  int batch, in_h, in_w, cin;   /* tensor being read  */
  int out_h, out_w, cout;       /* tensor being written (for wgrad: the un-gathered operand Q) */
  int ksize, stride, pad;
  int gather;                   /* 0 conv-like, 1 transposed-like */
  long long w_cin_stride;       /* elements between consecutive GEMM-input channels in w  */
  long long w_cout_stride;      /* elements between consecutive GEMM-output channels in w */
  int dtype;                    /* NSG_F32 / NSG_BF16: storage type of in/out/add/mask/xhat tensors */
} nsg_conv_geom;

typedef struct {
  int mode;                 /* NSG_PRO_* */
  const float* scale;       /* [channels] when affine */
  const float* shift;
} nsg_prologue;

typedef struct {
  const float* bias;        /* [cout] or null */
  const void* add_src;      /* same shape as out, added before masking, or null */
  int mask_mode;            /* NSG_MASK_* */
  const void* mask_src;     /* same shape as out */
  const float* mscale;      /* [cout] for NSG_MASK_AFFINE */
  const float* mshift;
  int stats_mode;           /* NSG_STATS_* */
  double* stats;            /* [2*cout], accumulated (+=) */
  const void* xhat_src;     /* NSG_STATS_BNBWD: pre-BN tensor h, same shape as out */
  const float* mean;        /* [cout] */
  const float* invstd;      /* [cout] */
  int tanh_out;             /* output activation: 0 none, 1 tanh (models.py:183), 2 ReLU (nsg_conv_gemm_tma only:
                             * the in-place ReLU that opens the next ResBlock, models.py:149) */
} nsg_epilogue;

/* BatchNorm fused into the producing convolution (nsg_conv_bn_tma; replaces the conv -> nn.BatchNorm2d -> ReLU
 * [-> residual add] chains of models.py:150-158 and their autograd mirror by ONE launch each):
 *   NSG_BNF_EVAL   running statistics folded into the output affine of the ordinary (streaming) kernel:
 *                  out = epilogue(gamma*(conv + bias - running_mean)/sqrt(running_var + eps) + beta), where epilogue =
 *                  the nsg_epilogue add_src / mask / ReLU as usual.
 *   NSG_BNF_TRAIN  batch statistics: every tile's fp32 accumulator stays in TMEM, the CTAs of the persistent grid sum
 *                  their moments into epi->stats (NSG_STATS_MOMENTS, zeroed by the caller), meet at `barrier`, then
 *                  out2 = [relu]([res +] gamma*(h - mean)*invstd + beta) with h = conv + bias; `out` (nullable) receives h
 *                  (the backward pass reads it); running statistics / num_batches_tracked updated as torch does;
 *                  coef (nullable, [4][64]) receives mean | invstd | scale | shift.
 *   NSG_BNF_BWD    g = epilogue(conv) (the data gradient: + add_src, ReLU mask), epi->stats (NSG_STATS_BNBWD with
 *                  xhat_src = h, mean, invstd of the BatchNorm being differentiated) summed over the grid, barrier, then
 *                  out2 = gamma*invstd*(g - sum_g/count - xhat*sum_gx/count); `out` (nullable) receives g;
 *                  dgamma += sum_gx, dbeta += sum_g (nullable).
 * TRAIN / BWD need every tile of a CTA to keep its accumulator: at most 8 tiles of 128 pixels per CTA of the
 * SM-count-sized grid (nsg_conv_bn_tma_fits(g) == 1), else NSG_E_UNSUPPORTED -- run nsg_conv_gemm_tma + nsg_bn_act64 /
 * nsg_bn_bwd64.  `barrier` is ONE unsigned counter per launch, zero before the launch.  The grid barrier relies on all
 * CTAs being resident (grid <= SM count, one CTA per SM): do not run two such launches concurrently on one device
 * from different streams. */
/* Cross-GPU statistics for synchronised BatchNorm (one process per GPU of ONE node): every rank owns an exchange
 * area of NSG_PEER_AREA_BYTES in peer-mapped memory (torch.distributed._symmetric_memory / cuMemCreate + cuMemMap),
 * zero-initialised, whose address AS MAPPED ON THIS GPU is areas[rank].  A launch that carries an nsg_peer_sync sums
 * its 128 fp64 statistics over the ranks through those areas (direct NVLink stores + flags, ranks added in rank order:
 * identical bits everywhere) -- no collective launch.  `slot` selects one of NSG_PEER_SLOTS exchange slots; a slot may
 * be used once between two points at which all ranks synchronise (the step's gradient all-reduce), by the same launch
 * on every rank.  A rank that never arrives traps the waiting kernels after ~40 s instead of hanging the device. */
#define NSG_PEER_MAX 8
#define NSG_PEER_SLOTS 32
#define NSG_PEER_AREA_BYTES (NSG_PEER_SLOTS * NSG_PEER_MAX * 128 * 8 + NSG_PEER_SLOTS * NSG_PEER_MAX * 4)
typedef struct {
  int world, rank, slot;
  void* areas[NSG_PEER_MAX];
} nsg_peer_sync;
/* sums[0..127] (fp64, device) := sum over ranks; one small launch (the stand-alone form of the exchange) */
int nsg_peer_allreduce_stats(double* sums, const nsg_peer_sync* peer, void* stream);

#define NSG_BNF_NONE 0
#define NSG_BNF_TRAIN 1
#define NSG_BNF_EVAL 2
#define NSG_BNF_BWD 3
typedef struct {
  int mode;
  int relu_out;             /* TRAIN: ReLU on out2 */
  const float* gamma;       /* [64] */
  const float* beta;        /* [64] (TRAIN / EVAL) */
  float eps, momentum;
  double count;             /* elements per channel behind the statistics (pixels, x world size for synchronised BN) */
  float* running_mean;      /* [64]; TRAIN: updated (nullable), EVAL: read */
  float* running_var;
  long long* num_batches_tracked;
  float* coef;              /* TRAIN: [4][64] published coefficients (nullable) */
  const void* res;          /* TRAIN: residual, same shape as the output (nullable) */
  unsigned int* barrier;    /* TRAIN / BWD */
  float* dgamma;            /* BWD (nullable) */
  float* dbeta;
  void* out2;               /* TRAIN / BWD: second output, 32-byte aligned */
  const nsg_peer_sync* peer; /* TRAIN / BWD, nullable: sum the statistics over the GPUs of the node inside the barrier
                              * (`count` must then be the GLOBAL element count and `barrier` TWO zeroed counters) */
} nsg_bn_fuse;

/* CUDA-core fp32-accumulate kernel, any shape.  `impl` must be 0 (kept for ABI stability). */
int nsg_conv_gemm(const nsg_conv_geom* g, const void* in, const float* w, const nsg_prologue* pro,
                  const nsg_epilogue* epi, void* out, int impl, void* stream);

/* weight gradient:  dw[t + a*w_cin_stride + b*w_cout_stride] += sum_m proP(P[gather0(m, t), a]) * proQ(Q[m, b])
 * with m over Q's pixels and the conv-like gather into P (geometry fields: in_* = P, out_* = Q).
 *   Conv2d:           P = x,  Q = dy  (a = ci, b = co)      ConvTranspose2d:  P = dy, Q = x  (a = co, b = ci)
 * dw is ACCUMULATED with float atomics; q_colsum (nullable, [cout]) += sum_m proQ(Q[m, b]) (= dbias for Conv2d). */
int nsg_conv_wgrad(const nsg_conv_geom* g, const void* p, const nsg_prologue* pro_p, const void* q,
                   const nsg_prologue* pro_q, float* dw, float* q_colsum, int impl, void* stream);

/* tcgen05 / TMEM tensor-core kernels for the 64 -> 64 channel layers (cin == cout == 64; other shapes return
 * NSG_E_UNSUPPORTED).  Same contract as nsg_conv_gemm / nsg_conv_wgrad, except that the forward/dgrad kernel
 * reads the weights from a caller-owned packed bf16 image (hi + lo halves per tap, UMMA K-major SWIZZLE_128B)
 * written by nsg_conv_pack_weights for the SAME geometry (the two strides select the GEMM orientation, so a
 * layer has one image for its forward and one for its data gradient).  `passes` selects operand precision:
 *   1  bf16 x bf16                      (bf16 mode: what torch.autocast(bfloat16) would feed cuDNN)
 *   2  activation x (w_hi + w_lo)       (bf16 activations, fp32-faithful weights)
 *   3  (a_hi + a_lo) x (w_hi + w_lo) without the lo*lo term: fp32 activations to ~2^-16 relative
 * Accumulation is fp32 in TMEM in every case.  nsg_conv_pack_bytes < 0 = shape not served. */
long long nsg_conv_pack_bytes(const nsg_conv_geom* g);
int nsg_conv_pack_weights(const nsg_conv_geom* g, const float* w, void* packed, void* stream);
int nsg_conv_gemm_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_prologue* pro,
                     const nsg_epilogue* epi, void* out, int passes, void* stream);
int nsg_conv_wgrad_tc(const nsg_conv_geom* g, const void* p, const nsg_prologue* pro_p, const void* q,
                      const nsg_prologue* pro_q, float* dw, float* q_colsum, int passes, void* stream);

/* TMA-fed persistent tcgen05 kernels (bf16 tensors, cin == cout == 64; the production path of the bf16 train step).
 * They read OPERAND-READY activations -- the BatchNorm/ReLU that precedes a convolution in models.py:148-183 has
 * already been applied by nsg_bn_act64 / an epilogue mask -- so there is no prologue argument: the tensor core
 * consumes exactly what the TMA unit wrote (zero padding = the unit's out-of-bounds fill).
 *   nsg_conv_pack_many     bf16 K-major SWIZZLE_128B weight images (k*k * 8192 bytes each, 128-byte aligned) of up
 *                          to 24 (layer, orientation) pairs in one launch; strides as in nsg_conv_geom
 *   nsg_conv_gemm_tma      same contract as nsg_conv_gemm_tc with passes = 1 and no prologue; `out` 32-byte aligned
 *                          (rows leave as whole 32-byte sectors)
 *   nsg_conv_wgrad_tma     dw_packed[t][a][b] (fp32, k*k*64*64, 16-byte aligned) += sum_m P[gather0(m,t),a]*Q[m,b];
 *                          q_colsum (nullable) as in nsg_conv_wgrad.  The split-K partials of all CTAs meet in the
 *                          one image through vector atomics.
 *   nsg_conv_wgrad_tma_sliced  the same sum left as S = nsg_conv_wgrad_slices(g) partial images
 *                          dw_slices[s][t][a][b] (each OVERWRITTEN, plain stores, no atomics; q_colsum still +=):
 *                          the form the training step uses -- faster, and the gradient no longer depends on CTA order
 *   nsg_conv_wgrad_slices  S for that geometry on the current device (0: geometry not served)
 *   nsg_wgrad_unpack_many  dw[t + a*w_cin_stride + b*w_cout_stride] += sum_{s < nslices[i]} dw_packed[s][t][a][b] for
 *                          up to 24 layers; nslices null = one image each */
/* The single-channel ends of the network (models.py:165 Conv2d(1 -> 64, 4, 2, 1), models.py:182
 * ConvTranspose2d(64 -> 1, 4, 2, 1)) on the tensor cores, bf16 tensors, geometry k4 / s2 / p1 only (else
 * NSG_E_UNSUPPORTED; the CUDA-core nsg_conv_gemm / nsg_conv_wgrad serve every other shape).  packed_w are 8 KB images
 * written by nsg_conv_pack_many with ntaps = -1 (rows = the 64 channels, K = the 16 taps: nsg_c1_conv_tc) or
 * ntaps = -2 (rows = the 16 taps, K = the 64 channels: nsg_c1_convT_tc); w_cin_stride / w_cout_stride there are the
 * element strides of w along K and along rows.
 *   nsg_c1_conv_tc   out[m, 0..63] = epilogue(bias + sum_t in1[gather0(m, t)] * w[t, 0..63])  (encoder.0 fwd, decoder.6 dgrad)
 *   nsg_c1_convT_tc  out1[m] = [tanh](bias + sum_{valid t} sum_c in[gather1(m, t), c] * w[c, t])   (decoder.6 fwd)
 *   nsg_c1_wgrad_tc  dw[t + b*w_cout_stride] += sum_m p1[gather0(m, t)] * q[m, b]; q_colsum[b] += sum_m q[m, b] */
int nsg_c1_conv_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi, void* out,
                   void* stream);
/* nsg_c1_conv_tc with an eval-mode BatchNorm (bn->mode == NSG_BNF_EVAL) folded into its output affine (models.py:165-167
 * in eval mode: one launch instead of two) */
int nsg_c1_conv_bn_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi,
                      const nsg_bn_fuse* bn, void* out, void* stream);
int nsg_c1_convT_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const float* bias, int tanh_out,
                    void* out, void* stream);
int nsg_c1_wgrad_tc(const nsg_conv_geom* g, const void* p, const void* q, float* dw, float* q_colsum, void* stream);

/* development aid: enable != 0 makes CTA 0 of every following nsg_conv_gemm_tma launch write clock64 stamps of its
 * pipeline milestones; stamps32 (nullable, HOST pointer to 32 values) receives the last launch's stamps (synchronises). */
int nsg_debug_timeline(int enable, unsigned long long* stamps32);
int nsg_conv_pack_many(int n, const float* const* w, void* const* packed, const int* ntaps,
                       const long long* w_cin_stride, const long long* w_cout_stride, void* stream);
int nsg_conv_gemm_tma(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi, void* out,
                      void* stream);
int nsg_conv_bn_tma(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi,
                    const nsg_bn_fuse* bn, void* out, void* stream);
int nsg_conv_bn_tma_fits(const nsg_conv_geom* g);
int nsg_conv_wgrad_tma(const nsg_conv_geom* g, const void* p, const void* q, float* dw_packed, float* q_colsum,
                       void* stream);
/* Programmatic dependent launch (every kernel of this library): 1 (default) lets a kernel become resident and run its
 * data-independent prologue while its predecessor in the stream drains; every kernel waits for ALL earlier work in
 * the stream before its first dependent access, so results are identical.  One thing is read above that wait: the
 * packed weight images of nsg_conv_gemm_tma, and only when the library knows they are settled -- it tracks
 * nsg_conv_pack_many per stream and requires at least one of its own launches between the pack and the consumer.
 * A caller that fills packed_w by other means (its own kernel) in the launch right before nsg_conv_gemm_tma on the
 * same stream must either put any other launch / event / copy in between or call nsg_pdl(0).
 * Returns the previous setting; any value other than 0 / 1 only reads.  Environment NSG_PDL=0 sets the default. */
int nsg_pdl(int enable);
int nsg_conv_wgrad_slices(const nsg_conv_geom* g);
int nsg_conv_wgrad_tma_sliced(const nsg_conv_geom* g, const void* p, const void* q, float* dw_slices, float* q_colsum,
                              void* stream);
int nsg_wgrad_unpack_many(int n, const float* const* dw_packed, float* const* dw, const int* ntaps, const int* nslices,
                          const long long* w_cin_stride, const long long* w_cout_stride, void* stream);

/* Streaming companions for bf16 tensors with 64 channels (fused_elem.cu), `pixels` = N*H*W:
 *   nsg_bn_act64   out = [relu]([res +] BN(h)); train (stats != null: [sum | sumsq] over `count` elements, as
 *                  accumulated by a NSG_STATS_MOMENTS epilogue): batch statistics, running statistics updated as
 *                  torch does; eval (stats == null): running statistics.  mean/invstd/scale/shift (nullable) are
 *                  published for the backward pass.                         nn.BatchNorm2d + ReLU, models.py:151-158
 *   nsg_bn_bwd64   dh = gamma*invstd*(g - sum_g/M - xhat*sum_gx/M) from stats = [sum g | sum g*xhat]
 *                  (NSG_STATS_BNBWD); dgamma += sum_gx; dbeta += sum_g; dbias_prev (nullable) += column sums of dh
 *   nsg_relu64     out = max(x, 0)
 *   nsg_vq_bwd_stats64  dz = g_st + coef*(z - zq)  (vq_st backward clone + commitment term, vector_quantization.py:50-52,
 *                  train.py:133) with the NSG_STATS_BNBWD reductions of the BatchNorm that produced z fused (g_st, zq, stats nullable) */
int nsg_bn_act64(const void* h, const void* res, void* out, const double* stats, double count, const float* gamma,
                 const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                 long long* num_batches_tracked, float* mean, float* invstd, float* scale, float* shift, int relu_out,
                 long long pixels, void* stream);
int nsg_bn_bwd64(const void* g, const void* h, void* dh, const double* stats, double count, const float* mean,
                 const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* dbias_prev,
                 long long pixels, void* stream);
/* The same two kernels with the cross-GPU statistic exchange INSIDE them (synchronised BatchNorm, one process per GPU):
 * CTA 0 swaps `stats` with the other ranks over peer memory (nsg_peer_sync above) while the tensor loads of every CTA
 * are in flight, writes the global sums back into `stats` and opens `gate` (one zeroed unsigned counter per launch);
 * `count` must be the GLOBAL element count.  peer == null or world <= 1: exactly nsg_bn_act64 / nsg_bn_bwd64. */
int nsg_bn_act64_sync(const void* h, const void* res, void* out, double* stats, double count, const float* gamma,
                      const float* beta, float eps, float momentum, float* running_mean, float* running_var,
                      long long* num_batches_tracked, float* mean, float* invstd, float* scale, float* shift,
                      int relu_out, long long pixels, const nsg_peer_sync* peer, unsigned int* gate, void* stream);
int nsg_bn_bwd64_sync(const void* g, const void* h, void* dh, double* stats, double count, const float* mean,
                      const float* invstd, const float* gamma, float* dgamma, float* dbeta, float* dbias_prev,
                      long long pixels, const nsg_peer_sync* peer, unsigned int* gate, void* stream);
int nsg_relu64(const void* x, void* out, long long n, void* stream);
int nsg_vq_bwd_stats64(const void* g_st, const void* z, const void* zq, float coef, void* dz, const void* h,
                       const float* mean, const float* invstd, double* stats, long long pixels, void* stream);

/* ---------------------------------------------------------------- batch norm (nn.BatchNorm2d, models.py:151,154,166,180) */
/* stats (sum, sumsq over `count` elements per channel) -> mean, invstd, scale = gamma*invstd,
 * shift = beta - mean*scale; running stats updated with momentum (unbiased variance), as torch does. */
int nsg_bn_finalize(const double* stats, int channels, double count, const float* gamma, const float* beta,
                    float eps, float momentum, float* running_mean, float* running_var,
                    long long* num_batches_tracked, float* mean, float* invstd, float* scale, float* shift,
                    void* stream);
/* eval mode: scale/shift from running statistics */
int nsg_bn_eval_coeffs(int channels, const float* gamma, const float* beta, const float* running_mean,
                       const float* running_var, float eps, float* scale, float* shift, void* stream);
/* y = [relu](scale*h + shift) [+ relu(res)]: ResBlock output  relu(x) + BN(h)  (models.py:158) */
int nsg_bn_apply(const void* h, const float* scale, const float* shift, const void* res, int res_relu,
                 int relu_out, void* y, long long pixels, int channels, int dtype, void* stream);
/* sums for BN backward: stats[c] += sum g, stats[C+c] += sum g*xhat   (g optionally ReLU-masked by bn(h) > 0) */
int nsg_bn_bwd_reduce(const void* g, const void* h, const float* mean, const float* invstd, const float* scale,
                      const float* shift, int relu_mask, double* stats, long long pixels, int channels,
                      int dtype, void* stream);
/* dh = gamma*invstd*(g - sum_g/M - xhat*sum_gx/M); dgamma += sum_gx; dbeta += sum_g;  dbias_prev += sum dh */
int nsg_bn_bwd_apply(const void* g, const void* h, const float* mean, const float* invstd, const float* gamma,
                     const float* scale, const float* shift, int relu_mask, const double* stats, double count,
                     void* dh, float* dgamma, float* dbeta, long long pixels, int channels, int dtype,
                     void* stream);
/* dx = (g1 + g2) * (x > 0)   (ResBlock input gradient: skip + conv path through the in-place ReLU) */
int nsg_relu_bwd_add(const void* g1, const void* g2, const void* x, void* dx, long long n, int dtype,
                     void* stream);
/* out[c] += sum over pixels of g[p, c]   (bias gradients) */
int nsg_colsum(const void* g, float* out, long long pixels, int channels, int dtype, void* stream);

/* ---------------------------------------------------------------- vector quantisation (vector_quantization.py) */
/* vq / vq_st forward (vector_quantization.py:6-23, 34-44) fused with the two VQ loss terms' common value
 * (train.py:131,133).  z (N,D) activations [dtype], codebook (K,D) fp32.
 *   indices : int64 [N]  argmin_k |z-e_k|^2, lowest index on ties            (required)
 *   codes   : [N,D] [dtype] = codebook[indices]                              (nullable)
 *   sqerr   : double[1] += sum (z - e)^2                                     (nullable)
 * mode 0 = exact fp32 CUDA-core search; 1 = tcgen05 split-bf16 tensor-core search with exact fp32 re-scoring of
 * the candidates (D in {64, 128, 256}; K <= 8192, 7168 at D = 256; fp32 z whose hi + lo tile images no longer fit in
 * shared memory -- D = 256, or D = 128 with K > 7168 -- runs the exact CUDA-core search inside the same call).  mode 1 needs a caller-owned scratch buffer `ws` of
 * nsg_vq_workspace_bytes(k, d, 1) bytes (negative = shape not served by that mode); mode 0 needs none. */
long long nsg_vq_workspace_bytes(int k, int d, int mode);
int nsg_vq_forward(const void* z, const float* codebook, long long n, int k, int d, int dtype, int mode,
                   long long* indices, void* codes, double* sqerr, void* ws, long long ws_bytes, void* stream);
/* index_add_ / codebook gradient (vector_quantization.py:58-61; autograd of models.py:137-138) and the EMA
 * extension's statistics: table[idx[n], :] += alpha * (rows[n,:] - (sub ? sub[n,:] : 0)); counts[idx[n]] += 1. */
int nsg_scatter_add(const void* rows, const void* sub, const long long* indices, long long n, int k, int d,
                    float alpha, float* table, float* counts, int dtype, void* stream);
/* development aid: 1 (default) = K <= 512 uses the persistent kernel with the codebook resident in shared memory,
 * 0 = always the per-tile streaming kernel.  Returns the previous setting; other values only query. */
int nsg_vq_resident(int enable);
/* development aid: 0 (default) = block-aggregated scatter kernel where the shape is served (n >= 4096, d % 4 == 0,
 * d <= 256): time independent of how skewed the assignment is; 1 = always the per-element vector-atomic kernel (faster
 * for uniform assignments, up to 12x slower for skewed ones).  Returns the previous setting; other values only query. */
int nsg_scatter_mode(int mode);
/* EMA codebook update (extension -- NOT in the reference; van den Oord et al. 2017, appendix A.1):
 *   ema_count <- decay*ema_count + (1-decay)*counts;   ema_sum <- decay*ema_sum + (1-decay)*sums;
 *   codebook[k,:] = ema_sum[k,:] / ((ema_count[k] + eps) / (sum ema_count + k*eps) * sum ema_count)
 * counts [k] / sums [k,d] are one step's nsg_scatter_add outputs (table = sums with alpha 1, counts); every array is
 * fp32; scratch is one caller-owned double. d % 4 == 0. */
int nsg_ema_update(const float* counts, const float* sums, float* ema_count, float* ema_sum, float* codebook, int k,
                   int d, float decay, float eps, double* scratch, void* stream);
/* out[n,:] = table[indices[n],:]: nn.Embedding lookup / index_select (models.py:137-138, 194); out-of-range -> 0 */
int nsg_gather_rows(const float* table, const long long* indices, long long n, int k, int d, void* out, int dtype,
                    void* stream);
/* encoder-side gradient at the bottleneck: dz = g_st + coef*(z - zq)   (vq_st backward clone + commitment term);
 * g_st and zq are nullable (treated as 0). */
int nsg_vq_backward_inputs(const void* g_st, const void* z, const void* zq, float coef, void* dz, long long n,
                           int dtype, void* stream);

/* ---------------------------------------------------------------- losses (train.py:118-134) */
/* recon = mean over (B,H,W) of (pad(x_tilde) - c)^2 where x_tilde has width wt <= w and is zero padded;
 * loss_sum[0] += sum of squares;  dpre (nullable) = d loss_recons / d (pre-tanh), i.e.
 * 2*(x_tilde - c)/(B*H*W) * (1 - x_tilde^2), shape (B,H,wt). x_tilde, c, dpre are fp32. */
int nsg_recon_loss(const float* x_tilde, const float* c, int batch, int h, int w, int wt, double* loss_sum,
                   float* dpre, void* stream);

/* The same for a bf16 x_tilde (the TMA engine's decoder output), with the conversions around it folded in: dpre
 * (nullable) is written as bf16, x_tilde_f32 (nullable, (B,H,wt)) receives x_tilde widened to fp32, dpre_sum (nullable,
 * one float) += the sum of the written dpre values = the gradient of the output layer's bias (models.py:182). */
int nsg_recon_loss_bf16(const void* x_tilde, const float* c, int batch, int h, int w, int wt, double* loss_sum,
                        void* dpre, float* x_tilde_f32, float* dpre_sum, void* stream);
/* losses3 = [sums[0]*recon_scale, sums[1]*vq_scale, sums[1]*vq_scale]: loss_recons, loss_vq, loss_commit
 * (train.py:129-134; the last two are the same number before beta) from the accumulated squared-error sums */
int nsg_loss_finalize(const double* sums, double recon_scale, double vq_scale, double* losses3, void* stream);

/* out = g * (1 - y^2): gradient through the decoder's final nn.Tanh (models.py:183), fp32 */
int nsg_tanh_bwd(const float* g, const float* y, float* out, long long n, void* stream);

/* ---------------------------------------------------------------- input pipeline (dataloader.py:410-415) */
/* Device half of collate_fn for the mel conditioning features: `frames` holds the kept frames of the batch's utterances
 * back to back, time-major ((sum of lengths) x n_mels fp32, cropped on the host as dataloader.py:356-362 prescribes);
 * offsets[b] .. offsets[b+1] (batch + 1 ints, in frames) delimit item b.  Writes the zero-padded, transposed batch
 * (batch, n_mels, t_out) as fp32 (out_f32, nullable) and as bf16 (out_bf16, nullable; = the (B, n_mels, T, 1) NHWC
 * tensor the bf16 encoder reads).  Frames beyond t_out are dropped, items shorter than t_out are zero-padded. */
int nsg_mel_collate(const float* frames, const int* offsets, int batch, int n_mels, int t_out, float* out_f32,
                    void* out_bf16, void* stream);

/* ---------------------------------------------------------------- other model families (models.py:64-118, 218-341; row N4) */
/* Stride-1 convolution with a RECTANGULAR kernel, any channel counts, any zero padding, fp32 NHWC: the layers of the
 * reference's VAE (5 x 5, models.py:74,85) and GatedPixelCNN (vertical (k/2+1) x k and horizontal 1 x (k/2+1) stacks,
 * models.py:241-253) that the square <= 4 x 4 entry points above do not cover.
 *   gather 0: out[n,oh,ow,co] = bias[co] + sum in[n, oh - pad_h + kh, ow - pad_w + kw, ci] * w[kh*kw_n + kw + ci*w_cin_stride + co*w_cout_stride]
 *   gather 1: the same with ih = oh + pad_h - kh, iw = ow + pad_w - kw  (ConvTranspose2d forward / Conv2d data gradient)
 * out_h / out_w may be SMALLER than the full result: only the leading rows / columns are computed (the crop of
 * models.py:264,268).  add_src (nullable, same shape as out) is added before the activation; act: 0 none, 1 tanh,
 * 2 ReLU; stats (nullable, act == 0): double[2*cout] += per-channel sum and sum of squares (BatchNorm moments). */
typedef struct {
  int batch, in_h, in_w, cin;
  int out_h, out_w, cout;
  int kh, kw, pad_h, pad_w;
  int gather;
  long long w_cin_stride, w_cout_stride;
} nsg_rect_geom;
int nsg_conv_rect(const nsg_rect_geom* geom, const float* in, const float* w, const float* bias, const float* add_src,
                  int act, double* stats, float* out, void* stream);
/* dw[tap + a*w_cin_stride + b*w_cout_stride] += sum over q's pixels of p[gather-0 tap pixel, a] * q[pixel, b]; geom:
 * in_* / cin describe p, out_* / cout describe q (swap the operands for a transposed convolution, as nsg_conv_wgrad);
 * q_colsum (nullable) += per-channel sums of q (the bias gradient). */
int nsg_conv_rect_wgrad(const nsg_rect_geom* geom, const float* p, const float* q, float* dw, float* q_colsum,
                        void* stream);

/* out[m][c] = tanh(a) * sigmoid(b), (a | b) = the two channel halves of x[m][2C] + add2[m][2C] + cond[item(m)][2C]
 * (GatedActivation, models.py:218-225, with the class-conditional bias of :262-270); add2 / cond nullable;
 * item(m) = m / pixels_per_item. */
int nsg_gated_act(const float* x, const float* add2, const float* cond, long long pixels, int channels,
                  long long pixels_per_item, float* out, void* stream);

/* Reparameterisation + KL term of the convolutional VAE (models.py:103-110): stats[m][2Z] = (mu | logvar) channel halves;
 * z[m][c] = mu + exp(logvar / 2) * eps[m][c];  kl_sum[0] += sum over m, c of KL(N(mu, std) || N(0, 1))
 * = 0.5 (std^2 + mu^2 - 1 - logvar).  Backward: dstats = (dz + kl_scale mu | dz eps std / 2 + kl_scale (std^2 - 1) / 2). */
int nsg_vae_reparam_kl(const float* stats, const float* eps, long long pixels, int z_channels, float* z, double* kl_sum,
                       void* stream);
int nsg_vae_reparam_kl_bwd(const float* stats, const float* eps, const float* dz, float kl_scale, long long pixels,
                           int z_channels, float* dstats, void* stream);

/* sum[0] += sum (a - b)^2; grad (nullable) = grad_scale * 2 (a - b) (* (1 - a^2) when through_tanh: a = tanh(pre),
 * grad is then with respect to pre)   -- loss.py:22-24 mse_loss(size_average=False) and its backward */
int nsg_sqerr_sum(const float* a, const float* b, long long n, double* sum, float grad_scale, int through_tanh,
                  float* grad, void* stream);

/* One categorical draw per row by inverse CDF: p = softmax(logits[row]) over `classes` contiguous floats at
 * logits + row * row_stride + offset; out[row * out_stride] = the first class whose cumulative probability exceeds
 * u[row] (u in [0, 1); the last class when rounding leaves the total below u)  -- GatedPixelCNN.generate,
 * models.py:334-339, with the uniform draws supplied by the caller. */
int nsg_softmax_sample(const float* logits, long long rows, int classes, long long row_stride, long long offset,
                       const float* u, long long* out, long long out_stride, void* stream);

/* ---------------------------------------------------------------- mel <-> audio (audio_tacotron.py) */
/* All fp32, frame-major device buffers: a spectrum is [n_frames][n_fft/2 + 1] (complex64 = interleaved re, im), a mel
 * spectrogram handed OUT is [n_frames][n_mels]; a mel spectrogram handed IN (nsg_mel_to_linear) keeps the layout the
 * reference passes around, (n_mels, n_frames) row-major (main.py:152-181).  n_fft: a power of two in [16, 4096]. */

/* librosa.stft(y, n_fft, hop) as audio_tacotron.py:154-158 calls it: centred frames over the reflect-padded signal,
 * periodic Hann window, n_frames = 1 + n_samples / hop.  preemph != 0: the transform of lfilter([1, -preemph], [1], y)
 * (audio_tacotron.py:22-25) without materialising it.  spec: [n_frames][n_fft/2 + 1][2]. */
int nsg_stft(const float* y, long long n_samples, int n_fft, int hop, float preemph, float* spec, void* stream);

/* audio_tacotron.py:70-78 melspectrogram with the signal-normalisation branch the reference's hparams select
 * (allow_clipping_in_normalization, not symmetric): pre-emphasis, STFT, |.|, basis_t ([n_fft/2 + 1][n_mels] =
 * librosa.filters.mel transposed), 20 log10(max(10^(min_level_db/20), .)) - ref_level_db, scaled into [0, max_abs] and
 * clipped.  mel: [n_frames][n_mels]. */
int nsg_melspectrogram(const float* wav, long long n_samples, int n_fft, int hop, float preemph, const float* basis_t,
                       int n_mels, float min_level_db, float ref_level_db, float max_abs, float* mel, void* stream);

/* librosa.istft(D, hop) as audio_tacotron.py:160-161 calls it: irfft, synthesis window, overlap-add, division by the
 * window sum-of-squares envelope (where > tiny(float32)), n_fft/2 trimmed at both ends -> hop * (n_frames - 1) samples.
 * workspace: nsg_istft_workspace_bytes(n_frames, n_fft) bytes, caller-owned. */
long long nsg_istft_workspace_bytes(int n_frames, int n_fft);
int nsg_istft(const float* spec, int n_frames, int n_fft, int hop, void* workspace, float* y, void* stream);

/* nsg_istft without any allocation: the frame scratch lives in a file-scope device array inside the library.  (No
 * counterpart in the reference: added for callers that may not call cudaMalloc
 * -- code under CUDA graph capture, serving loops, harnesses that forbid it.)  Limit: n_frames * n_fft <=
 * nsg_istft_static_max_floats() = 2^20 (4 MiB held by every GPU that loads the library, until exit); larger inputs
 * are rejected with NSG_E_INVALID and a message naming nsg_istft.  The entry point calls no cudaMalloc / cudaFree or
 * allocating library routine.  Every call on a device shares the array: calls that overlap on the GPU (different
 * streams) race -- serialise them.  Nothing in the scratch has to start at zero (each call overwrites what it reads).
 * Same results as nsg_istft, bit for bit. */
long long nsg_istft_static_max_floats(void);
int nsg_istft_static(const float* spec, int n_frames, int n_fft, int hop, float* y, void* stream);

/* audio_tacotron.py:142-152 _griffin_lim: y = istft(mag * exp(2 pi i angles0)); `iters` times { angles =
 * exp(1j * angle(stft(y))); y = istft(mag * angles) }.  mag: [n_frames][n_fft/2 + 1] target magnitudes (the reference
 * passes S ** power); angles0: same shape, the np.random.rand(*S.shape) of :146 (fractions of a turn) -- explicit so
 * that a caller (and the parity tests) own the randomness.  ONE launch: a persistent grid of <= one CTA per SM with a
 * grid barrier per iteration (launched with the device's greatest priority, see "Launch policy" in INTEGRATION.md).
 * workspace: nsg_griffin_lim_workspace_bytes() bytes, 256-byte aligned, caller-owned.  y: hop * (n_frames - 1). */
long long nsg_griffin_lim_workspace_bytes(int n_frames, int n_fft);
int nsg_griffin_lim(const float* mag, const float* angles0, int n_frames, int n_fft, int hop, int iters, void* workspace,
                    float* y, void* stream);

/* audio_tacotron.py:101-107 + the `** hparams.power` of :115: mag[t][k] = max(1e-10, sum_m inv_basis_t[m][k] *
 * 10^((clip(mel[m][t], 0, max_abs) * -min_level_db / max_abs + min_level_db + ref_level_db) / 20)) ^ power.
 * mel: (n_mels, n_frames) row-major; inv_basis_t: [n_mels][n_bins] = pinv(mel basis) transposed. */
int nsg_mel_to_linear(const float* mel, int n_mels, int n_frames, const float* inv_basis_t, int n_bins,
                      float min_level_db, float ref_level_db, float max_abs, float power, float* mag, void* stream);

/* scipy.signal.lfilter([1], [1, -k], x) (audio_tacotron.py:27-30): y[i] = x[i] + k y[i-1]; x and y must not overlap */
int nsg_inv_preemphasis(const float* x, long long n_samples, float k, float* y, void* stream);

/* ---------------------------------------------------------------- optimiser (torch.optim.Adam, main.py:124) */
/* one Adam step over a flat fp32 buffer.  step_count is a device scalar incremented AFTER use (graphs replay
 * correctly): by the last block of the same launch when `ticket` (one zero-initialised unsigned counter owned by this
 * optimiser, zero again afterwards) is given, else by a second one-thread launch.  grad_scale multiplies grads (DP
 * averaging).  zero_grads != 0 clears each gradient after it has been used (the next step's optimizer.zero_grad(),
 * main.py/train.py:116, for free). */
int nsg_adam_step(float* params, float* grads, float* exp_avg, float* exp_avg_sq, long long n, float lr,
                  float beta1, float beta2, float eps, float grad_scale, long long* step_count, int zero_grads,
                  unsigned int* ticket, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NSG_B200_H_ */
