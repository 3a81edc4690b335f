// C-ABI plumbing: error text, device check, dispatch between the CUDA-core and tcgen05 kernels.
#include <cstdlib>
#include <mutex>
#include <cstring>

#include "common.cuh"

namespace nsg {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int cuda_fail(cudaError_t e, const char* what) {
  set_error("%s: CUDA error %d (%s)", what, (int)e, cudaGetErrorString(e));
  return NSG_E_CUDA;
}

int num_sms() {
  static int cached = 0;
  if (cached == 0) {
    int dev = 0, n = 0;
    if (cudaGetDevice(&dev) == cudaSuccess &&
        cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) == cudaSuccess && n > 0)
      cached = n;
    else
      cached = 148;
  }
  return cached;
}

int conv_gemm_simt(const nsg_conv_geom* cg, const void* in, const float* w, const nsg_prologue* pro,
                   const nsg_epilogue* epi, void* out, cudaStream_t stream);
int conv_wgrad_simt(const nsg_conv_geom* cg, const void* p, const nsg_prologue* pro_p, const void* q,
                    const nsg_prologue* pro_q, float* dw, float* q_colsum, cudaStream_t stream);
long long conv_pack_bytes(const nsg_conv_geom* cg);
int conv_pack_weights(const nsg_conv_geom* cg, const float* w, void* packed, cudaStream_t stream);
int conv_gemm_tc(const nsg_conv_geom* cg, const void* in, const void* packed, const nsg_prologue* pro,
                 const nsg_epilogue* epi, void* out, int passes, cudaStream_t stream);
int conv_wgrad_tc(const nsg_conv_geom* cg, const void* p, const nsg_prologue* pro_p, const void* q,
                  const nsg_prologue* pro_q, float* dw, float* q_colsum, int passes, cudaStream_t stream);

int conv_bn_tma_fits(const nsg_conv_geom* cg);
int peer_allreduce_stats(double* sums, const nsg_peer_sync* p, cudaStream_t stream);
int conv_gemm_tma(const nsg_conv_geom* cg, const void* in, const void* packed, const nsg_epilogue* epi,
                  const nsg_bn_fuse* bnf, void* out,
                  cudaStream_t stream);
int debug_timeline(int enable, unsigned long long* host_out);
int conv_c1_tc(const nsg_conv_geom* cg, const void* in, const void* packed, const nsg_epilogue* epi,
               const nsg_bn_fuse* bnf, void* out, cudaStream_t stream);
int convT_c1_tc(const nsg_conv_geom* cg, const void* in, const void* packed, const float* bias, int tanh_out, void* out,
                cudaStream_t stream);
int wgrad_c1_tc(const nsg_conv_geom* cg, const void* p, const void* q, float* dw, float* q_colsum, cudaStream_t stream);
int conv_pack_many(int n, const float* const* w, void* const* dst, const int* ntaps, const long long* in_stride,
                   const long long* out_stride, cudaStream_t stream);
int conv_wgrad_tma(const nsg_conv_geom* cg, const void* p, const void* q, float* dwp, float* q_colsum, int sliced,
                   cudaStream_t stream);
int conv_wgrad_slices(const nsg_conv_geom* cg);
int wgrad_unpack_many(int n, const float* const* src, float* const* dw, const int* ntaps, const int* nslices,
                      const long long* a_stride, const long long* b_stride, cudaStream_t stream);

int greatest_priority() {
  static int cached = 1;
  if (cached == 1) {
    int least = 0, greatest = 0;
    cached = cudaDeviceGetStreamPriorityRange(&least, &greatest) == cudaSuccess ? greatest : 0;
  }
  return cached;
}

int g_pdl = getenv("NSG_PDL") ? atoi(getenv("NSG_PDL")) : 1;     // common.cuh: programmatic dependent launch

namespace {
bool g_notes_full = false;
struct StreamNote {
  cudaStream_t s;
  bool used, packed;
  unsigned long long since_pack;     // library launches on s after the last pack on s
};
StreamNote g_notes[16];
std::mutex g_notes_mu;
StreamNote* note_for(cudaStream_t s, bool create) {
  for (auto& n : g_notes)
    if (n.used && n.s == s) return &n;
  if (!create) return nullptr;
  for (auto& n : g_notes)
    if (!n.used) {
      n = StreamNote{s, true, false, 0};
      return &n;
    }
  return nullptr;       // table full: callers treat the stream as never settled
}
}  // namespace
void pdl_note_launch(cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_notes_mu);
  if (StreamNote* n = note_for(s, false)) ++n->since_pack;
}
void pdl_note_pack(cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_notes_mu);
  if (StreamNote* n = note_for(s, true)) {
    n->packed = true;
    n->since_pack = 0;
  } else {
    for (auto& m : g_notes) m.since_pack = 0, m.packed = true;     // cannot track: unsettle everything, forever safe
    g_notes_full = true;
  }
}
bool pdl_weights_settled(cudaStream_t s) {
  std::lock_guard<std::mutex> lk(g_notes_mu);
  if (g_notes_full) return false;
  StreamNote* n = note_for(s, false);
  return !n || !n->packed || n->since_pack >= 1;
}
}  // namespace nsg

// pinned, device-mapped words the kernels' bounded waits report into before trapping (one buffer per process)
static volatile unsigned int* g_hang_host = nullptr;
namespace nsg {
int conv_tc_hang_init(unsigned int*);
int conv_tma_hang_init(unsigned int*);
int vq_tc_hang_init(unsigned int*);
int fused_elem_hang_init(unsigned int*);
}
static int hang_init() {
  if (g_hang_host) return NSG_OK;
  unsigned int* h = nullptr;
  cudaError_t e = cudaHostAlloc(reinterpret_cast<void**>(&h), 8 * sizeof(unsigned int), cudaHostAllocMapped);
  if (e != cudaSuccess) return nsg::cuda_fail(e, "cudaHostAlloc(hang report)");
  memset(h, 0, 8 * sizeof(unsigned int));
  unsigned int* d = nullptr;
  e = cudaHostGetDevicePointer(reinterpret_cast<void**>(&d), h, 0);
  if (e != cudaSuccess) return nsg::cuda_fail(e, "cudaHostGetDevicePointer(hang report)");
  if (int rc = nsg::conv_tc_hang_init(d)) return rc;
  if (int rc = nsg::conv_tma_hang_init(d)) return rc;
  if (int rc = nsg::vq_tc_hang_init(d)) return rc;
  if (int rc = nsg::fused_elem_hang_init(d)) return rc;
  g_hang_host = h;
  return NSG_OK;
}

using namespace nsg;

extern "C" {

const char* nsg_version(void) { return "nsg_b200 0.1 (sm_100a)"; }
const char* nsg_last_error(void) { return g_err; }

int nsg_device_ok(char* name, int name_len) {
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    cuda_fail(e, "cudaGetDevice");
    return 0;
  }
  cudaDeviceProp prop;
  e = cudaGetDeviceProperties(&prop, dev);
  if (e != cudaSuccess) {
    cuda_fail(e, "cudaGetDeviceProperties");
    return 0;
  }
  if (name && name_len > 0) {
    strncpy(name, prop.name, (size_t)name_len - 1);
    name[name_len - 1] = 0;
  }
  if (prop.major != 10) {
    set_error("device '%s' is sm_%d%d; this library contains sm_100a code only", prop.name, prop.major, prop.minor);
    return 0;
  }
  if (hang_init() != NSG_OK) return 0;
  return 1;
}

/* Where a kernel gave up waiting (see tc_common.cuh "bounded waits"): out[0] != 0 when a report is present, then
 * out[1] = translation unit (1 conv_tc.cu, 2 conv_tma.cu, 3 vq_tc.cu, 4 fused_elem.cu), out[2] = source line, out[3] = block
 * (x | y << 20), out[4] = thread, out[5] = barrier parity / counter value, out[6] = grid size.  clear != 0 resets it. */
int nsg_hang_report(unsigned int* out, int clear) {
  if (!g_hang_host) {
    if (out) memset(out, 0, 8 * sizeof(unsigned int));
    return NSG_OK;
  }
  if (out) memcpy(out, (const void*)g_hang_host, 8 * sizeof(unsigned int));
  if (clear) memset((void*)g_hang_host, 0, 8 * sizeof(unsigned int));
  return NSG_OK;
}

static int check_geom(const nsg_conv_geom* g) {
  NSG_REQUIRE(g, "conv: null geometry");
  NSG_REQUIRE(g->batch > 0 && g->in_h > 0 && g->in_w > 0 && g->cin > 0 && g->out_h > 0 && g->out_w > 0 &&
                  g->cout > 0,
              "conv: non-positive dimension");
  NSG_REQUIRE(g->gather == 0 || g->gather == 1, "conv: gather must be 0 or 1");
  return NSG_OK;
}

int nsg_conv_gemm(const nsg_conv_geom* g, const void* in, const float* w, const nsg_prologue* pro,
                  const nsg_epilogue* epi, void* out, int impl, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(in && w && out, "conv_gemm: null pointer");
  NSG_REQUIRE(impl == 0, "conv_gemm: impl must be 0 (the tensor-core kernel is nsg_conv_gemm_tc)");
  return conv_gemm_simt(g, in, w, pro, epi, out, (cudaStream_t)stream);
}

long long nsg_conv_pack_bytes(const nsg_conv_geom* g) { return g ? conv_pack_bytes(g) : -1; }

int nsg_conv_pack_weights(const nsg_conv_geom* g, const float* w, void* packed, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(w && packed, "conv_pack_weights: null pointer");
  return conv_pack_weights(g, w, packed, (cudaStream_t)stream);
}

int nsg_conv_gemm_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_prologue* pro,
                     const nsg_epilogue* epi, void* out, int passes, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(in && packed_w && out, "conv_gemm_tc: null pointer");
  return conv_gemm_tc(g, in, packed_w, pro, epi, out, passes, (cudaStream_t)stream);
}

int nsg_conv_wgrad_tc(const nsg_conv_geom* g, const void* p, const nsg_prologue* pro_p, const void* q,
                      const nsg_prologue* pro_q, float* dw, float* q_colsum, int passes, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(p && q && dw, "conv_wgrad_tc: null pointer");
  return conv_wgrad_tc(g, p, pro_p, q, pro_q, dw, q_colsum, passes, (cudaStream_t)stream);
}

int nsg_conv_wgrad(const nsg_conv_geom* g, const void* p, const nsg_prologue* pro_p, const void* q,
                   const nsg_prologue* pro_q, float* dw, float* q_colsum, int impl, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(p && q && dw, "conv_wgrad: null pointer");
  NSG_REQUIRE(impl == 0, "conv_wgrad: impl must be 0 (the tensor-core kernel is nsg_conv_wgrad_tc)");
  return conv_wgrad_simt(g, p, pro_p, q, pro_q, dw, q_colsum, (cudaStream_t)stream);
}

int nsg_conv_gemm_tma(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi, void* out,
                      void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(in && packed_w && out, "conv_gemm_tma: null pointer");
  return conv_gemm_tma(g, in, packed_w, epi, nullptr, out, (cudaStream_t)stream);
}

int nsg_conv_bn_tma(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi,
                    const nsg_bn_fuse* bn, void* out, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(in && packed_w, "conv_bn_tma: null pointer");
  return conv_gemm_tma(g, in, packed_w, epi, bn, out, (cudaStream_t)stream);
}

int nsg_peer_allreduce_stats(double* sums, const nsg_peer_sync* peer, void* stream) {
  return peer_allreduce_stats(sums, peer, (cudaStream_t)stream);
}

int nsg_conv_bn_tma_fits(const nsg_conv_geom* g) {
  if (check_geom(g)) return 0;
  return conv_bn_tma_fits(g);
}

int nsg_c1_conv_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi, void* out,
                   void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(in && packed_w && out, "c1_conv_tc: null pointer");
  return conv_c1_tc(g, in, packed_w, epi, nullptr, out, (cudaStream_t)stream);
}

int nsg_c1_conv_bn_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const nsg_epilogue* epi,
                      const nsg_bn_fuse* bn, void* out, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(in && packed_w && out, "c1_conv_bn_tc: null pointer");
  return conv_c1_tc(g, in, packed_w, epi, bn, out, (cudaStream_t)stream);
}

int nsg_c1_convT_tc(const nsg_conv_geom* g, const void* in, const void* packed_w, const float* bias, int tanh_out,
                    void* out, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(in && packed_w && out, "c1_convT_tc: null pointer");
  return convT_c1_tc(g, in, packed_w, bias, tanh_out, out, (cudaStream_t)stream);
}

int nsg_c1_wgrad_tc(const nsg_conv_geom* g, const void* p, const void* q, float* dw, float* q_colsum, void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(p && q && dw, "c1_wgrad_tc: null pointer");
  return wgrad_c1_tc(g, p, q, dw, q_colsum, (cudaStream_t)stream);
}

int nsg_debug_timeline(int enable, unsigned long long* stamps32) { return debug_timeline(enable, stamps32); }

int nsg_conv_pack_many(int n, const float* const* w, void* const* packed, const int* ntaps,
                       const long long* w_cin_stride, const long long* w_cout_stride, void* stream) {
  NSG_REQUIRE(n == 0 || (w && packed && ntaps && w_cin_stride && w_cout_stride), "conv_pack_many: null pointer");
  return conv_pack_many(n, w, packed, ntaps, w_cin_stride, w_cout_stride, (cudaStream_t)stream);
}

int nsg_conv_wgrad_tma(const nsg_conv_geom* g, const void* p, const void* q, float* dw_packed, float* q_colsum,
                       void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(p && q && dw_packed, "conv_wgrad_tma: null pointer");
  return conv_wgrad_tma(g, p, q, dw_packed, q_colsum, 0, (cudaStream_t)stream);
}

int nsg_pdl(int enable) {
  const int old = g_pdl;
  if (enable == 0 || enable == 1) g_pdl = enable;
  return old;
}

int nsg_conv_wgrad_slices(const nsg_conv_geom* g) {
  if (check_geom(g)) return 0;
  return conv_wgrad_slices(g);
}

int nsg_conv_wgrad_tma_sliced(const nsg_conv_geom* g, const void* p, const void* q, float* dw_slices, float* q_colsum,
                              void* stream) {
  int rc = check_geom(g);
  if (rc) return rc;
  NSG_REQUIRE(p && q && dw_slices, "conv_wgrad_tma_sliced: null pointer");
  return conv_wgrad_tma(g, p, q, dw_slices, q_colsum, 1, (cudaStream_t)stream);
}

int nsg_wgrad_unpack_many(int n, const float* const* dw_packed, float* const* dw, const int* ntaps, const int* nslices,
                          const long long* w_cin_stride, const long long* w_cout_stride, void* stream) {
  NSG_REQUIRE(n == 0 || (dw_packed && dw && ntaps && w_cin_stride && w_cout_stride), "wgrad_unpack_many: null pointer");
  return wgrad_unpack_many(n, dw_packed, dw, ntaps, nslices, w_cin_stride, w_cout_stride, (cudaStream_t)stream);
}

}  // extern "C"
