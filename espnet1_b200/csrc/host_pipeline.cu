// Host-buffer entry point (bench.py `e2e`): logits arrive in HOST memory, the
// loss (and optionally the gradient) go back to HOST memory.  Utterance chunks
// are copied H2D on a copy stream while the previous chunk runs loss_fwd_grad
// on the compute stream; gradient chunks return on a third stream.
#include <mutex>
#include <vector>

#include "ctc_ws.cuh"

namespace b200ctc {

struct Arena {
  void* logits = nullptr; size_t logits_cap = 0;
  void* grad = nullptr; size_t grad_cap = 0;
  void* ws = nullptr; size_t ws_cap = 0;
  void* small = nullptr; size_t small_cap = 0;  // targets, offsets, lengths, nll
  cudaStream_t s_copy = nullptr, s_comp = nullptr, s_back = nullptr;
  std::vector<cudaEvent_t> ev;
  int device = -1;
};
static Arena g_arena;
static std::mutex g_mu;

static bool grow(void** p, size_t* cap, size_t need) {
  if (*cap >= need) return true;
  if (*p) cudaFree(*p);
  *p = nullptr; *cap = 0;
  if (cudaMalloc(p, need) != cudaSuccess) return false;
  *cap = need;
  return true;
}

static void release_locked() {
  Arena& a = g_arena;
  if (a.logits) cudaFree(a.logits);
  if (a.grad) cudaFree(a.grad);
  if (a.ws) cudaFree(a.ws);
  if (a.small) cudaFree(a.small);
  for (auto e : a.ev) cudaEventDestroy(e);
  if (a.s_copy) cudaStreamDestroy(a.s_copy);
  if (a.s_comp) cudaStreamDestroy(a.s_comp);
  if (a.s_back) cudaStreamDestroy(a.s_back);
  a = Arena();
}

#define CU_TRY(expr)                                                   \
  do {                                                                 \
    cudaError_t e_ = (expr);                                           \
    if (e_ != cudaSuccess) {                                           \
      set_error("host pipeline: %s: %s", #expr, cudaGetErrorString(e_)); \
      return B200CTC_ERR_DEVICE;                                       \
    }                                                                  \
  } while (0)

}  // namespace b200ctc

using namespace b200ctc;

extern "C" void b200ctc_host_pipeline_release(void) {
  std::lock_guard<std::mutex> lk(g_mu);
  release_locked();
}

extern "C" int b200ctc_loss_fwd_bwd_host(const void* logits_host, int dtype,
                                         const int64_t* targets_host,
                                         const int64_t* in_lens_host,
                                         const int64_t* tgt_lens_host, int B, int T, int V,
                                         int blank, int zero_infinity, float gscale,
                                         float* nll_host, void* grad_host, void** grad_dev_out,
                                         int chunk_utts) {
  B2_REQUIRE(logits_host && in_lens_host && tgt_lens_host && nll_host, "loss_host: null pointer");
  B2_REQUIRE(B > 0 && T > 0 && V > 0, "loss_host: bad shape");
  B2_REQUIRE(dtype >= 0 && dtype <= 2, "loss_host: unknown dtype %d", dtype);
  std::lock_guard<std::mutex> lk(g_mu);
  Arena& a = g_arena;
  int dev = 0;
  CU_TRY(cudaGetDevice(&dev));
  if (a.device != dev) { release_locked(); a.device = dev; }
  if (!a.s_copy) {
    CU_TRY(cudaStreamCreateWithFlags(&a.s_copy, cudaStreamNonBlocking));
    CU_TRY(cudaStreamCreateWithFlags(&a.s_comp, cudaStreamNonBlocking));
    CU_TRY(cudaStreamCreateWithFlags(&a.s_back, cudaStreamNonBlocking));
  }
  if (chunk_utts <= 0) chunk_utts = (B + 7) / 8;
  if (chunk_utts > B) chunk_utts = B;
  const int nchunk = (B + chunk_utts - 1) / chunk_utts;
  // host-side label bookkeeping
  std::vector<int64_t> offs(B + 1, 0);
  int64_t lmax = 0;
  for (int b = 0; b < B; ++b) {
    const int64_t l = tgt_lens_host[b] < 0 ? 0 : tgt_lens_host[b];
    offs[b + 1] = offs[b] + l;
    if (l > lmax) lmax = l;
  }
  B2_REQUIRE(targets_host || offs[B] == 0, "loss_host: null targets");
  B2_REQUIRE(2 * lmax + 1 <= 8192, "loss_host: target too long");
  const size_t esz = dtype_size(dtype);
  const size_t utt_bytes = (size_t)T * V * esz;
  const LossLayout wl = make_loss_layout(chunk_utts, T, V, (int)lmax);
  const size_t n_tgt = (size_t)offs[B];
  const size_t small_need = 8 * (n_tgt + (size_t)(B + 1) + 2 * (size_t)B) + 4 * (size_t)B + 256;
  if (!grow(&a.logits, &a.logits_cap, utt_bytes * B) || !grow(&a.grad, &a.grad_cap, utt_bytes * B) ||
      !grow(&a.ws, &a.ws_cap, wl.total) || !grow(&a.small, &a.small_cap, small_need)) {
    set_error("loss_host: device allocation failed");
    return B200CTC_ERR_DEVICE;
  }
  while ((int)a.ev.size() < 2 * nchunk + 1) {
    cudaEvent_t e;
    CU_TRY(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
    a.ev.push_back(e);
  }
  int64_t* d_tgt = static_cast<int64_t*>(a.small);
  int64_t* d_off = d_tgt + n_tgt;
  int64_t* d_il = d_off + (B + 1);
  int64_t* d_tl = d_il + B;
  float* d_nll = reinterpret_cast<float*>(d_tl + B);
  if (n_tgt) CU_TRY(cudaMemcpyAsync(d_tgt, targets_host, 8 * n_tgt, cudaMemcpyHostToDevice, a.s_comp));
  CU_TRY(cudaMemcpyAsync(d_off, offs.data(), 8 * (size_t)(B + 1), cudaMemcpyHostToDevice, a.s_comp));
  CU_TRY(cudaMemcpyAsync(d_il, in_lens_host, 8 * (size_t)B, cudaMemcpyHostToDevice, a.s_comp));
  CU_TRY(cudaMemcpyAsync(d_tl, tgt_lens_host, 8 * (size_t)B, cudaMemcpyHostToDevice, a.s_comp));
  // offsets are consumed relative to each chunk's first utterance: the kernels
  // index targets[offs[b]] with the GLOBAL offsets, so pass the global arrays
  // shifted by the chunk start.
  int rc = B200CTC_OK;
  for (int c = 0; c < nchunk && rc == B200CTC_OK; ++c) {
    const int b0 = c * chunk_utts;
    const int nb = (b0 + chunk_utts <= B) ? chunk_utts : B - b0;
    char* dl = static_cast<char*>(a.logits) + utt_bytes * b0;
    char* dg = static_cast<char*>(a.grad) + utt_bytes * b0;
    CU_TRY(cudaMemcpyAsync(dl, static_cast<const char*>(logits_host) + utt_bytes * b0,
                           utt_bytes * nb, cudaMemcpyHostToDevice, a.s_copy));
    CU_TRY(cudaEventRecord(a.ev[2 * c], a.s_copy));
    CU_TRY(cudaStreamWaitEvent(a.s_comp, a.ev[2 * c], 0));
    rc = b200ctc_loss_fwd_grad(dl, dtype, d_tgt, d_off + b0, d_il + b0, d_tl + b0, nb, T, V,
                               (int)lmax, blank, zero_infinity, gscale, d_nll + b0, nullptr, 0.f,
                               dg, a.ws, a.ws_cap, a.s_comp);
    if (rc) break;
    if (grad_host) {
      CU_TRY(cudaEventRecord(a.ev[2 * c + 1], a.s_comp));
      CU_TRY(cudaStreamWaitEvent(a.s_back, a.ev[2 * c + 1], 0));
      CU_TRY(cudaMemcpyAsync(static_cast<char*>(grad_host) + utt_bytes * b0, dg, utt_bytes * nb,
                             cudaMemcpyDeviceToHost, a.s_back));
    }
  }
  if (rc == B200CTC_OK)
    CU_TRY(cudaMemcpyAsync(nll_host, d_nll, 4 * (size_t)B, cudaMemcpyDeviceToHost, a.s_comp));
  CU_TRY(cudaStreamSynchronize(a.s_comp));
  CU_TRY(cudaStreamSynchronize(a.s_back));
  CU_TRY(cudaStreamSynchronize(a.s_copy));
  if (grad_dev_out) *grad_dev_out = a.grad;
  return rc;
}
