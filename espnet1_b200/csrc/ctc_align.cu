// Forced alignment (espnet/nets/pytorch_backend/ctc.py:163-226): Viterbi over the
// blank-interleaved label lattice with back-pointers, then back-tracking.
// The reference keeps the lattice in fp64 (numpy default) with a finite
// log-zero of -1e11 and lets index -1 wrap to the last state for s = 0; both
// are reproduced so that the integer alignment is identical.
#include <mutex>

#include "common.cuh"

namespace b200ctc {

__global__ void __launch_bounds__(1024) align_kernel(const float* __restrict__ lp,
                                                     const int64_t* __restrict__ target, int T,
                                                     int V, int L, int blank,
                                                     int64_t* __restrict__ out,
                                                     int* __restrict__ back) {
  extern __shared__ double dl[];  // [2][S]
  const int S = 2 * L + 1;
  const int NT = blockDim.x;
  const double logzero = -100000000000.0;
  auto label = [&](int s) -> int { return (s & 1) ? (int)target[(s - 1) >> 1] : blank; };
  for (int s = threadIdx.x; s < S; s += NT) {
    dl[s] = (s < 2) ? (double)lp[label(s)] : logzero;  // delta[0,s] = lp[0][y_int[s]], s<2
  }
  __syncthreads();
  for (int t = 1; t < T; ++t) {
    const double* cur = dl + ((t - 1) & 1) * S;
    double* nxt = dl + (t & 1) * S;
    for (int s = threadIdx.x; s < S; s += NT) {
      const int ls = label(s);
      const int s1 = s >= 1 ? s - 1 : S - 1;  // numpy's [-1] wrap for s = 0
      double best = cur[s];
      int arg = s;
      if (cur[s1] > best) { best = cur[s1]; arg = s1; }
      if (!(ls == blank || s < 2 || ls == label(s - 2))) {
        if (cur[s - 2] > best) { best = cur[s - 2]; arg = s - 2; }
      }
      nxt[s] = best + (double)lp[(size_t)t * V + ls];
      back[(size_t)t * S + s] = arg;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const double* fin = dl + ((T - 1) & 1) * S;
    int s = (S >= 2 && fin[S - 2] > fin[S - 1]) ? S - 2 : S - 1;
    for (int t = T - 1; t >= 0; --t) {
      out[t] = label(s);
      if (t > 0) s = back[(size_t)t * S + s];
    }
  }
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" size_t b200ctc_align_workspace_bytes(int T, int tgt_len) {
  if (T <= 0 || tgt_len < 0) return 0;
  return (size_t)T * (2 * (size_t)tgt_len + 1) * sizeof(int) + 64;
}

extern "C" int b200ctc_forced_align(const float* lp, const int64_t* target, int T, int V,
                                    int tgt_len, int blank, int64_t* out_ids, void* ws,
                                    size_t ws_bytes, b200ctc_stream_t stream) {
  B2_REQUIRE(lp && target && out_ids && ws, "forced_align: null pointer");
  B2_REQUIRE(T > 0 && V > 0 && tgt_len >= 1 && blank >= 0 && blank < V, "forced_align: bad shape");
  if (ws_bytes < b200ctc_align_workspace_bytes(T, tgt_len)) {
    set_error("forced_align: workspace too small");
    return B200CTC_ERR_WORKSPACE;
  }
  const int S = 2 * tgt_len + 1;
  const size_t smem = 2 * (size_t)S * sizeof(double);
  B2_REQUIRE(smem <= 200 * 1024, "forced_align: target too long (%d)", tgt_len);
  int nt = (S + 31) / 32 * 32;
  if (nt > 1024) nt = 1024;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  align_kernel<<<1, nt, smem, static_cast<cudaStream_t>(stream)>>>(lp, target, T, V, tgt_len, blank,
                                                                  out_ids, static_cast<int*>(ws));
  count_launch();
  return check_launch("forced_align");
}

// ---- allocation-free variant ------------------------------------------------------------
// b200ctc_forced_align_static: the same kernel with its back-pointer scratch in a file-scope
// device array instead of a caller-provided workspace, for callers that cannot allocate (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No
// counterpart in the reference.
//  * the array costs B200CTC_ALIGN_STATIC_MAX_BYTES on every GPU that loads the library,
//    until the process exits;
//  * in HOST code `g_align_scratch` is only a stand-in for the symbol, not a device address:
//    the address is resolved once per device (with that device current) by
//    cudaGetSymbolAddress below and that pointer is what the kernel receives;
//  * all calls on a device share the array and it keeps the previous call's contents.  Nothing
//    has to start at zero: the kernel writes back[t][s] for every t in [1,T), s in [0,S) before
//    thread 0 back-tracks through exactly those words, so stale contents are never read;
//  * calls that overlap on one GPU (two streams, two host threads) RACE on the array: the
//    caller must serialise them (one stream, or events).
static __device__ __align__(256) unsigned char g_align_scratch[B200CTC_ALIGN_STATIC_MAX_BYTES];

namespace b200ctc {
constexpr int kAlignMaxDevices = 64;
static std::mutex g_align_mu;
static void* g_align_ptr[kAlignMaxDevices] = {};

// device address of g_align_scratch on the CURRENT device (resolved once per device)
static int align_scratch_ptr(int** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kAlignMaxDevices) {
    set_error("forced_align_static: no usable CUDA device (%s)",
              e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
    return B200CTC_ERR_DEVICE;
  }
  std::lock_guard<std::mutex> lk(g_align_mu);
  if (!g_align_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_align_scratch);
    if (e != cudaSuccess || !p) {
      set_error("forced_align_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return B200CTC_ERR_DEVICE;
    }
    g_align_ptr[dev] = p;
  }
  *out = static_cast<int*>(g_align_ptr[dev]);
  return B200CTC_OK;
}
}  // namespace b200ctc

extern "C" int b200ctc_forced_align_static(const float* lp, const int64_t* target, int T, int V,
                                           int tgt_len, int blank, int64_t* out_ids,
                                           b200ctc_stream_t stream) {
  B2_REQUIRE(lp && target && out_ids, "forced_align_static: null pointer");
  B2_REQUIRE(T > 0 && V > 0 && tgt_len >= 1 && blank >= 0 && blank < V,
             "forced_align_static: bad shape");
  const int S = 2 * tgt_len + 1;
  const size_t need = (size_t)T * (size_t)S * sizeof(int);
  if (need > (size_t)B200CTC_ALIGN_STATIC_MAX_BYTES) {
    set_error("forced_align_static: T=%d x (2*tgt_len+1)=%d back-pointers need %zu bytes, the static "
              "scratch holds %zu (B200CTC_ALIGN_STATIC_MAX_BYTES); use b200ctc_forced_align with a "
              "caller workspace", T, S, need, (size_t)B200CTC_ALIGN_STATIC_MAX_BYTES);
    return B200CTC_ERR_WORKSPACE;
  }
  const size_t smem = 2 * (size_t)S * sizeof(double);
  B2_REQUIRE(smem <= 200 * 1024, "forced_align_static: target too long (%d)", tgt_len);
  int* back = nullptr;
  const int rc = align_scratch_ptr(&back);
  if (rc) return rc;
  int nt = (S + 31) / 32 * 32;
  if (nt > 1024) nt = 1024;
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  align_kernel<<<1, nt, smem, static_cast<cudaStream_t>(stream)>>>(lp, target, T, V, tgt_len, blank,
                                                                  out_ids, back);
  count_launch();
  return check_launch("forced_align_static");
}
