// Forced alignment (espnet/nets/pytorch_backend/ctc.py:163-226): Viterbi over the
// blank-interleaved label lattice with back-pointers, then back-tracking.
// The reference keeps the lattice in fp64 (numpy default) with a finite
// log-zero of -1e11 and lets index -1 wrap to the last state for s = 0; both
// are reproduced so that the integer alignment is identical.
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
