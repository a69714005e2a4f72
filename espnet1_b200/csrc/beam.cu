// Beam pruning of the label-synchronous search, fused.
//
// One step of BatchBeamSearch.search (espnet/nets/batch_beam_search.py:206-283)
// after the scorers have run is, in the reference, a chain of ~50 small launches
// and a host loop: weighted sum of the scorers' (n, V) matrices, + running score,
// topk over (beam x V), // and % by V, per-scorer score gathers, prefix append,
// ended-hypothesis masks.  Here it is ONE kernel, one CTA per utterance:
//
//   w[h, v] = ((0 + w_1 s_1[h,v]) + w_2 s_2[h,v]) + ... + run[h]     (same fp32
//             operation order as the reference: separate multiply and add, no FMA)
//   top-k of w over the utterance's (W slots x V) entries, sorted, ties -> lower index
//   for each winner j: prev slot, token, new running score (-inf if the slot ended or
//   the utterance is finished), per-scorer cumulative scores, the prefix row
//   yseq_out[j] = yseq_in[prev] ++ token, and a packed fp32 record for the host.
//
// The same selection core serves the pre-beam (top-P tokens of every hypothesis row,
// :228-234).  Selection: the segment's weighted scores are evaluated once into shared
// memory (<= 56 K values); k rounds of block-wide arg-max (warp shuffles + one
// shared-memory hop), the owner of a winner retires it.  Latency-bound by design: a
// segment is 5 000 to 56 K values, L2-resident; there is one segment per CTA.
#include "common.cuh"

namespace b200ctc {

constexpr int kBeamMaxScorers = B200CTC_BEAM_MAX_SCORERS;
constexpr int kBeamMaxK = 64;  // winners per segment

struct BeamScores {
  const float* p[kBeamMaxScorers];
  long long rs[kBeamMaxScorers], cs[kBeamMaxScorers];  // row / column strides in elements
  float w[kBeamMaxScorers];
  int K;
};

// Phase 1: every thread evaluates its strided share of the segment (rows of V entries
// starting at row0; element e = i*NT + tid), 8 elements x K matrices of independent loads
// in flight, each element accumulating its terms in scorer order (the reference's fp32
// operation order); the values go to shared memory, the thread keeps its own maximum.
// Phase 2: k rounds of block-wide arg-max (warp shuffles + one shared-memory hop, ties ->
// lower index); the owner of a winner retires it and rescans its share.
template <int NT>
__device__ __forceinline__ void segment_topk(const BeamScores& S, long long row0, int V, int E,
                                             const float* __restrict__ run, int k, float* vals,
                                             float* res_v, int* res_e, float* sm_v, int* sm_e) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  constexpr int NWARP = NT / 32, UN = 8;
  float bv = kNegInf;
  int be = 0x7fffffff;
  for (int e0 = tid; e0 < E; e0 += NT * UN) {
    float acc[UN];
    long long row[UN];
    int col[UN];
#pragma unroll
    for (int j = 0; j < UN; ++j) {
      const int e = e0 + j * NT;
      const int ec = e < E ? e : 0;
      const int h = ec / V;
      row[j] = row0 + h;
      col[j] = ec - h * V;
      acc[j] = 0.f;
    }
#pragma unroll 1
    for (int q = 0; q < S.K; ++q) {
      const float* __restrict__ p = S.p[q];
      const long long rs = S.rs[q], cs = S.cs[q];
      const float wq = S.w[q];
      float x[UN];
#pragma unroll
      for (int j = 0; j < UN; ++j) x[j] = __ldg(p + row[j] * rs + (long long)col[j] * cs);
#pragma unroll
      for (int j = 0; j < UN; ++j) acc[j] = __fadd_rn(acc[j], __fmul_rn(wq, x[j]));
    }
#pragma unroll
    for (int j = 0; j < UN; ++j) {
      const int e = e0 + j * NT;
      if (e < E) {
        float x = acc[j];
        if (run) x = __fadd_rn(x, __ldg(run + row[j]));
        if (!(x == x)) x = kNegInf;  // NaN never wins (and NaN is free to mark retired winners)
        vals[e] = x;
        if (x > bv || (x == bv && e < be)) { bv = x; be = e; }  // -inf entries are candidates too
      }
    }
  }
  for (int r = 0; r < k; ++r) {
    float cv = bv;
    int ce = be;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, cv, o);
      const int oe = __shfl_xor_sync(0xffffffffu, ce, o);
      if (ov > cv || (ov == cv && oe < ce)) { cv = ov; ce = oe; }
    }
    if (lane == 0) { sm_v[wid] = cv; sm_e[wid] = ce; }
    __syncthreads();
    if (wid == 0) {
      cv = lane < NWARP ? sm_v[lane] : kNegInf;
      ce = lane < NWARP ? sm_e[lane] : 0x7fffffff;
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ov = __shfl_xor_sync(0xffffffffu, cv, o);
        const int oe = __shfl_xor_sync(0xffffffffu, ce, o);
        if (ov > cv || (ov == cv && oe < ce)) { cv = ov; ce = oe; }
      }
      if (lane == 0) { res_v[r] = cv; res_e[r] = ce < E ? ce : 0; }
    }
    __syncthreads();
    // retire the winner (NaN: compares false with everything, so it can never win again -
    // also when the finite candidates are exhausted and -inf entries are being handed out:
    // the k winners stay DISTINCT, lowest index first, like torch.topk) and rescan my share
    const int we = res_e[r];
    if ((we % NT) == tid) {
      vals[we] = __int_as_float(0x7fc00000);
      bv = kNegInf;
      be = 0x7fffffff;
      for (int e = tid; e < E; e += NT) {
        const float x = vals[e];
        if (x > bv || (x == bv && e < be)) { bv = x; be = e; }
      }
    }
  }
}

template <int NT>
__global__ void __launch_bounds__(NT) beam_select_kernel(
    const __grid_constant__ BeamScores S, const float* __restrict__ run, const float* __restrict__ cum_in,
    const int64_t* __restrict__ yseq_in, long long ys_in_stride, int cur_len,
    const unsigned char* __restrict__ active, const int* __restrict__ maxlen, int step_i, int W,
    int V, int k, int eos, float* __restrict__ out_val, int64_t* __restrict__ out_flat,
    float* __restrict__ out_run, float* __restrict__ cum_out, int64_t* __restrict__ yseq_out,
    long long ys_out_stride, float* __restrict__ pack) {
  extern __shared__ float vals[];  // the segment's W*V weighted scores
  __shared__ float res_v[kBeamMaxK], sm_v[32];
  __shared__ int res_e[kBeamMaxK], sm_e[32];
  const int u = blockIdx.x, U = gridDim.x, tid = threadIdx.x;
  const int E = W * V;
  const long long n_in = (long long)U * W, n_out = (long long)U * k;
  segment_topk<NT>(S, (long long)u * W, V, E, run, k, vals, res_v, res_e, sm_v, sm_e);
  __syncthreads();
  const bool act = (active ? active[u] != 0 : true) && !(maxlen && maxlen[u] - 1 == step_i);
  if (tid < k) {
    const int j = tid, e = res_e[j];
    const float val = res_v[j];
    const int prev = e / V, tok = e - prev * V;
    const long long row = (long long)u * W + prev, o = (long long)u * k + j;
    out_val[o] = val;
    out_flat[o] = e;
    const bool alive = act && tok != eos && val > kNegInf;
    out_run[o] = alive ? val : kNegInf;
    pack[o] = val;
    pack[n_out + o] = (float)prev;
    pack[2 * n_out + o] = (float)tok;
    for (int q = 0; q < S.K; ++q) {
      const float c = __fadd_rn(cum_in[(long long)q * n_in + row],
                                S.p[q][row * S.rs[q] + (long long)tok * S.cs[q]]);
      cum_out[(long long)q * n_out + o] = c;
      pack[(3 + q) * n_out + o] = c;
    }
    yseq_out[o * ys_out_stride + cur_len] = tok;
  }
  for (int j = 0; j < k; ++j) {  // prefix rows: yseq_out[j] = yseq_in[prev_j] ++ tok_j
    const int prev = res_e[j] / V;
    const int64_t* src = yseq_in + ((long long)u * W + prev) * ys_in_stride;
    int64_t* dst = yseq_out + ((long long)u * k + j) * ys_out_stride;
    for (int c = tid; c < cur_len; c += NT) dst[c] = src[c];
  }
}

template <int NT>
__global__ void __launch_bounds__(NT) beam_prebeam_kernel(const __grid_constant__ BeamScores S, int V,
                                                          int k, int64_t* __restrict__ out_ids) {
  extern __shared__ float vals[];
  __shared__ float res_v[kBeamMaxK], sm_v[32];
  __shared__ int res_e[kBeamMaxK], sm_e[32];
  const long long row = blockIdx.x;
  segment_topk<NT>(S, row, V, V, nullptr, k, vals, res_v, res_e, sm_v, sm_e);
  __syncthreads();
  if (threadIdx.x < k) out_ids[row * k + threadIdx.x] = res_e[threadIdx.x];
}

constexpr long long kBeamMaxSegment = 56 * 1024;  // fp32 values of one segment in shared memory

template <typename Kern>
static int beam_smem(Kern kern, size_t bytes) {
  if (bytes > 48 * 1024 &&
      cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes) != cudaSuccess) {
    set_error("beam: cannot reserve %zu bytes of shared memory", bytes);
    cudaGetLastError();
    return B200CTC_ERR_DEVICE;
  }
  return B200CTC_OK;
}

static int fill_scores(BeamScores& S, const float* const* ptrs, const int64_t* strides,
                       const float* weights, int K) {
  B2_REQUIRE(K >= 1 && K <= kBeamMaxScorers, "beam: 1..%d score matrices (got %d)", kBeamMaxScorers, K);
  S.K = K;
  for (int q = 0; q < K; ++q) {
    B2_REQUIRE(ptrs[q] != nullptr, "beam: score matrix %d is NULL", q);
    S.p[q] = ptrs[q];
    S.rs[q] = strides[2 * q];
    S.cs[q] = strides[2 * q + 1];
    S.w[q] = weights[q];
  }
  return B200CTC_OK;
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" int b200ctc_beam_select(const float* const* score_ptrs, const int64_t* strides,
                                   const float* weights, int K, const float* run_score,
                                   const float* cum_in, const int64_t* yseq_in,
                                   int64_t yseq_in_stride, int cur_len, const unsigned char* active,
                                   const int* maxlen, int step_i, int U, int W, int V, int k, int eos,
                                   float* out_val, int64_t* out_flat, float* out_run, float* cum_out,
                                   int64_t* yseq_out, int64_t yseq_out_stride, float* pack,
                                   b200ctc_stream_t stream) {
  BeamScores S;
  if (int rc = fill_scores(S, score_ptrs, strides, weights, K)) return rc;
  B2_REQUIRE(U >= 1 && W >= 1 && V >= 1 && k >= 1 && k <= kBeamMaxK, "beam_select: bad sizes U=%d W=%d V=%d k=%d", U, W, V, k);
  const long long E = (long long)W * V;
  B2_REQUIRE(E <= kBeamMaxSegment, "beam_select: beam x vocabulary = %lld exceeds %lld", E, kBeamMaxSegment);
  B2_REQUIRE(cur_len >= 0 && cur_len < yseq_out_stride && cur_len <= yseq_in_stride, "beam_select: prefix buffers too short");
  auto st = static_cast<cudaStream_t>(stream);
  const size_t smem = (size_t)E * 4;
#define GO(NT)                                                                                     \
  do {                                                                                             \
    if (int rc = beam_smem(beam_select_kernel<NT>, smem)) return rc;                               \
    beam_select_kernel<NT><<<U, NT, smem, st>>>(S, run_score, cum_in, yseq_in, yseq_in_stride,     \
                                                cur_len, active, maxlen, step_i, W, V, k, eos,     \
                                                out_val, out_flat, out_run, cum_out, yseq_out,     \
                                                yseq_out_stride, pack);                            \
  } while (0)
  if (E <= 4096) GO(256);
  else GO(1024);
#undef GO
  count_launch();
  return check_launch("beam_select");
}

extern "C" int b200ctc_beam_prebeam(const float* const* score_ptrs, const int64_t* strides,
                                    const float* weights, int K, int64_t n_rows, int V, int k,
                                    int64_t* out_ids, b200ctc_stream_t stream) {
  BeamScores S;
  if (int rc = fill_scores(S, score_ptrs, strides, weights, K)) return rc;
  B2_REQUIRE(n_rows >= 1 && V >= 1 && k >= 1 && k <= kBeamMaxK && k <= V, "beam_prebeam: bad sizes rows=%lld V=%d k=%d", (long long)n_rows, V, k);
  B2_REQUIRE(V <= kBeamMaxSegment, "beam_prebeam: vocabulary %d exceeds %lld", V, kBeamMaxSegment);
  auto st = static_cast<cudaStream_t>(stream);
  const size_t smem = (size_t)V * 4;
#define GO(NT)                                                                          \
  do {                                                                                  \
    if (int rc = beam_smem(beam_prebeam_kernel<NT>, smem)) return rc;                   \
    beam_prebeam_kernel<NT><<<(unsigned)n_rows, NT, smem, st>>>(S, V, k, out_ids);      \
  } while (0)
  if (V <= 16384) GO(256);
  else GO(1024);
#undef GO
  count_launch();
  return check_launch("beam_prebeam");
}
