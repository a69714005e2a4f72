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

__device__ __forceinline__ unsigned cluster_ctarank() {
  unsigned r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// read one word of CTA `rank`'s shared memory (same offset as `p` in mine)
__device__ __forceinline__ unsigned ld_dsmem_u32(const void* p, int rank) {
  unsigned a = (unsigned)__cvta_generic_to_shared(p), ra, v;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(ra) : "r"(a), "r"(rank));
  asm volatile("ld.shared::cluster.u32 %0, [%1];" : "=r"(v) : "r"(ra) : "memory");
  return v;
}
__device__ __forceinline__ float ld_dsmem_f32(const float* p, int rank) {
  return __uint_as_float(ld_dsmem_u32(p, rank));
}
__device__ __forceinline__ int ld_dsmem_i32(const int* p, int rank) { return (int)ld_dsmem_u32(p, rank); }

struct BeamScores {
  const float* p[kBeamMaxScorers];
  long long rs[kBeamMaxScorers], cs[kBeamMaxScorers];  // row / column strides in elements
  float w[kBeamMaxScorers];
  int K;
};

// The k best of n <= 2048 DISTINCT candidates (value, index; index 0x7fffffff = empty slot) by
// rank counting: every thread takes candidates, counts how many beat each (the candidate list
// is read by all threads in step: shared-memory broadcasts) and files it under its rank.  No
// serial rounds: one warp selecting k times over the list kept the other warps at a barrier
// for a third of the kernel.  Ends with a barrier; res_* hold the winners in order.
template <int NT>
__device__ __forceinline__ void rank_select(const float* cv, const int* ce, int lists, int per, int k,
                                            float* res_v, int* res_e) {
  // The candidates come as `lists` lists of `per` entries, each SORTED (best first, empty slots
  // last): a candidate's rank is its position in its own list plus, per other list, the number
  // of entries that beat it - a binary search (<= 7 steps) instead of a scan.
  const int tid = threadIdx.x, n = lists * per;
  if (tid < k) { res_v[tid] = kNegInf; res_e[tid] = 0; }  // fewer than k candidates: like the first version
  __syncthreads();
  for (int i = tid; i < n; i += NT) {
    const float x = cv[i];
    const int e = ce[i];
    if (e == 0x7fffffff) continue;
    const int own = i / per;
    int rank = i - own * per;
    for (int l = 0; l < lists && rank < k; ++l) {
      if (l == own) continue;
      const float* lv = cv + l * per;
      const int* le = ce + l * per;
      int lo = 0, hi = per;  // first entry of list l that does NOT beat (x, e)
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const float y = lv[mid];
        const int f = le[mid];
        if (f != 0x7fffffff && (y > x || (y == x && f < e))) lo = mid + 1; else hi = mid;
      }
      rank += lo;
    }
    B2_DEV_CHECK(rank >= 0);
    if (rank < k) { res_v[rank] = x; res_e[rank] = e; }
  }
  __syncthreads();
}

// Phase 1: every thread evaluates its strided share of the segment (rows of V entries
// starting at row0; element e = i*NT + tid), 8 elements x K matrices of independent loads
// in flight, each element accumulating its terms in scorer order (the reference's fp32
// operation order); the values go to shared memory, the thread keeps its own maximum.
// Phase 2: the k best, ties -> lower index (two levels, see below).
// `off`: the CTA owns elements [off, off + E) of a larger segment (a cluster shares one
// utterance, below); vals[] is indexed locally, the reported winners res_e[] carry LOCAL
// indices too (the caller adds `off`; ties inside a slice are by index either way).
// ONE_ROW: the segment is one row of V entries (the pre-beam): no row / column split at all.
template <int NT, bool ONE_ROW = false>
__device__ __forceinline__ void segment_topk(const BeamScores& S, long long row0, int V, int E,
                                             const float* __restrict__ run, int k, float* vals,
                                             float* res_v, int* res_e, int off = 0) {
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  constexpr int NWARP = NT / 32, UN = 8;
  // e / V by multiply-high: exact for e, V < 2^16 (segments hold at most 54 K values)
  const unsigned v_magic = V > 1 ? (unsigned)(0x100000000ull / (unsigned)V) + 1u : 0u;
  // the thread's best AND second best: a retired winner is replaced without a rescan of the
  // share (a dependent chain of E/NT shared-memory reads - it, not the barriers, was most of a
  // round); only a thread that wins twice rescans
  float bv = kNegInf, sv = kNegInf;
  int be = 0x7fffffff, se = 0x7fffffff;
  bool stale = false;  // the second best is unknown (it was promoted)
  auto offer = [&](float x, int e) {
    if (x > bv || (x == bv && e < be)) { sv = bv; se = be; bv = x; be = e; }
    else if (x > sv || (x == sv && e < se)) { sv = x; se = e; }
  };
  for (int e0 = tid; e0 < E; e0 += NT * UN) {
    float acc[UN];
    long long row[UN];
    int col[UN];
#pragma unroll
    for (int j = 0; j < UN; ++j) {
      const int e = e0 + j * NT;
      const int ec = (e < E ? e : 0) + off;
      const int h = ONE_ROW ? 0 : (V > 1 ? (int)__umulhi((unsigned)ec, v_magic) : ec);
      row[j] = row0 + h;
      col[j] = ec - h * V;
      acc[j] = 0.f;
    }
    // scorer q + 1's loads are issued before scorer q's values are consumed (two rounds of
    // loads in flight; the accumulation order over q stays the reference's)
    float x[UN];
    auto fetch = [&](int q, float* dst) {
      const float* __restrict__ p = S.p[q];
      const long long rs = S.rs[q], cs = S.cs[q];
#pragma unroll
      for (int j = 0; j < UN; ++j) dst[j] = __ldg(p + row[j] * rs + (long long)col[j] * cs);
    };
    fetch(0, x);
#pragma unroll 1
    for (int q = 0; q < S.K; ++q) {
      const float wq = S.w[q];
      float nx[UN];
      if (q + 1 < S.K) fetch(q + 1, nx);
#pragma unroll
      for (int j = 0; j < UN; ++j) acc[j] = __fadd_rn(acc[j], __fmul_rn(wq, x[j]));
#pragma unroll
      for (int j = 0; j < UN; ++j) x[j] = nx[j];
    }
#pragma unroll
    for (int j = 0; j < UN; ++j) {
      const int e = e0 + j * NT;
      if (e < E) {
        float x = acc[j];
        if (run) x = __fadd_rn(x, __ldg(run + row[j]));
        if (!(x == x)) x = kNegInf;  // NaN never wins (and NaN is free to mark retired winners)
        vals[e] = x;
        offer(x, e);  // -inf entries are candidates too
      }
    }
  }
  // Phase 2, two levels.  (a) every WARP selects the k best of its own lanes' shares: k rounds
  // of shuffle-only arg-max, the owner of a winner retires it (NaN: compares false with
  // everything, so it can never win again - also when the finite candidates are exhausted and
  // -inf entries are being handed out: the winners stay DISTINCT, lowest index first, like
  // torch.topk) and rescans its share.  No block barrier inside the rounds: the first version
  // (k block-wide rounds, two __syncthreads each) spent 1.5 us per round - 23 us of a pre-beam
  // launch, 40 us of a pruning launch.  (b) the global k best are among the NWARP*k warp winners
  // (one total order - value descending, index ascending - at both levels): one warp merges them.
  float* wl_v = vals + ((E + 31) & ~31);  // NWARP * k warp winners, behind the segment
  int* wl_e = reinterpret_cast<int*>(wl_v + NWARP * k);
  for (int r = 0; r < k; ++r) {
    float cv = bv;
    int ce = be;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, cv, o);
      const int oe = __shfl_xor_sync(0xffffffffu, ce, o);
      if (ov > cv || (ov == cv && oe < ce)) { cv = ov; ce = oe; }
    }
    B2_DEV_CHECK(ce == 0x7fffffff || (ce >= 0 && ce < E));
    if (lane == 0) { wl_v[wid * k + r] = cv; wl_e[wid * k + r] = ce; }
    if (ce == be && ce != 0x7fffffff) {  // mine: retire it
      vals[ce] = __int_as_float(0x7fc00000);
      if (!stale) {
        bv = sv; be = se; stale = true;
      } else {
        bv = sv = kNegInf;
        be = se = 0x7fffffff;
        for (int e = tid; e < E; e += NT) offer(vals[e], e);  // NaN (retired) never enters
        stale = false;
      }
    }
  }
  __syncthreads();
  rank_select<NT>(wl_v, wl_e, NWARP, k, k, res_v, res_e);
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
  __shared__ float res_v[kBeamMaxK];
  __shared__ int res_e[kBeamMaxK];
  griddep_wait_chain();  // programmatic stream serialisation: the scorers' kernels come right before
  const int u = blockIdx.x, U = gridDim.x, tid = threadIdx.x;
  const int E = W * V;
  const long long n_in = (long long)U * W, n_out = (long long)U * k;
  segment_topk<NT>(S, (long long)u * W, V, E, run, k, vals, res_v, res_e);
  __syncthreads();
  const bool act = (active ? active[u] != 0 : true) && !(maxlen && maxlen[u] - 1 == step_i);
  if (tid < k) {
    const int j = tid, e = res_e[j];
    const float val = res_v[j];
    B2_DEV_CHECK(e >= 0 && e < W * V);
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
  // prefix rows: yseq_out[j] = yseq_in[prev_j] ++ tok_j - one flat pass (k dependent
  // load -> store round trips in a row cost more than the selection)
  for (int i = tid; i < k * cur_len; i += NT) {
    const int j = i / cur_len, cpos = i - j * cur_len;
    const int prev = res_e[j] / V;
    yseq_out[((long long)u * k + j) * ys_out_stride + cpos] =
        yseq_in[((long long)u * W + prev) * ys_in_stride + cpos];
  }
}

// The same step with one utterance's segment shared by a thread-block CLUSTER of C CTAs: with
// one CTA per utterance 32 utterances occupy 32 of 148 SMs and a CTA walks 50 K values alone
// (62 us per position, as long as the prefix scoring itself).  Every CTA of the cluster takes a
// contiguous slice of the W*V entries, finds its own k best (the global k best are among them:
// one total order - value descending, index ascending - everywhere), and CTA 0 merges the C
// sorted lists out of its neighbours' shared memory (DSMEM) and writes the step's outputs.
template <int NT>
__global__ void __launch_bounds__(NT) beam_select_cluster_kernel(
    const __grid_constant__ BeamScores S, const float* __restrict__ run, const float* __restrict__ cum_in,
    const int64_t* __restrict__ yseq_in, long long ys_in_stride, int cur_len,
    const unsigned char* __restrict__ active, const int* __restrict__ maxlen, int step_i, int W,
    int V, int k, int eos, float* __restrict__ out_val, int64_t* __restrict__ out_flat,
    float* __restrict__ out_run, float* __restrict__ cum_out, int64_t* __restrict__ yseq_out,
    long long ys_out_stride, float* __restrict__ pack, int C, int slice) {
  extern __shared__ float vals[];  // this CTA's slice of the weighted scores
  __shared__ float res_v[kBeamMaxK];
  __shared__ int res_e[kBeamMaxK];
  __shared__ float mg_v[8 * kBeamMaxK];  // CTA 0: the C candidate lists
  __shared__ int mg_e[8 * kBeamMaxK];
  griddep_wait_chain();
  const int c = (int)cluster_ctarank(), u = blockIdx.x / C, U = gridDim.x / C, tid = threadIdx.x;
  const int E = W * V;
  const int off = c * slice, mine = min(slice, E - off);  // the launcher keeps mine >= k
  B2_DEV_CHECK(mine >= k && C <= 8);
  const long long n_in = (long long)U * W, n_out = (long long)U * k;
  segment_topk<NT>(S, (long long)u * W, V, mine, run, k, vals, res_v, res_e, off);
  __syncthreads();
  cluster_sync_all();  // every CTA's res_v / res_e are complete and visible cluster-wide
  if (c == 0) {
    for (int i = tid; i < C * k; i += NT) {
      const int src = i / k, j = i - src * k;
      mg_v[i] = ld_dsmem_f32(&res_v[j], src);
      mg_e[i] = ld_dsmem_i32(&res_e[j], src) + src * slice;
    }
  }
  cluster_sync_all();  // the neighbours may exit: their lists have been copied
  if (c != 0) return;
  __syncthreads();
  rank_select<NT>(mg_v, mg_e, C, k, k, res_v, res_e);
  const bool act = (active ? active[u] != 0 : true) && !(maxlen && maxlen[u] - 1 == step_i);
  if (tid < k) {
    const int j = tid, e = res_e[j];
    const float val = res_v[j];
    B2_DEV_CHECK(e >= 0 && e < W * V);
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
      const float cq = __fadd_rn(cum_in[(long long)q * n_in + row],
                                 S.p[q][row * S.rs[q] + (long long)tok * S.cs[q]]);
      cum_out[(long long)q * n_out + o] = cq;
      pack[(3 + q) * n_out + o] = cq;
    }
    yseq_out[o * ys_out_stride + cur_len] = tok;
  }
  // prefix rows: yseq_out[j] = yseq_in[prev_j] ++ tok_j - one flat pass (k dependent
  // load -> store round trips in a row cost more than the selection)
  for (int i = tid; i < k * cur_len; i += NT) {
    const int j = i / cur_len, cpos = i - j * cur_len;
    const int prev = res_e[j] / V;
    yseq_out[((long long)u * k + j) * ys_out_stride + cpos] =
        yseq_in[((long long)u * W + prev) * ys_in_stride + cpos];
  }
}

template <int NT>
__global__ void __launch_bounds__(NT) beam_prebeam_kernel(const __grid_constant__ BeamScores S, int V,
                                                          int k, int64_t* __restrict__ out_ids) {
  extern __shared__ float vals[];
  __shared__ float res_v[kBeamMaxK];
  __shared__ int res_e[kBeamMaxK];
  griddep_wait_chain();
  const long long row = blockIdx.x;
  segment_topk<NT, true>(S, row, V, V, nullptr, k, vals, res_v, res_e);
  __syncthreads();
  if (threadIdx.x < k) out_ids[row * k + threadIdx.x] = res_e[threadIdx.x];
}

constexpr long long kBeamMaxSegment = 54 * 1024;  // fp32 values of one segment in shared memory
constexpr size_t kBeamSmemBudget = 227 * 1024 - 6 * 1024;  // dynamic; the rest: static arrays
// dynamic shared memory of a selection CTA: its segment + the warps' winner lists
static size_t topk_smem(long long E, int nt, int k) {
  return (size_t)((E + 31) & ~31LL) * 4 + (size_t)(nt / 32) * k * 8;
}

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
  // few utterances x a large segment: a cluster of C CTAs per utterance (see the kernel)
  static const int cluster_on = tuning_int("B200CTC_BEAM_CLUSTER", 1);
  // (one CTA of 512 threads per SM at 90 registers: all clusters in ONE wave of 148 SMs)
  const int C = (!cluster_on || E < 16384) ? 1 : U <= 18 ? 8 : U <= 37 ? 4 : U <= 74 ? 2 : 1;
  if (C > 1) {
    constexpr int NT = 512;
    const int slice = (int)((E + C - 1) / C);
    const size_t smem_c = topk_smem(slice, NT, k);
    if (int rc = beam_smem(beam_select_cluster_kernel<NT>, smem_c)) return rc;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(U * C));
    cfg.blockDim = dim3(NT);
    cfg.dynamicSmemBytes = smem_c;
    cfg.stream = st;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = (unsigned)C;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled(5) ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, beam_select_cluster_kernel<NT>, S, run_score, cum_in, yseq_in,
                                       (long long)yseq_in_stride, cur_len, active, maxlen, step_i, W, V, k, eos,
                                       out_val, out_flat, out_run, cum_out, yseq_out,
                                       (long long)yseq_out_stride, pack, C, slice);
    if (e != cudaSuccess) {
      set_error("beam_select: cluster launch: %s", cudaGetErrorString(e));
      cudaGetLastError();
      return B200CTC_ERR_DEVICE;
    }
    count_launch();
    return check_launch("beam_select");
  }
  // most threads whose winner lists still fit beside the segment
  const int nt = E <= 4096 ? 256 : topk_smem(E, 1024, k) <= kBeamSmemBudget ? 1024
               : topk_smem(E, 512, k) <= kBeamSmemBudget ? 512 : 256;
  const size_t smem = topk_smem(E, nt, k);
  B2_REQUIRE(smem <= kBeamSmemBudget, "beam_select: beam x vocabulary = %lld with k = %d does not fit shared memory", E, k);
#define GO(NT)                                                                                     \
  do {                                                                                             \
    if (int rc = beam_smem(beam_select_kernel<NT>, smem)) return rc;                               \
    launch_pdl(beam_select_kernel<NT>, dim3(U), dim3(NT), smem, st, 5, S, run_score, cum_in, yseq_in,   \
               (long long)yseq_in_stride, cur_len, active, maxlen, step_i, W, V, k, eos, out_val,  \
               out_flat, out_run, cum_out, yseq_out, (long long)yseq_out_stride, pack);            \
  } while (0)
  if (nt == 256) GO(256);
  else if (nt == 512) GO(512);
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
  const int nt = V <= 16384 ? 256 : topk_smem(V, 1024, k) <= kBeamSmemBudget ? 1024
               : topk_smem(V, 512, k) <= kBeamSmemBudget ? 512 : 256;
  const size_t smem = topk_smem(V, nt, k);
  B2_REQUIRE(smem <= kBeamSmemBudget, "beam_prebeam: vocabulary %d with k = %d does not fit shared memory", V, k);
#define GO(NT)                                                                          \
  do {                                                                                  \
    if (int rc = beam_smem(beam_prebeam_kernel<NT>, smem)) return rc;                   \
    launch_pdl(beam_prebeam_kernel<NT>, dim3((unsigned)n_rows), dim3(NT), smem, st, 5, S, V, k, out_ids); \
  } while (0)
  if (nt == 256) GO(256);
  else if (nt == 512) GO(512);
  else GO(1024);
#undef GO
  count_launch();
  return check_launch("beam_prebeam");
}
