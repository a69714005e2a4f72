// Batched CTC prefix scoring (CTCPrefixScoreTH, espnet/nets/ctc_prefix_score.py:11-269).
// One thread per (hypothesis, candidate token): the whole frame loop of the
// reference (:156-161, ~6 kernel launches per frame there) runs inside ONE
// kernel launch per beam step; psi is accumulated on the fly.
// Arithmetic follows the reference formulas in fp32 with its finite log-zero
// (-1e10, sticky under fp32 addition) in the max-shift form m + log(1 + e^-|a-b|).
//
// The step is a dependent chain of T log-sum-exps per thread (latency-bound: ~60 cycles
// per frame), so everything else is kept off it: the frame's inputs (x_t[k], x_t[blank],
// phi_{t-1}) are fetched kPF frames ahead into registers, and with a pre-beam
// (snum << V) x_t[k] comes from a TIME-CONTIGUOUS copy of the posteriors (B,V,T) made
// once per utterance batch - one cache line then serves 32 frames of a thread instead
// of one.  Full-vocabulary scoring reads the (B,T,V) rows directly (coalesced over k).
#include "common.cuh"

namespace b200ctc {

__device__ __forceinline__ float lse2f(float a, float b) {
  const float m = fmaxf(a, b);
  return m + logf(expf(a - m) + expf(b - m));
}
// same value up to rounding on the MUFU units (ex2 / lg2): 2 transcendental ops per call
__device__ __forceinline__ float lse2_mufu(float a, float b) {
  return fmaxf(a, b) + __logf(1.f + __expf(-fabsf(a - b)));
}

// (B,T,V) -> (B,V,T) tiled transpose (32x32 tiles through shared memory)
// frames [tfirst, T) only; `ld` = time stride of the copy (>= T: room for streaming appends)
__global__ void __launch_bounds__(256) prefix_transpose_kernel(const float* __restrict__ in,
                                                               float* __restrict__ out, int T,
                                                               int V, int tfirst, int ld) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, v0 = blockIdx.x * 32, t0 = tfirst + blockIdx.y * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;  // 8 rows per pass
  const float* src = in + (size_t)b * T * V;
  float* dst = out + (size_t)b * V * ld;
  for (int r = ty; r < 32; r += 8)
    if (t0 + r < T && v0 + tx < V) tile[r][tx] = src[(size_t)(t0 + r) * V + v0 + tx];
  __syncthreads();
  for (int r = ty; r < 32; r += 8)
    if (v0 + r < V && t0 + tx < T) dst[(size_t)(v0 + r) * ld + t0 + tx] = tile[tx][r];
}

// frames t >= xlens[b]: every label logzero, blank 0   (:47-50)
__global__ void __launch_bounds__(256) prefix_pad_kernel(float* __restrict__ lp,
                                                         const int64_t* __restrict__ xlens, int T,
                                                         int V, int blank, float logzero) {
  const int t = blockIdx.x, b = blockIdx.y;
  if (t < xlens[b]) return;
  float* row = lp + ((size_t)b * T + t) * V;
  for (int v = threadIdx.x; v < V; v += 256) row[v] = (v == blank) ? 0.f : logzero;
}

// r0[t,0,:] = logzero ; r0[t,1,b*n_hyps+h] = sum_{tau<=t} lp[b,tau,blank]  (:83-91)
// sequential fp32 accumulation (what torch.cumsum does on CPU); one warp per b
__global__ void __launch_bounds__(32) prefix_init_kernel(const float* __restrict__ lp, int B, int T,
                                                         int V, int n_hyps, int blank,
                                                         float logzero, float* __restrict__ r0) {
  const int b = blockIdx.x, lane = threadIdx.x;
  const size_t n_bh = (size_t)B * n_hyps;
  float acc = 0.f;
  for (int t0 = 0; t0 < T; t0 += 32) {
    const int t = t0 + lane;
    const float x = t < T ? lp[((size_t)b * T + t) * V + blank] : 0.f;
    float mine = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float xi = __shfl_sync(0xffffffffu, x, i);
      if (t0 + i < T) acc = (t0 + i == 0) ? xi : acc + xi;
      if (i == lane) mine = acc;
    }
    if (t < T) {
      for (int h = 0; h < n_hyps; ++h) {
        r0[((size_t)t * 2 + 0) * n_bh + (size_t)b * n_hyps + h] = logzero;
        r0[((size_t)t * 2 + 1) * n_bh + (size_t)b * n_hyps + h] = mine;
      }
    }
  }
}

// f_arg[h] = sum_t att_w[h,t] * t ; min / max over h   (:145-147)
// one warp per hypothesis row, 32 warps; a row's loads are all in flight before the
// (fixed-order) accumulation starts
constexpr int kWinWarps = 32;
__global__ void __launch_bounds__(kWinWarps * 32) prefix_window_kernel(
    const float* __restrict__ att_w, int n_bh, int T, float* __restrict__ minmax) {
  __shared__ float smin[kWinWarps], smax[kWinWarps];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float lo = -kNegInf, hi = kNegInf;
  for (int h = wid; h < n_bh; h += kWinWarps) {
    const float* row = att_w + (size_t)h * T;
    float acc = 0.f;
    for (int t0 = lane; t0 < T; t0 += 32 * 16) {
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = (t0 + 32 * j < T) ? __ldg(row + t0 + 32 * j) : 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (t0 + 32 * j < T) acc = fmaf(v[j], (float)(t0 + 32 * j), acc);
    }
    acc = warp_sum(acc);
    lo = fminf(lo, acc);
    hi = fmaxf(hi, acc);
  }
  if (lane == 0) { smin[wid] = lo; smax[wid] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < kWinWarps; ++i) { lo = fminf(lo, smin[i]); hi = fmaxf(hi, smax[i]); }
    minmax[0] = lo;
    minmax[1] = hi;
  }
}

// Per-utterance windows (opt-in extension; the reference takes ONE window over the
// whole batch, :144-153): the same formula applied to the hypotheses of each utterance alone -
// i.e. what the reference computes when the utterance is decoded by itself.  Everything stays
// on the device (no read-back): fprev / fnew (B,2) int32 = {f_min, f_max}, win (B,2) = {start, end}.
__global__ void __launch_bounds__(kWinWarps * 32) prefix_window_utt_kernel(
    const float* __restrict__ att_w, int n_hyps, int T, int margin, int out_len,
    const int* __restrict__ fprev, int* __restrict__ fnew, int* __restrict__ win) {
  __shared__ float smin[kWinWarps], smax[kWinWarps];
  const int b = blockIdx.x, lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float lo = -kNegInf, hi = kNegInf;
  for (int h = wid; h < n_hyps; h += kWinWarps) {
    const float* row = att_w + ((size_t)b * n_hyps + h) * T;
    float acc = 0.f;
    for (int t0 = lane; t0 < T; t0 += 32 * 16) {  // same accumulation order as the global kernel
      float v[16];
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = (t0 + 32 * j < T) ? __ldg(row + t0 + 32 * j) : 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (t0 + 32 * j < T) acc = fmaf(v[j], (float)(t0 + 32 * j), acc);
    }
    acc = warp_sum(acc);
    lo = fminf(lo, acc);
    hi = fmaxf(hi, acc);
  }
  if (lane == 0) { smin[wid] = lo; smax[wid] = hi; }
  __syncthreads();
  if (threadIdx.x == 0) {
    for (int i = 1; i < kWinWarps; ++i) { lo = fminf(lo, smin[i]); hi = fmaxf(hi, smax[i]); }
    const int pmin = fprev ? fprev[2 * b] : 0, pmax = fprev ? fprev[2 * b + 1] : 1;  // (:95)
    const int f_min = max((int)lo, pmin), f_max = max((int)hi, pmax);  // int(): toward zero
    fnew[2 * b] = f_min;
    fnew[2 * b + 1] = f_max;
    win[2 * b] = min(pmax, max(max(f_min - margin, out_len), 1));
    win[2 * b + 1] = min(f_max + margin, T);
  }
}

// r_sum[t,h] = logsumexp(r_prev[t,0,h], r_prev[t,1,h])   (:131)
__global__ void __launch_bounds__(256) prefix_rsum_kernel(const float* __restrict__ r_prev, int T,
                                                          int n_bh, float* __restrict__ rsum) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)T * n_bh) return;
  const size_t t = i / n_bh, h = i - t * n_bh;
  rsum[i] = lse2f(r_prev[(t * 2 + 0) * n_bh + h], r_prev[(t * 2 + 1) * n_bh + h]);
}

// pre-beam only: log_psi / scores / idmap defaults for tokens that are not scored
__global__ void __launch_bounds__(256) prefix_fill_kernel(
    const float* __restrict__ rsum, const float* __restrict__ s_prev,
    const int64_t* __restrict__ end_frames, int n_bh, int n_hyps, int V, int blank, int eos,
    float logzero, float* __restrict__ log_psi, float* __restrict__ scores,
    int64_t* __restrict__ idmap) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)n_bh * V) return;
  const int h = (int)(i / V), v = (int)(i - (size_t)h * V);
  float p = logzero;
  if (v == eos && v != blank) p = rsum[(size_t)end_frames[h / n_hyps] * n_bh + h];
  log_psi[i] = p;
  scores[i] = p - (s_prev ? s_prev[i] : 0.f);
  idmap[i] = -1;
}

constexpr int kPF = 16;  // frames fetched ahead of the recursion
#ifndef B200CTC_PSI_BLOCKED
#define B200CTC_PSI_BLOCKED 0  // 1: psi reduced per group of kPF frames (tuning build, see profiles/README.md)
#endif

// TC: lp_t is the time-contiguous copy (B,V,T); else lp rows (B,T,V) are read directly
template <bool TC>
__global__ void __launch_bounds__(128) prefix_step_kernel(
    const float* __restrict__ lp, const float* __restrict__ lp_t, const float* __restrict__ r_prev,
    const float* __restrict__ rsum, const float* __restrict__ s_prev,
    const int64_t* __restrict__ last_ids, const int64_t* __restrict__ scoring_ids,
    const int64_t* __restrict__ end_frames, int n_bh, int n_hyps, int T, int V, int snum,
    int out_len, int start, int end, int blank, int eos, float logzero, float* __restrict__ r,
    float* __restrict__ log_psi, float* __restrict__ scores, int64_t* __restrict__ idmap, int ldt,
    const int* __restrict__ win_utt) {
  const size_t idx = (size_t)blockIdx.x * 128 + threadIdx.x;
  if (idx >= (size_t)n_bh * snum) return;
  const int h = (int)(idx / snum);
  const int b = h / n_hyps;
  if (win_utt) {  // per-utterance window instead of the batch-wide one
    start = win_utt[2 * b];
    end = win_utt[2 * b + 1];
  }
  const int k = scoring_ids ? (int)scoring_ids[idx] : (int)(idx - (size_t)h * snum);
  const bool same = (k == (int)last_ids[h]);
  // per-frame input streams: x_t[k], x_t[blank], phi_{t-1} (= r^b_prev if k repeats the
  // last label, else logsumexp(r^n_prev, r^b_prev), :132-141) with their strides over t
  const float* xk = TC ? lp_t + ((size_t)b * V + k) * ldt : lp + (size_t)b * T * V + k;
  const float* xb = TC ? lp_t + ((size_t)b * V + blank) * ldt : lp + (size_t)b * T * V + blank;
  const size_t xs = TC ? 1 : (size_t)V;
  const float* ph = same ? r_prev + n_bh + h : rsum + h;
  const size_t phs = same ? 2 * (size_t)n_bh : (size_t)n_bh;
  const size_t plane = (size_t)n_bh * snum;  // one (t, n|b) plane of r
  float* rout = r + idx;
  // frames before the window stay logzero; r^n_0 = x_0[k] for the empty prefix (:129-130)
  float rn = logzero, rb = logzero;
  for (int t = 0; t < start; ++t) {
    const float v0 = (t == 0 && out_len == 0) ? xk[0] : logzero;
    __stcs(rout + ((size_t)t * 2 + 0) * plane, v0);
    __stcs(rout + ((size_t)t * 2 + 1) * plane, logzero);
    if (t == start - 1) rn = v0;
  }
  // psi = logsumexp({phi_{t-1} + x_t[k]}_{t in window} U {r^n_{start-1}})  (:164-179),
  // accumulated on line as (running max pm, sum ps of e^(term - pm))
  float pm = rn, ps = 1.f;
  float cx[kPF], cb[kPF], cp[kPF];
  auto fetch = [&](int t0) {  // clamped: always a valid address, values beyond `end` unused
#pragma unroll
    for (int j = 0; j < kPF; ++j) {
      const int t = (t0 + j < end) ? t0 + j : end - 1;
      cx[j] = __ldg(xk + (size_t)t * xs);
      cb[j] = __ldg(xb + (size_t)t * xs);
      cp[j] = __ldg(ph + (size_t)(t - 1) * phs);
    }
  };
  if (start < end) fetch(start);
  for (int t0 = start; t0 < end; t0 += kPF) {
    float x_k[kPF], x_b[kPF], phi[kPF];
#pragma unroll
    for (int j = 0; j < kPF; ++j) { x_k[j] = cx[j]; x_b[j] = cb[j]; phi[j] = cp[j]; }
    if (t0 + kPF < end) fetch(t0 + kPF);  // next group: in flight while this one is computed
#if B200CTC_PSI_BLOCKED
    // psi is a plain reduction over the frames (no recursion): one maximum and one rescale
    // per group of kPF terms instead of a running-maximum update per frame, and none of it
    // sits on the r^n / r^b chain below
    {
      float term[kPF], tm = -__builtin_huge_valf();
#pragma unroll
      for (int j = 0; j < kPF; ++j) {
        term[j] = (t0 + j < end) ? phi[j] + x_k[j] : -__builtin_huge_valf();
        tm = fmaxf(tm, term[j]);
      }
      const float m2 = fmaxf(pm, tm);
      float acc = ps * __expf(pm - m2);
#pragma unroll
      for (int j = 0; j < kPF; ++j) acc += __expf(term[j] - m2);
      ps = acc;
      pm = m2;
    }
#endif
#pragma unroll
    for (int j = 0; j < kPF; ++j) {
      const int t = t0 + j;
      if (t < end) {
        const float rn_new = lse2_mufu(rn, phi[j]) + x_k[j];
        const float rb_new = lse2_mufu(rn, rb) + x_b[j];
#if !B200CTC_PSI_BLOCKED
        const float term = phi[j] + x_k[j];
        const float m2 = fmaxf(pm, term);
        ps = ps * __expf(pm - m2) + __expf(term - m2);
        pm = m2;
#endif
        __stcs(rout + ((size_t)t * 2 + 0) * plane, rn_new);
        __stcs(rout + ((size_t)t * 2 + 1) * plane, rb_new);
        rn = rn_new;
        rb = rb_new;
      }
    }
  }
  for (int t = end; t < T; ++t) {
    __stcs(rout + ((size_t)t * 2 + 0) * plane, logzero);
    __stcs(rout + ((size_t)t * 2 + 1) * plane, logzero);
  }
  float psi = pm + __logf(ps);
  const size_t o = (size_t)h * V + k;
  if (scoring_ids) {
    idmap[o] = (int64_t)(idx - (size_t)h * snum);
    if (k == eos || k == blank) return;  // set by prefix_fill_kernel (:181-185)
  } else {
    if (k == eos) psi = rsum[(size_t)end_frames[b] * n_bh + h];
    if (k == blank) psi = logzero;
  }
  log_psi[o] = psi;
  scores[o] = psi - (s_prev ? s_prev[o] : 0.f);
}

// index_select_state (:189-220)
__global__ void __launch_bounds__(256) prefix_select_r_kernel(
    const float* __restrict__ r, const int64_t* __restrict__ idmap,
    const int64_t* __restrict__ best_ids, int n_hyps, int W, int T, int V, int snum, int n_bh,
    float* __restrict__ r_new) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  const size_t n_new = (size_t)(n_bh / n_hyps) * W;
  if (i >= (size_t)T * 2 * n_new) return;
  const size_t tn = i / n_new;
  const int hn = (int)(i - tn * n_new);
  const int b = hn / W;
  const int64_t best = best_ids[hn];
  const int hyp = (int)(best / V) + b * n_hyps;
  const int label = (int)(best % V);
  int col = label;
  if (idmap) {
    const int64_t c = idmap[(size_t)hyp * V + label];
    col = c < 0 ? 0 : (int)c;
  }
  r_new[i] = r[(tn * n_bh + hyp) * snum + col];
}
__global__ void __launch_bounds__(256) prefix_select_s_kernel(
    const float* __restrict__ log_psi, const int64_t* __restrict__ best_ids, int n_hyps, int W,
    int V, int n_new, float* __restrict__ s_new) {
  const size_t i = (size_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (size_t)n_new * V) return;
  const int hn = (int)(i / V);
  const int b = hn / W;
  s_new[i] = log_psi[(size_t)b * n_hyps * V + best_ids[hn]];
}

static inline unsigned blocks(size_t n, int per) { return (unsigned)((n + per - 1) / per); }

}  // namespace b200ctc

using namespace b200ctc;

extern "C" int b200ctc_prefix_pad(float* lp, const int64_t* xlens, int B, int T, int V, int blank,
                                  float logzero, b200ctc_stream_t stream) {
  B2_REQUIRE(lp && xlens, "prefix_pad: null pointer");
  B2_REQUIRE(B >= 0 && T >= 0 && V > 0 && blank >= 0 && blank < V, "prefix_pad: bad shape");
  if (B == 0 || T == 0) return B200CTC_OK;
  B2_REQUIRE(B <= 65535, "prefix_pad: batch too large");
  prefix_pad_kernel<<<dim3(T, B), 256, 0, static_cast<cudaStream_t>(stream)>>>(lp, xlens, T, V,
                                                                               blank, logzero);
  count_launch();
  return check_launch("prefix_pad");
}

extern "C" int b200ctc_prefix_init_state(const float* lp, int B, int T, int V, int n_hyps,
                                         int blank, float logzero, float* r0,
                                         b200ctc_stream_t stream) {
  B2_REQUIRE(lp && r0, "prefix_init_state: null pointer");
  B2_REQUIRE(B > 0 && T > 0 && V > 0 && n_hyps > 0, "prefix_init_state: bad shape");
  prefix_init_kernel<<<B, 32, 0, static_cast<cudaStream_t>(stream)>>>(lp, B, T, V, n_hyps, blank,
                                                                      logzero, r0);
  count_launch();
  return check_launch("prefix_init_state");
}

extern "C" int b200ctc_prefix_window(const float* att_w, int n_bh, int T, float* minmax,
                                     b200ctc_stream_t stream) {
  B2_REQUIRE(att_w && minmax && n_bh > 0 && T > 0, "prefix_window: bad argument");
  prefix_window_kernel<<<1, kWinWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(att_w, n_bh, T,
                                                                                    minmax);
  count_launch();
  return check_launch("prefix_window");
}

extern "C" int b200ctc_prefix_window_utt(const float* att_w, int B, int n_hyps, int T, int margin,
                                         int out_len, const int* fprev, int* fnew, int* win,
                                         b200ctc_stream_t stream) {
  B2_REQUIRE(att_w && fnew && win && B > 0 && n_hyps > 0 && T > 0 && margin >= 0 && out_len >= 0,
             "prefix_window_utt: bad argument");
  prefix_window_utt_kernel<<<B, kWinWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
      att_w, n_hyps, T, margin, out_len, fprev, fnew, win);
  count_launch();
  return check_launch("prefix_window_utt");
}

extern "C" int b200ctc_prefix_transpose(const float* lp, float* lp_t, int B, int T, int V, int t0,
                                        int lp_t_ld, b200ctc_stream_t stream) {
  B2_REQUIRE(lp && lp_t, "prefix_transpose: null pointer");
  B2_REQUIRE(B >= 0 && T >= 0 && V > 0, "prefix_transpose: bad shape");
  B2_REQUIRE(t0 >= 0 && t0 <= T && lp_t_ld >= T, "prefix_transpose: bad range t0=%d T=%d ld=%d", t0, T,
             lp_t_ld);
  if (B == 0 || T == t0) return B200CTC_OK;
  const int nt = T - t0;
  B2_REQUIRE(B <= 65535 && (nt + 31) / 32 <= 65535, "prefix_transpose: shape too large");
  prefix_transpose_kernel<<<dim3((V + 31) / 32, (nt + 31) / 32, B), 256, 0,
                            static_cast<cudaStream_t>(stream)>>>(lp, lp_t, T, V, t0, lp_t_ld);
  count_launch();
  return check_launch("prefix_transpose");
}

extern "C" int b200ctc_prefix_step(const float* lp, const float* lp_t, const float* r_prev,
                                   const float* s_prev,
                                   const int64_t* last_ids, const int64_t* scoring_ids,
                                   const int64_t* end_frames, int B, int n_hyps, int T, int V,
                                   int snum, int out_len, int start, int end, int blank, int eos,
                                   float logzero, float* r, float* log_psi, float* scores,
                                   int64_t* idmap, float* rsum_ws, int lp_t_ld, const int* win_utt,
                                   b200ctc_stream_t stream) {
  B2_REQUIRE(lp && r_prev && last_ids && end_frames && r && log_psi && scores && rsum_ws,
             "prefix_step: null pointer");
  B2_REQUIRE(!lp_t || lp_t_ld >= T, "prefix_step: lp_t_ld=%d < T=%d", lp_t_ld, T);
  B2_REQUIRE(B > 0 && n_hyps > 0 && T > 0 && V > 0 && snum > 0, "prefix_step: bad shape");
  B2_REQUIRE(scoring_ids ? idmap != nullptr : snum == V,
             "prefix_step: scoring_ids/idmap/snum mismatch (snum=%d V=%d)", snum, V);
  B2_REQUIRE(win_utt || (start >= 1 && start <= T && end <= T), "prefix_step: bad window [%d,%d)", start,
             end);
  B2_REQUIRE(blank >= 0 && blank < V && eos >= 0 && eos < V, "prefix_step: bad blank / eos id");
  const int n_bh = B * n_hyps;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  prof_mark(4, 0, st);
  prefix_rsum_kernel<<<blocks((size_t)T * n_bh, 256), 256, 0, st>>>(r_prev, T, n_bh, rsum_ws);
  count_launch();
  if (scoring_ids) {
    prefix_fill_kernel<<<blocks((size_t)n_bh * V, 256), 256, 0, st>>>(
        rsum_ws, s_prev, end_frames, n_bh, n_hyps, V, blank, eos, logzero, log_psi, scores, idmap);
    count_launch();
  }
  // the time-contiguous copy pays off when few of the V columns are touched per hypothesis
  if (lp_t && scoring_ids)
    prefix_step_kernel<true><<<blocks((size_t)n_bh * snum, 128), 128, 0, st>>>(
        lp, lp_t, r_prev, rsum_ws, s_prev, last_ids, scoring_ids, end_frames, n_bh, n_hyps, T, V,
        snum, out_len, start, end, blank, eos, logzero, r, log_psi, scores, idmap, lp_t_ld, win_utt);
  else
    prefix_step_kernel<false><<<blocks((size_t)n_bh * snum, 128), 128, 0, st>>>(
        lp, lp_t, r_prev, rsum_ws, s_prev, last_ids, scoring_ids, end_frames, n_bh, n_hyps, T, V,
        snum, out_len, start, end, blank, eos, logzero, r, log_psi, scores, idmap, lp_t_ld, win_utt);
  count_launch();
  prof_mark(4, 1, st);
  return check_launch("prefix_step");
}

extern "C" int b200ctc_prefix_select(const float* r, const float* log_psi, const int64_t* idmap,
                                     const int64_t* best_ids, int B, int n_hyps, int W, int T,
                                     int V, int snum, float* r_new, float* s_new,
                                     b200ctc_stream_t stream) {
  B2_REQUIRE(r && log_psi && best_ids && r_new && s_new, "prefix_select: null pointer");
  B2_REQUIRE(B > 0 && n_hyps > 0 && W > 0 && T > 0 && V > 0 && snum > 0, "prefix_select: bad shape");
  const int n_bh = B * n_hyps, n_new = B * W;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  prefix_select_r_kernel<<<blocks((size_t)T * 2 * n_new, 256), 256, 0, st>>>(
      r, idmap, best_ids, n_hyps, W, T, V, snum, n_bh, r_new);
  prefix_select_s_kernel<<<blocks((size_t)n_new * V, 256), 256, 0, st>>>(log_psi, best_ids, n_hyps,
                                                                         W, V, n_new, s_new);
  count_launch(2);
  return check_launch("prefix_select");
}
