// Batched CTC prefix scoring (CTCPrefixScoreTH, espnet/nets/ctc_prefix_score.py:11-269).
// The whole frame loop of the reference (:156-161, ~6 kernel launches per frame there) runs
// inside ONE kernel launch per beam step; psi is accumulated on the fly.
// Arithmetic follows the reference formulas in fp32 with its finite log-zero
// (-1e10, sticky under fp32 addition) in the max-shift form m + log(1 + e^-|a-b|).
//
// Two kernels serve the recursion:
//  * prefix_step_kernel - one thread per (hypothesis, token) walks the frames: a dependent
//    chain of T log-sum-exps per thread, its inputs (x_t[k], x_t[blank], phi_{t-1}) fetched
//    kPF frames ahead into registers.  Full-vocabulary scoring (1.6 M threads at C3, bound by
//    the write of r) and the fallback of the kernel below.
//  * prefix_step_scan_kernel - with a pre-beam (snum << V, 4 800 pairs at C3) that chain ran on
//    1 % of the machine's threads; the recursion is an affine map per frame, so a WARP takes a
//    pair, every lane a chunk of frames, and a warp scan of the chunk maps makes the frame
//    loop parallel in time (see the kernel).  x_t[k] comes from a TIME-CONTIGUOUS copy of the
//    posteriors (B,V,T) made once per utterance batch.
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

// (B,T,V) -> (B,V,T) tiled transpose: 64x64 tiles through shared memory, 256-byte runs on both
// sides (32x32 tiles moved 4 KB per CTA: 80 K CTAs at C3, 1.4 TB/s)
// frames [tfirst, T) only; `ld` = time stride of the copy (>= T: room for streaming appends)
constexpr int kTrTile = 64;
__global__ void __launch_bounds__(256) prefix_transpose_kernel(const float* __restrict__ in,
                                                               float* __restrict__ out, int T,
                                                               int V, int tfirst, int ld) {
  __shared__ float tile[kTrTile][kTrTile + 1];
  const int b = blockIdx.z, v0 = blockIdx.x * kTrTile, t0 = tfirst + blockIdx.y * kTrTile;
  const int tx = threadIdx.x & 63, ty = threadIdx.x >> 6;  // 4 rows per pass
  const float* src = in + (size_t)b * T * V;
  float* dst = out + (size_t)b * V * ld;
#pragma unroll 4
  for (int r = ty; r < kTrTile; r += 4)
    if (t0 + r < T && v0 + tx < V) tile[r][tx] = __ldcs(src + (size_t)(t0 + r) * V + v0 + tx);
  __syncthreads();
#pragma unroll 4
  for (int r = ty; r < kTrTile; r += 4)
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
// Accumulated sequentially in DOUBLE and rounded per element - what torch.cumsum does for a
// float tensor on the CPU (its accumulation type there is double), i.e. what the reference's
// decoding computes.  An fp32 running sum is ~1e-3 off at T = 500 (the sum reaches -2000,
// ulp 1.2e-4), which then sits in every score of the first position
// (tools/exp/prefix_err.py: 7.7e-4 against 4.6e-5 for the reference's arithmetic).  One warp per b.
__global__ void __launch_bounds__(32) prefix_init_kernel(const float* __restrict__ lp, int B, int T,
                                                         int V, int n_hyps, int blank,
                                                         float logzero, float* __restrict__ r0) {
  const int b = blockIdx.x, lane = threadIdx.x;
  const size_t n_bh = (size_t)B * n_hyps;
  double acc = 0.0;
  for (int t0 = 0; t0 < T; t0 += 32) {
    const int t = t0 + lane;
    const float x = t < T ? lp[((size_t)b * T + t) * V + blank] : 0.f;
    float mine = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float xi = __shfl_sync(0xffffffffu, x, i);
      if (t0 + i < T) acc += (double)xi;
      if (i == lane) mine = (float)acc;
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

// Two passes in ONE launch:
//  r_sum[t,h] = logsumexp(r_prev[t,0,h], r_prev[t,1,h])   (:131)
//  pre-beam only: log_psi / scores / idmap defaults for tokens that are not scored (:181-185).
// The fill's one dependent entry (eos: r_sum at the utterance's last frame) is recomputed from
// r_prev with the same expression instead of being read back, so the passes need no ordering.
// (rows of 256-wide blocks: no per-thread index divisions)
__global__ void __launch_bounds__(256) prefix_prepare_kernel(
    const float* __restrict__ r_prev, const float* __restrict__ s_prev,
    const int64_t* __restrict__ end_frames, int T, int n_bh, int n_hyps, int V, int blank, int eos,
    float logzero, int fill, float* __restrict__ rsum, float* __restrict__ log_psi,
    float* __restrict__ scores, int64_t* __restrict__ idmap) {
  griddep_wait_chain();  // launched with programmatic stream serialisation: inputs come from the previous kernel
  // blocks [0, T * cb1): r_sum rows (t, 256 hypotheses); the rest: fill rows (h, 256 tokens) -
  // one 32-bit division per BLOCK
  const int cb1 = (n_bh + 255) / 256, cb2 = (V + 255) / 256;
  const int blk = blockIdx.x;
  if (blk < T * cb1) {
    const int row = blk / cb1, col = (blk - row * cb1) * 256 + threadIdx.x;
    const size_t t = row;
    if (col < n_bh)
      rsum[t * n_bh + col] = lse2f(r_prev[(t * 2 + 0) * n_bh + col], r_prev[(t * 2 + 1) * n_bh + col]);
    return;
  }
  const int fb = blk - T * cb1;
  const int row = fb / cb2, col = (fb - row * cb2) * 256 + threadIdx.x;
  if (fill && row < n_bh && col < V) {  // row = h, col = v
    const int h = row, v = col;
    const size_t i = (size_t)h * V + v;
    float p = logzero;
    if (v == eos && v != blank) {
      const size_t t = (size_t)end_frames[h / n_hyps];
      p = lse2f(r_prev[(t * 2 + 0) * n_bh + h], r_prev[(t * 2 + 1) * n_bh + h]);
    }
    log_psi[i] = p;
    scores[i] = p - (s_prev ? s_prev[i] : 0.f);
    idmap[i] = -1;
  }
}

constexpr int kPF = 16;  // frames fetched ahead of the recursion

// TC: lp_t is the time-contiguous copy (B,V,T); else lp rows (B,T,V) are read directly.
// AH: groups of kPF frames in flight ahead of the recursion.  With a pre-beam the launch is
// small (n_bh * snum = 4 800 threads at C3) and its rows are cold in DRAM: the launcher then
// uses 32-thread CTAs (one warp per SM - the per-frame loads of a warp touch 32 different
// rows, so four warps on one SM queue behind one load/store unit), AH = 2, and - when the
// copy's time stride allows 16-byte loads - frame groups ALIGNED to kPF so that a thread's
// 16 frames of x_t[k] are four LDG.128 instead of sixteen LDG.32 (frames outside [start, end)
// of the first and last group are fetched and ignored).  The arithmetic per frame is unchanged.
template <bool TC, int AH>
__global__ void __launch_bounds__(128) prefix_step_kernel(
    const float* __restrict__ lp, const float* __restrict__ lp_t, const float* __restrict__ r_prev,
    const float* __restrict__ rsum, const float* __restrict__ s_prev,
    const int64_t* __restrict__ last_ids, const int64_t* __restrict__ scoring_ids,
    const int64_t* __restrict__ end_frames, int n_bh, int n_hyps, int T, int V, int snum,
    int out_len, int start, int end, int blank, int eos, float logzero, float* __restrict__ r,
    float* __restrict__ log_psi, float* __restrict__ scores, int64_t* __restrict__ idmap, int ldt,
    const int* __restrict__ win_utt, int vec, long long last_stride) {
  const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= (size_t)n_bh * snum) return;
  const int h = (int)(idx / snum);
  const int b = h / n_hyps;
  if (win_utt) {  // per-utterance window instead of the batch-wide one
    start = win_utt[2 * b];
    end = win_utt[2 * b + 1];
  }
  const int k = scoring_ids ? (int)scoring_ids[idx] : (int)(idx - (size_t)h * snum);
  const bool same = (k == (int)last_ids[(size_t)h * last_stride]);
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
  float cx[AH][kPF], cb[AH][kPF], cp[AH][kPF];
  const bool aligned = TC && vec != 0;
  auto fetch = [&](int a, int g) {  // clamped: always a valid address, frames outside the window unused
    if (aligned && g + kPF <= ldt) {
      const float4* pk = reinterpret_cast<const float4*>(xk + g);
      const float4* pb = reinterpret_cast<const float4*>(xb + g);
#pragma unroll
      for (int j = 0; j < kPF / 4; ++j) {
        const float4 vk = __ldg(pk + j), vb = __ldg(pb + j);
        cx[a][4 * j] = vk.x; cx[a][4 * j + 1] = vk.y; cx[a][4 * j + 2] = vk.z; cx[a][4 * j + 3] = vk.w;
        cb[a][4 * j] = vb.x; cb[a][4 * j + 1] = vb.y; cb[a][4 * j + 2] = vb.z; cb[a][4 * j + 3] = vb.w;
      }
    } else {
#pragma unroll
      for (int j = 0; j < kPF; ++j) {
        const int t = min(max(g + j, start), end - 1);
        cx[a][j] = __ldg(xk + (size_t)t * xs);
        cb[a][j] = __ldg(xb + (size_t)t * xs);
      }
    }
#pragma unroll
    for (int j = 0; j < kPF; ++j) {
      const int t = min(max(g + j, start), end - 1);  // start >= 1: t - 1 >= 0
      cp[a][j] = __ldg(ph + (size_t)(t - 1) * phs);
    }
  };
  const int g0 = aligned ? (start & ~(kPF - 1)) : start;
#pragma unroll
  for (int a = 0; a < AH; ++a)
    if (g0 + a * kPF < end) fetch(a, g0 + a * kPF);
  for (int gq = g0; gq < end; gq += AH * kPF) {
#pragma unroll
    for (int a = 0; a < AH; ++a) {
      const int t0 = gq + a * kPF;
      if (t0 < end) {
        float x_k[kPF], x_b[kPF], phi[kPF];
#pragma unroll
        for (int j = 0; j < kPF; ++j) { x_k[j] = cx[a][j]; x_b[j] = cb[a][j]; phi[j] = cp[a][j]; }
        if (t0 + AH * kPF < end) fetch(a, t0 + AH * kPF);  // in flight while this group is computed
#pragma unroll
        for (int j = 0; j < kPF; ++j) {
          const int t = t0 + j;
          if (t >= start && t < end) {
            const float rn_new = lse2_mufu(rn, phi[j]) + x_k[j];
            const float rb_new = lse2_mufu(rn, rb) + x_b[j];
            const float term = phi[j] + x_k[j];
            const float m2 = fmaxf(pm, term);
            ps = ps * __expf(pm - m2) + __expf(term - m2);
            pm = m2;
            __stcs(rout + ((size_t)t * 2 + 0) * plane, rn_new);
            __stcs(rout + ((size_t)t * 2 + 1) * plane, rb_new);
            rn = rn_new;
            rb = rb_new;
          }
        }
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
    if (k == eos || k == blank) return;  // set by prefix_prepare_kernel (:181-185)
  } else {
    if (k == eos) psi = rsum[(size_t)end_frames[b] * n_bh + h];
    if (k == blank) psi = logzero;
  }
  log_psi[o] = psi;
  scores[o] = psi - (s_prev ? s_prev[o] : 0.f);
}

// ---------------------------------------------------------------------------------------
// The same step with the frame loop PARALLEL IN TIME, for the pre-beam case (few candidates).
//
// With a pre-beam the step is tiny and serial: n_bh * snum = 4 800 threads at C3 each walking
// 500 dependent frames (~250 cycles per frame with one warp per scheduler and nothing to hide
// its latencies behind): 66 us per position, a third of a decoding position, on 1 % of the
// machine's threads.  But the recursion is linear in the probability domain,
//     R^n_t = (R^n_{t-1} + Phi_{t-1}) X_t[k]        R^b_t = (R^n_{t-1} + R^b_{t-1}) X_t[blank]
// (:156-161; Phi does not depend on the state), i.e. an affine map of (R^n, R^b) with a
// lower-triangular matrix.  Such maps compose into maps of the same shape,
//     (a, b, d, c_n, c_b):  R^n' = a R^n + c_n,   R^b' = b R^n + d R^b + c_b,
// so a WARP takes one (hypothesis, candidate) pair, every lane a chunk of ceil(len/32) frames:
//   1. the lane composes the maps of its frames (in the log domain, all terms non-negative in
//      the probability domain: sums only, no cancellation) and its share of psi;
//   2. a warp scan of the 32 chunk maps gives every lane the state at the start of its chunk;
//   3. from there the lane runs the reference recursion itself over its frames, so every
//      r_t comes out of the same formula as in the serial kernel (only the chunk-start states
//      see a different summation order; differences are at rounding level, tests as before).
// One CTA = one hypothesis and up to 16 of its candidates: phi_{t-1} (both variants) and
// x_t[blank] are staged once per CTA, x_t[k] per warp, all coalesced; results are staged in
// shared memory (r^n over x_t[k]) and written as rows of r.  Chunks are stored with an odd
// stride so that the lanes' chunk walks are bank-conflict free.
constexpr float kScanZero = -1e30f;  // log of 0 for the map coefficients (finite: no inf - inf)
constexpr int kScanCand = 16;        // candidates (warps) per CTA

struct ScanMap { float a, b, d, cn, cb; };

// `m` applied AFTER `e`
__device__ __forceinline__ ScanMap scan_compose(const ScanMap& e, const ScanMap& m) {
  ScanMap o;
  o.a = m.a + e.a;
  o.d = m.d + e.d;
  o.b = lse2_mufu(m.b + e.a, m.d + e.b);
  o.cn = lse2_mufu(m.a + e.cn, m.cn);
  o.cb = lse2_mufu(lse2_mufu(m.b + e.cn, m.d + e.cb), m.cb);
  return o;
}

__global__ void __launch_bounds__(32 * kScanCand, 2) prefix_step_scan_kernel(
    const float* __restrict__ lp_t, const float* __restrict__ r_prev, const float* __restrict__ rsum,
    const float* __restrict__ s_prev, const int64_t* __restrict__ last_ids,
    const int64_t* __restrict__ scoring_ids, int n_bh, int n_hyps, int T, int V, int snum, int out_len,
    int start, int end, int blank, int eos, float logzero, float* __restrict__ r,
    float* __restrict__ log_psi, float* __restrict__ scores, int64_t* __restrict__ idmap, int ldt,
    const int* __restrict__ win_utt, int cap /* frames a staged array can hold */, long long last_stride,
    int time_major /* r as (n_bh, snum, 2, T): the position call's private layout */) {
  extern __shared__ float sm[];
  griddep_wait_chain();
  const int h = blockIdx.y, c0 = blockIdx.x * kScanCand;
  const int ncand = min(kScanCand, snum - c0);
  const int b = h / n_hyps, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, NT = blockDim.x;
  if (win_utt) {
    start = win_utt[2 * b];
    end = win_utt[2 * b + 1];
  }
  const int len = max(end - start, 0);
  const int c = (len + 31) / 32, cs = c | 1;  // frames per lane; odd chunk stride in shared memory
  const int S = cap + 1;                      // per-array stride (odd multiple pattern across arrays)
  float* phiA = sm;                 // logsumexp(r^n_prev, r^b_prev)[t-1]
  float* phiB = sm + S;             // r^b_prev[t-1]
  float* xbl = sm + 2 * S;          // x_t[blank]
  float* xk_all = sm + 3 * S;       // per warp: x_t[k], later r^n_t
  float* rb_all = xk_all + (NT >> 5) * S;  // per warp: r^b_t
  __shared__ float v0s[kScanCand];
  // i / c and e / ncand by multiply-high: exact for operands below 2^16 (frames, candidates)
  const unsigned c_magic = c > 1 ? (unsigned)(0x100000000ull / (unsigned)c) + 1u : 0u;
  auto slot = [&](int i) {
    const int L = c > 1 ? (int)__umulhi((unsigned)i, c_magic) : i;
    B2_DEV_CHECK(i >= 0 && i < len && L == i / (c > 0 ? c : 1) && i + L * (cs - c) < cap);
    return i + L * (cs - c);
  };

  // ---- stage the CTA's inputs: frame i of the window is absolute frame start + i
  const float* xb_row = lp_t + ((size_t)b * V + blank) * ldt + start;
  for (int i = tid; i < len; i += NT) {
    const int sl = slot(i);
    const size_t tp = (size_t)(start + i - 1);  // start >= 1
    phiA[sl] = __ldg(rsum + tp * n_bh + h);
    phiB[sl] = __ldg(r_prev + (tp * 2 + 1) * n_bh + h);
    xbl[sl] = __ldg(xb_row + i);
  }
  const bool have = wid < ncand;
  const size_t idx = (size_t)h * snum + c0 + wid;
  int k = 0;
  bool same = false;
  float* xk = xk_all + wid * S;
  float* rbo = rb_all + wid * S;
  if (have) {
    k = (int)scoring_ids[idx];
    B2_DEV_CHECK(k >= 0 && k < V && idx < (size_t)n_bh * snum);
    same = (k == (int)last_ids[(size_t)h * last_stride]);
    const float* xk_row = lp_t + ((size_t)b * V + k) * ldt;
    for (int i = lane; i < len; i += 32) xk[slot(i)] = __ldg(xk_row + start + i);
    if (lane == 0) v0s[wid] = (out_len == 0) ? __ldg(xk_row) : logzero;  // r^n_0 of the empty prefix (:129-130)
  }
  __syncthreads();

  if (have) {
    const float* phi = same ? phiB : phiA;
    const int base = lane * cs, i0 = lane * c;
    const int nmine = max(0, min(c, len - i0));
    // r^n / r^b at frame start-1: logzero, except r^n_0 = x_0[k] when the window starts at frame 1
    const float rn0 = (start == 1) ? ((out_len == 0) ? __ldg(lp_t + ((size_t)b * V + k) * ldt) : logzero) : logzero;
    const float rb0 = logzero;
    // ---- 1. this lane's chunk as one affine map, and its share of psi
    ScanMap m = {0.f, kScanZero, 0.f, kScanZero, kScanZero};
    float pm = (lane == 0) ? rn0 : kScanZero, ps = (lane == 0) ? 1.f : 0.f;
    for (int j = 0; j < nmine; ++j) {
      const float x_k = xk[base + j], x_b = xbl[base + j], ph = phi[base + j];
      const float nb = x_b + lse2_mufu(m.a, m.b);
      const float ncb = x_b + lse2_mufu(m.cn, m.cb);
      m.cn = x_k + lse2_mufu(m.cn, ph);
      m.a += x_k;
      m.d += x_b;
      m.b = nb;
      m.cb = ncb;
      const float term = ph + x_k;
      const float m2 = fmaxf(pm, term);
      ps = ps * __expf(pm - m2) + __expf(term - m2);
      pm = m2;
    }
    // ---- 2. inclusive scan of the maps over the lanes, then shift by one lane
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      ScanMap e;
      e.a = __shfl_up_sync(0xffffffffu, m.a, o);
      e.b = __shfl_up_sync(0xffffffffu, m.b, o);
      e.d = __shfl_up_sync(0xffffffffu, m.d, o);
      e.cn = __shfl_up_sync(0xffffffffu, m.cn, o);
      e.cb = __shfl_up_sync(0xffffffffu, m.cb, o);
      if (lane >= o) m = scan_compose(e, m);
    }
    ScanMap pre;
    pre.a = __shfl_up_sync(0xffffffffu, m.a, 1);
    pre.b = __shfl_up_sync(0xffffffffu, m.b, 1);
    pre.d = __shfl_up_sync(0xffffffffu, m.d, 1);
    pre.cn = __shfl_up_sync(0xffffffffu, m.cn, 1);
    pre.cb = __shfl_up_sync(0xffffffffu, m.cb, 1);
    float rn = rn0, rb = rb0;
    if (lane > 0) {
      rn = lse2_mufu(pre.a + rn0, pre.cn);
      rb = lse2_mufu(lse2_mufu(pre.b + rn0, pre.d + rb0), pre.cb);
    }
    // ---- 3. the reference recursion over my frames from that state; r^n replaces x_t[k]
    for (int j = 0; j < nmine; ++j) {
      const float x_k = xk[base + j], x_b = xbl[base + j], ph = phi[base + j];
      const float rn_new = lse2_mufu(rn, ph) + x_k;
      const float rb_new = lse2_mufu(rn, rb) + x_b;
      xk[base + j] = rn_new;
      rbo[base + j] = rb_new;
      rn = rn_new;
      rb = rb_new;
    }
    // ---- psi = logsumexp over the window's terms and r^n_{start-1} (:164-179)
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, pm, o), os = __shfl_xor_sync(0xffffffffu, ps, o);
      const float m2 = fmaxf(pm, om);
      ps = ps * __expf(pm - m2) + os * __expf(om - m2);
      pm = m2;
    }
    if (lane == 0) {
      const size_t o = (size_t)h * V + k;
      idmap[o] = (int64_t)(c0 + wid);
      if (k != eos && k != blank) {  // those two are set by prefix_prepare_kernel (:181-185)
        const float psi = pm + __logf(ps);
        log_psi[o] = psi;
        scores[o] = psi - (s_prev ? s_prev[o] : 0.f);
      }
    }
    if (time_major) {
      // the pair's two rows of T frames are contiguous: the warp streams them out itself,
      // 128 bytes per store instruction, no CTA-wide pass and no partial-sector rows
      __syncwarp();
      float* rn_row = r + idx * 2 * (size_t)T;
      float* rb_row = rn_row + T;
      B2_DEV_CHECK(start >= 1 && end <= T);
      for (int t = lane; t < T; t += 32) {
        float vn = logzero, vb = logzero;
        if (t >= start && t < end) {
          const int sl = slot(t - start);
          vn = xk[sl];
          vb = rbo[sl];
        } else if (t == 0) {
          vn = (out_len == 0) ? __ldg(lp_t + ((size_t)b * V + k) * ldt) : logzero;
        }
        __stcs(rn_row + t, vn);
        __stcs(rb_row + t, vb);
      }
    }
  }
  if (time_major) return;
  __syncthreads();

  // ---- rows of r: (t, n|b) -> ncand contiguous values, one half-warp per row
  const size_t plane = (size_t)n_bh * snum;
  const int cc = tid & 15, nhw = NT >> 4;
  float* rdst = r + (size_t)h * snum + c0 + cc;
  for (int row = tid >> 4; row < 2 * T; row += nhw) {
    const int t = row >> 1, x = row & 1;
    if (cc < ncand) {
      float v = logzero;
      if (t >= start && t < end) {
        v = (x ? rb_all : xk_all)[cc * S + slot(t - start)];
      } else if (row == 0) {
        v = v0s[cc];
      }
      __stcs(rdst + (size_t)row * plane, v);
    }
  }
}

// index_select_state (:189-220): both gathers in one launch
__global__ void __launch_bounds__(256) prefix_select_kernel(
    const float* __restrict__ r, const float* __restrict__ log_psi, const int64_t* __restrict__ idmap,
    const int64_t* __restrict__ best_ids, int n_hyps, int W, int T, int V, int snum, int n_bh,
    float* __restrict__ r_new, float* __restrict__ s_new, int time_major) {
  griddep_wait_chain();
  // rows of 256-wide blocks - first the r gather's, then the s gather's - and 32-bit index
  // arithmetic (one division per block, none per thread)
  const int n_new = (n_bh / n_hyps) * W;
  const int rows_r = time_major ? 2 * n_new : 2 * T, cols_r = time_major ? T : n_new;
  const int cb_r = (cols_r + 255) / 256, cb_s = (V + 255) / 256;
  const int blk = blockIdx.x;
  const bool r_part = blk < rows_r * cb_r;
  const int fb = r_part ? blk : blk - rows_r * cb_r, cbk = r_part ? cb_r : cb_s;
  const int row = fb / cbk, c = (fb - row * cbk) * 256 + threadIdx.x;
  // the winner's (hypothesis, column of r) for new slot hn
  auto source = [&](int hn, int& hyp, int& col) {
    const unsigned best = (unsigned)best_ids[hn];  // < n_hyps * V
    B2_DEV_CHECK(best_ids[hn] >= 0 && best_ids[hn] < (int64_t)n_hyps * V);
    const unsigned ph = best / (unsigned)V;
    const int label = (int)(best - ph * (unsigned)V);
    hyp = (int)ph + (hn / W) * n_hyps;
    col = label;
    if (idmap) {
      const int64_t m = idmap[(size_t)hyp * V + label];
      col = m < 0 ? 0 : (int)m;
    }
    B2_DEV_CHECK(hyp >= 0 && hyp < n_bh && col >= 0 && col < snum);
  };
  if (!r_part) {  // row = hn, column = v: the winner's psi for every token (:206-209)
    if (row < n_new && c < V)
      s_new[(size_t)row * V + c] = log_psi[(size_t)(row / W) * n_hyps * V + best_ids[row]];
    return;
  }
  if (time_major) {
    // r is (n_bh, snum, 2, T): row = (hn, x), consecutive threads walk t (coalesced reads), the
    // (T,2,n_new) state is written with a stride - the scatter sits on the write side
    if (row < 2 * n_new && c < T) {
      const int hn = row >> 1, x = row & 1;
      int hyp, col;
      source(hn, hyp, col);
      r_new[((size_t)c * 2 + x) * n_new + hn] = r[(((size_t)hyp * snum + col) * 2 + x) * T + c];
    }
  } else if (row < 2 * T && c < n_new) {  // row = (t, x), column = hn
    int hyp, col;
    source(c, hyp, col);
    r_new[(size_t)row * n_new + c] = r[((size_t)row * n_bh + hyp) * snum + col];
  }
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
  B2_REQUIRE(B <= 65535 && (nt + kTrTile - 1) / kTrTile <= 65535, "prefix_transpose: shape too large");
  prefix_transpose_kernel<<<dim3((V + kTrTile - 1) / kTrTile, (nt + kTrTile - 1) / kTrTile, B), 256, 0,
                            static_cast<cudaStream_t>(stream)>>>(lp, lp_t, T, V, t0, lp_t_ld);
  count_launch();
  return check_launch("prefix_transpose");
}

// few candidates per hypothesis: the scan kernel takes the step (and its shared memory fits).
// It does ~3x the arithmetic of the serial kernel on all SMs; the serial kernel's latency-bound
// time barely grows with the number of pairs, so beyond ~12 k pairs it wins again (B200, snum 15,
// T=500: 9 600 pairs 79 vs 92 us per step, 14 400 pairs 110 vs 101; tools/exp/scan_crossover.py)
constexpr size_t kScanMaxPairs = 12288;
bool b200ctc::prefix_scan_applies(bool have_lp_t, bool have_ids, int n_bh, int snum, int T) {
  static const int scan_on = tuning_int("B200CTC_PREFIX_SCAN", 1);
  const int scan_warps = snum < kScanCand ? snum : kScanCand;
  const int scan_cap = 32 * (((T + 31) / 32) | 1);
  const size_t scan_smem = (size_t)(3 + 2 * scan_warps) * (scan_cap + 1) * sizeof(float);
  return have_lp_t && have_ids && scan_on && (size_t)n_bh * snum <= kScanMaxPairs && T >= 64 && T < 2048 &&
         n_bh <= 65535 && scan_smem <= 100 * 1024;
}

// b200ctc_prefix_step with the last token of hypothesis h at last_ids[h * last_stride] (the
// position call reads it straight out of the prefix buffer) and, optionally, r in the position
// call's private time-major layout (n_bh, snum, 2, T)
int b200ctc::prefix_step_impl(const float* lp, const float* lp_t, const float* r_prev, const float* s_prev,
                              const int64_t* last_ids, long long last_stride, const int64_t* scoring_ids,
                              const int64_t* end_frames, int B, int n_hyps, int T, int V, int snum,
                              int out_len, int start, int end, int blank, int eos, float logzero, float* r,
                              float* log_psi, float* scores, int64_t* idmap, float* rsum_ws, int lp_t_ld,
                              const int* win_utt, int r_time_major, b200ctc_stream_t stream) {
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
  {
    const size_t nblk = (size_t)T * blocks(n_bh, 256) + (scoring_ids ? (size_t)n_bh * blocks(V, 256) : 0);
    B2_REQUIRE(nblk < (1ull << 31), "prefix_step: shape too large");
    launch_pdl(prefix_prepare_kernel, dim3((unsigned)nblk), dim3(256), 0, st, 5, r_prev, s_prev,
               end_frames, T, n_bh, n_hyps, V, blank, eos, logzero, (int)(scoring_ids != nullptr), rsum_ws,
               log_psi, scores, idmap);
    count_launch();
  }
  // the time-contiguous copy pays off when few of the V columns are touched per hypothesis
  const size_t pairs = (size_t)n_bh * snum;
  const int scan_warps = snum < kScanCand ? snum : kScanCand;
  const int scan_cap = 32 * (((T + 31) / 32) | 1);
  const size_t scan_smem = (size_t)(3 + 2 * scan_warps) * (scan_cap + 1) * sizeof(float);
  const bool scan = prefix_scan_applies(lp_t != nullptr, scoring_ids != nullptr, n_bh, snum, T);
  B2_REQUIRE(scan || !r_time_major, "prefix_step: the time-major layout of r is the scan kernel's only");
  if (scan) {
    // few candidates: the frame loop parallel in time, one warp per (hypothesis, candidate)
    if (scan_smem > 48 * 1024 &&  // per device: set on every call (the process may use several)
        cudaFuncSetAttribute(prefix_step_scan_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int)scan_smem) != cudaSuccess) {
      set_error("prefix_step: cannot reserve %zu bytes of shared memory for the scan kernel", scan_smem);
      cudaGetLastError();
      return B200CTC_ERR_DEVICE;
    }
    launch_pdl(prefix_step_scan_kernel, dim3((snum + kScanCand - 1) / kScanCand, n_bh), dim3(32 * scan_warps),
               scan_smem, st, 5, lp_t, r_prev, (const float*)rsum_ws, s_prev, last_ids, scoring_ids, n_bh, n_hyps,
               T, V, snum, out_len, start, end, blank, eos, logzero, r, log_psi, scores, idmap, lp_t_ld, win_utt,
               scan_cap, last_stride, r_time_major);
  } else if (lp_t && scoring_ids) {
    // small launch: spread the warps over the SMs (see the kernel); 16-byte loads need a
    // 16-byte aligned copy whose time stride is a multiple of 4 frames
    static const int small_nt = tuning_int("B200CTC_PREFIX_SMALL_NT", 32);
    static const int ahead = tuning_int("B200CTC_PREFIX_AHEAD", 2);
    static const int vec_on = tuning_int("B200CTC_PREFIX_VEC", 1);
    const int nt = pairs <= (size_t)2 * 148 * small_nt ? small_nt : pairs <= (size_t)2 * 148 * 64 ? 64 : 128;
    const int vec = vec_on && lp_t_ld % 4 == 0 && (reinterpret_cast<uintptr_t>(lp_t) & 15) == 0;
    if (ahead >= 2 && nt < 128)
      prefix_step_kernel<true, 2><<<blocks(pairs, nt), nt, 0, st>>>(
          lp, lp_t, r_prev, rsum_ws, s_prev, last_ids, scoring_ids, end_frames, n_bh, n_hyps, T, V,
          snum, out_len, start, end, blank, eos, logzero, r, log_psi, scores, idmap, lp_t_ld, win_utt, vec,
          last_stride);
    else
      prefix_step_kernel<true, 1><<<blocks(pairs, nt), nt, 0, st>>>(
          lp, lp_t, r_prev, rsum_ws, s_prev, last_ids, scoring_ids, end_frames, n_bh, n_hyps, T, V,
          snum, out_len, start, end, blank, eos, logzero, r, log_psi, scores, idmap, lp_t_ld, win_utt, vec,
          last_stride);
  } else {
    prefix_step_kernel<false, 1><<<blocks(pairs, 128), 128, 0, st>>>(
        lp, lp_t, r_prev, rsum_ws, s_prev, last_ids, scoring_ids, end_frames, n_bh, n_hyps, T, V,
        snum, out_len, start, end, blank, eos, logzero, r, log_psi, scores, idmap, lp_t_ld, win_utt, 0,
        last_stride);
  }
  count_launch();
  prof_mark(4, 1, st);
  return check_launch("prefix_step");
}

extern "C" int b200ctc_prefix_step(const float* lp, const float* lp_t, const float* r_prev,
                                   const float* s_prev,
                                   const int64_t* last_ids, const int64_t* scoring_ids,
                                   const int64_t* end_frames, int B, int n_hyps, int T, int V,
                                   int snum, int out_len, int start, int end, int blank, int eos,
                                   float logzero, float* r, float* log_psi, float* scores,
                                   int64_t* idmap, float* rsum_ws, int lp_t_ld, const int* win_utt,
                                   b200ctc_stream_t stream) {
  return prefix_step_impl(lp, lp_t, r_prev, s_prev, last_ids, 1, scoring_ids, end_frames, B, n_hyps, T, V, snum,
                          out_len, start, end, blank, eos, logzero, r, log_psi, scores, idmap, rsum_ws,
                          lp_t_ld, win_utt, 0, stream);
}

int b200ctc::prefix_select_impl(const float* r, const float* log_psi, const int64_t* idmap,
                                const int64_t* best_ids, int B, int n_hyps, int W, int T, int V, int snum,
                                float* r_new, float* s_new, int r_time_major, b200ctc_stream_t stream) {
  B2_REQUIRE(r && log_psi && best_ids && r_new && s_new, "prefix_select: null pointer");
  B2_REQUIRE(B > 0 && n_hyps > 0 && W > 0 && T > 0 && V > 0 && snum > 0, "prefix_select: bad shape");
  const int n_bh = B * n_hyps, n_new = B * W;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int rows_r = r_time_major ? 2 * n_new : 2 * T, cols_r = r_time_major ? T : n_new;
  const size_t nblk = (size_t)rows_r * blocks(cols_r, 256) + (size_t)n_new * blocks(V, 256);
  B2_REQUIRE(nblk < (1ull << 31) && (long long)n_hyps * V < (1ll << 31), "prefix_select: shape too large");
  launch_pdl(prefix_select_kernel, dim3((unsigned)nblk), dim3(256), 0, st, 5, r, log_psi, idmap,
             best_ids, n_hyps, W, T, V, snum, n_bh, r_new, s_new, r_time_major);
  count_launch();
  return check_launch("prefix_select");
}

extern "C" int b200ctc_prefix_select(const float* r, const float* log_psi, const int64_t* idmap,
                                     const int64_t* best_ids, int B, int n_hyps, int W, int T,
                                     int V, int snum, float* r_new, float* s_new,
                                     b200ctc_stream_t stream) {
  return prefix_select_impl(r, log_psi, idmap, best_ids, B, n_hyps, W, T, V, snum, r_new, s_new, 0, stream);
}
