// CTC backward, row pass "B": the gradient w.r.t. the logits is produced in ONE
// streaming pass - each (b,t) row is read once (TMA bulk copy into a shared
// memory ring), softmax and the lattice occupancies are combined on chip and the
// (B,T,V) gradient is written exactly once:
//     grad[b,t,v] = g_b * ( softmax(x)[v] - sum_{s: l_s = v} gamma_t(s) )
// (= autograd of log_softmax + ctc_loss, espnet2/asr/ctc.py:54-55; the
// analytically-zero term softmax * sum_v g_lp is dropped.)
#include "rows_common.cuh"

namespace b200ctc {

template <typename T, int NT, int PER, bool TMA>
__global__ void __launch_bounds__(NT) grad_rows_kernel(
    const T* __restrict__ logits, T* __restrict__ grad, int B, int Tfr, int V, int blank,
    const UttInfo* __restrict__ utt, const int* __restrict__ lab, const int* __restrict__ dup,
    int Lp, int Sp, const float* __restrict__ lse2, const float* __restrict__ qt,
    const float* __restrict__ alpha, const float* __restrict__ beta,
    const float* __restrict__ gscale, int gstride, float gmul, int nstage, int stage_stride) {
  constexpr int N = Elems<T>::N;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t bars[kMaxRowStages];
  __shared__ float wp[32];
  unsigned char* stages = dsm;
  float* pbuf = reinterpret_cast<float*>(dsm + (size_t)nstage * stage_stride);  // [Lp]
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t nrows = (int64_t)B * Tfr;
  const uint32_t row_bytes = (uint32_t)V * sizeof(T);
  const int nchunk = V / N;
  const int ntail = V - nchunk * N;

  RowCursor pc;
  pc.init(blockIdx.x, gridDim.x, Tfr);
  uint32_t np = 0;
  auto issue_next = [&]() {
    while (pc.g < nrows && (pc.t >= utt[pc.b].T || utt[pc.b].zeroed)) pc.next();
    if (pc.g < nrows) {
      const int st = np % nstage;
      mbar_expect_tx(&bars[st], row_bytes);
      bulk_g2s(stages + (size_t)st * stage_stride, logits + pc.g * V, row_bytes, &bars[st]);
      ++np;
      pc.next();
    }
  };
  if (TMA) {
    if (tid == 0) {
      for (int i = 0; i < nstage; ++i) mbar_init(&bars[i], 1);
      fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0)
      for (int i = 0; i < nstage; ++i) issue_next();
  }

  uint32_t n = 0;
  RowCursor rc;
  rc.init(blockIdx.x, gridDim.x, Tfr);
  // per-row scalars are fetched one row ahead (their latency hides behind the current row)
  float l2_nx = rc.g < nrows ? lse2[rc.g] : 0.f;
  float q_nx = rc.g < nrows ? qt[rc.g] : 0.f;
  for (; rc.g < nrows; rc.next()) {
    const int b = rc.b, t = rc.t;
    const int64_t g = rc.g;
    const float l2 = l2_nx, q = q_nx;
    if (g + rc.stride < nrows) {
      l2_nx = lse2[g + rc.stride];
      q_nx = qt[g + rc.stride];
    }
    const UttInfo u = utt[b];
    T* gr = grad + g * V;
    const bool al16 = (reinterpret_cast<uintptr_t>(gr) & 15) == 0;
    if (t >= u.T || u.zeroed) {  // padded frame or zeroed utterance: gradient is 0
      float z[N];
#pragma unroll
      for (int e = 0; e < N; ++e) z[e] = 0.f;
      if (al16) {
        for (int i = tid; i < nchunk; i += NT) store16<T>(gr + i * N, z);
        if (tid < ntail) gr[nchunk * N + tid] = from_f32<T>(0.f);
      } else {
        for (int i = tid; i < V; i += NT) gr[i] = from_f32<T>(0.f);
      }
      continue;
    }
    const int L = u.L, S = 2 * L + 1;
    float gb = gmul * (gscale ? gscale[(size_t)b * gstride] : 1.f);
    float o_t = q;  // = mt + offA + offB - aend - lam  (finalize_kernel)
    if (u.infinite) {  // inf loss kept (!zero_infinity): the reference's gradient is NaN
      gb = __int_as_float(0x7fc00000);
      o_t = gb;
    }
    const float* al = alpha + (size_t)g * Sp;
    const float* be = beta + (size_t)g * Sp;
    const int* lb = lab + (size_t)b * Lp;
    constexpr int SM = 3;  // lattice positions per thread whose loads are issued before the wait
    float ab[SM];
#pragma unroll
    for (int j = 0; j < SM; ++j) {
      const int s = tid + j * NT;
      ab[j] = s < S ? al[s] + be[s] : kNegInf;
    }
    const unsigned char* stg;
    if (TMA) {
      const int st = n % nstage;
      stg = stages + (size_t)st * stage_stride;
      mbar_wait(&bars[st], (n / nstage) & 1);
    } else {
      stg = stages;
      const T* src = logits + g * V;
      T* dst = reinterpret_cast<T*>(stages);
      for (int i = tid; i < V; i += NT) dst[i] = src[i];
      __syncthreads();
    }
    const T* xs = reinterpret_cast<const T*>(stg);
    // ---- phase 1a: occupancies gamma_t(s) = 2^(alpha + beta - lp2(l_s) + o_t)
    float blank_part = 0.f;
    for (int s0 = 0; s0 < S; s0 += NT * SM) {
#pragma unroll
      for (int j = 0; j < SM; ++j) {
        const int s = s0 + tid + j * NT;
        if (s < S) {
          const float abv = s0 == 0 ? ab[j] : al[s] + be[s];
          const bool odd = s & 1;
          const int k = odd ? (s + 1) >> 1 : 0;
          const int v = odd ? lb[k] : blank;
          const float lp2 = fmaf(to_f32<T>(xs[v]), kLog2e, -l2);
          const float p = ex2(abv - lp2 + o_t);
          if (odd) pbuf[k] = p; else blank_part += p;
        }
      }
    }
    blank_part = warp_sum(blank_part);
    if (lane == 0) wp[wid] = blank_part;
    // ---- phase 1b: dense part, g_b * softmax(x), straight from registers
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = tid + j * NT;
      if (i < nchunk) {
        float x[N];
        unpack16<T>(reinterpret_cast<const uint4*>(stg)[i], x);
#pragma unroll
        for (int e = 0; e < N; ++e) x[e] = gb * ex2(fmaf(x[e], kLog2e, -l2));
        if (al16) store16<T>(gr + i * N, x);
        else
#pragma unroll
          for (int e = 0; e < N; ++e) gr[i * N + e] = from_f32<T>(x[e]);
      }
    }
    if (tid < ntail) {
      const int i = nchunk * N + tid;
      gr[i] = from_f32<T>(gb * ex2(fmaf(to_f32<T>(xs[i]), kLog2e, -l2)));
    }
    __syncthreads();
    // ---- phase 2: the <= L+1 columns that carry occupancy are overwritten by
    // their single owner thread (deterministic for repeated labels)
    const int* db = dup + (size_t)b * Lp;
    for (int k = tid + 1; k <= L; k += NT) {
      const int code = db[k];
      if (code & kDupFirst) {
        float acc = pbuf[k];
        for (int nx = code & kDupMask; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
        const int v = lb[k];
        const float sm = ex2(fmaf(to_f32<T>(xs[v]), kLog2e, -l2));
        gr[v] = from_f32<T>(gb * (sm - acc));
      }
    }
    if (tid == NT - 1) {
      float acc = 0.f;
      for (int w2 = 0; w2 < NT / 32; ++w2) acc += wp[w2];
      for (int nx = u.blank_head; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
      const float sm = ex2(fmaf(to_f32<T>(xs[blank]), kLog2e, -l2));
      gr[blank] = from_f32<T>(gb * (sm - acc));
    }
    __syncthreads();  // stage, pbuf and wp are free again
    if (TMA && tid == 0) issue_next();
    ++n;
  }
}

template <typename T, int NT>
static int launch_pass_b(const void* logits, void* grad, int B, int Tfr, int V, int blank,
                         const LossLayout& w, const unsigned char* ws, const float* gscale,
                         int gstride, float gmul, cudaStream_t st) {
  constexpr int PER = 40 / Elems<T>::N;
  const size_t extra = align_up(sizeof(float) * (size_t)w.Lp, 128);
  const RowLaunchCfg c = row_launch_cfg(logits, (size_t)V * sizeof(T), extra, NT, "B200CTC_B");
  B2_REQUIRE((size_t)c.nstage * c.stride + extra <= 226 * 1024,
             "ctc_loss bwd: V=%d too large for the shared-memory row path", V);
  const size_t smem = (size_t)c.nstage * c.stride + extra;
  const int64_t rows = (int64_t)B * Tfr;
  int64_t grid = (int64_t)sm_count() * c.ctas_per_sm;
  if (grid > rows) grid = rows;
  auto go = [&](auto k) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<(unsigned)grid, NT, smem, st>>>(
        static_cast<const T*>(logits), static_cast<T*>(grad), B, Tfr, V, blank,
        reinterpret_cast<const UttInfo*>(ws + w.utt), reinterpret_cast<const int*>(ws + w.lab),
        reinterpret_cast<const int*>(ws + w.dup), w.Lp, w.Sp,
        reinterpret_cast<const float*>(ws + w.lse2), reinterpret_cast<const float*>(ws + w.qt),
        reinterpret_cast<const float*>(ws + w.alpha), reinterpret_cast<const float*>(ws + w.beta),
        gscale, gstride, gmul, c.nstage, c.stride);
  };
  if (c.tma) go(grad_rows_kernel<T, NT, PER, true>);
  else go(grad_rows_kernel<T, NT, PER, false>);
  count_launch();
  return check_launch("ctc grad_rows");
}

template <typename T>
static int dispatch_pass_b(const void* logits, void* grad, int B, int Tfr, int V, int blank,
                           const LossLayout& w, const unsigned char* ws, const float* gscale,
                           int gstride, float gmul, cudaStream_t st) {
  const int need = (V + 39) / 40;
  if (need <= 128) return launch_pass_b<T, 128>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, st);
  if (need <= 256) return launch_pass_b<T, 256>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, st);
  if (need <= 512) return launch_pass_b<T, 512>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, st);
  if (need <= 1024) return launch_pass_b<T, 1024>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, st);
  set_error("ctc_loss bwd: V=%d exceeds the supported row size", V);
  return B200CTC_ERR_ARG;
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" int b200ctc_loss_bwd(const void* logits, int dtype, const int64_t* in_lens, int B, int T,
                                int V, int max_tgt_len, int blank, int zero_infinity,
                                const float* gscale, int gscale_stride, float gscale_mul,
                                void* grad, const void* ws_, size_t ws_bytes,
                                b200ctc_stream_t stream) {
  (void)in_lens; (void)zero_infinity;  // lengths and the zeroed flags live in the workspace
  B2_REQUIRE(B >= 0 && T >= 0 && V > 0 && max_tgt_len >= 0, "loss_bwd: bad shape");
  if (B == 0 || T == 0) return B200CTC_OK;
  B2_REQUIRE(logits && grad && ws_, "loss_bwd: null pointer");
  B2_REQUIRE(dtype >= 0 && dtype <= 2, "loss_bwd: unknown dtype %d", dtype);
  B2_REQUIRE(gscale_stride == 0 || gscale_stride == 1, "loss_bwd: gscale_stride must be 0 or 1");
  const LossLayout w = make_loss_layout(B, T, V, max_tgt_len);
  if (ws_bytes < w.total) {
    set_error("loss_bwd: workspace %zu < required %zu bytes", ws_bytes, w.total);
    return B200CTC_ERR_WORKSPACE;
  }
  const unsigned char* ws = static_cast<const unsigned char*>(ws_);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  int rc;
  prof_mark(3, 0, st);
  switch (dtype) {
    case B200CTC_F32:
      rc = dispatch_pass_b<float>(logits, grad, B, T, V, blank, w, ws, gscale, gscale_stride, gscale_mul, st);
      break;
    case B200CTC_F16:
      rc = dispatch_pass_b<__half>(logits, grad, B, T, V, blank, w, ws, gscale, gscale_stride, gscale_mul, st);
      break;
    default:
      rc = dispatch_pass_b<__nv_bfloat16>(logits, grad, B, T, V, blank, w, ws, gscale, gscale_stride, gscale_mul, st);
      break;
  }
  prof_mark(3, 1, st);
  return rc;
}
