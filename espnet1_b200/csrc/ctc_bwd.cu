// CTC gradient, row pass "B": the gradient w.r.t. the logits is produced in ONE
// streaming pass - each (b,t) row is read once (TMA bulk copy into a shared
// memory ring); its log-sum-exp, the soft-max and the lattice occupancies are
// combined on chip and the (B,T,V) gradient is written exactly once:
//     grad[b,t,v] = g_b * ( softmax(x)[v] - sum_{s: l_s = v} gamma_t(s) )
// (= autograd of log_softmax + ctc_loss, espnet2/asr/ctc.py:54-55; the
// analytically-zero term softmax * sum_v g_lp is dropped.)  The row's lse2 is
// also written out: in the eager mode (b200ctc_loss_fwd_grad) this is the only
// full read of the logits and the loss value is finished from it (nll_kernel).
#include "rows_common.cuh"

namespace b200ctc {

// WIDE (long lattices, S > 3*NT; TMA path only): the row's alpha, beta and the utterance's
// labels are staged in shared memory by the same TMA transaction as the logits row, so the
// occupancy loop (19 positions per thread at C4) reads shared memory instead of waiting
// on global loads one unrolled batch at a time.
// (plain __launch_bounds__(NT): an explicit minBlocks = 1 lets ptxas take 154 registers here and
//  the persistent grid then runs a second wave - measured: 414 us instead of 225 us)
template <typename T, int NT, int PER, bool TMA, bool WIDE>
__global__ void __launch_bounds__(NT) grad_rows_kernel(
    const T* __restrict__ logits, T* __restrict__ grad, int B, int Tfr, int V, int blank,
    const UttInfo* __restrict__ utt, const int* __restrict__ lab, const int* __restrict__ dup,
    int Lp, int Sp, int NWp, int NB, int wshift, float* __restrict__ lse2, const float* __restrict__ cmax,
    const float* __restrict__ alpha, const float* __restrict__ beta,
    const int* __restrict__ offA, const int* __restrict__ offB,
    const float* __restrict__ gscale, int gstride, float gmul, int nstage, int stage_stride,
    int pdl, const int* __restrict__ err, int l2mode) {
  constexpr int N = Elems<T>::N;
  constexpr int NWARP = NT / 32;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t bars[kMaxRowStages];
  __shared__ float wp[32];
  __shared__ float2 wms[2][32];  // per-warp (max, sum) of the row, double buffered
  unsigned char* stages = dsm;
  float* pbuf = reinterpret_cast<float*>(dsm + (size_t)nstage * stage_stride);  // [Lp]
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t nrows = (int64_t)B * Tfr;
  const uint32_t row_bytes = (uint32_t)V * sizeof(T);
  // WIDE stage: [logits row | alpha (Sp) | beta (Sp) | labels (Lp) | duplicate codes (Lp)]
  const uint32_t lat_off = (row_bytes + 127u) & ~127u;
  const uint32_t lat_bytes = 4u * (uint32_t)Sp, lab_bytes = 4u * (uint32_t)Lp;
  const int nchunk = V / N;
  const int ntail = V - nchunk * N;

  const uint64_t pol_stream = l2_policy((l2mode & 1) ? 1 : 0);
  RowCursor pc;
  pc.init(blockIdx.x, gridDim.x, Tfr);
  uint32_t np = 0;
  auto issue_next = [&]() {
    while (pc.g < nrows && (pc.t >= utt[pc.b].T || utt[pc.b].zeroed)) pc.next();
    if (pc.g < nrows) {
      const int st = np % nstage;
      unsigned char* dst = stages + (size_t)st * stage_stride;
      mbar_expect_tx(&bars[st], WIDE ? row_bytes + 2 * lat_bytes + 2 * lab_bytes : row_bytes);
      // the logits stream through once: first to be evicted, so that the lattices (written
      // evict_last by the forward kernels) are still in L2 when their row comes up
      bulk_g2s_hint(dst, logits + pc.g * V, row_bytes, &bars[st], pol_stream);
      if (WIDE) {
        bulk_g2s_hint(dst + lat_off, alpha + (size_t)pc.g * Sp, lat_bytes, &bars[st], pol_stream);
        bulk_g2s_hint(dst + lat_off + lat_bytes, beta + (size_t)pc.g * Sp, lat_bytes, &bars[st], pol_stream);
        bulk_g2s(dst + lat_off + 2 * lat_bytes, lab + (size_t)pc.b * Lp, lab_bytes, &bars[st]);
        bulk_g2s(dst + lat_off + 2 * lat_bytes + lab_bytes, dup + (size_t)pc.b * Lp, lab_bytes, &bars[st]);
      }
      ++np;
      pc.next();
    }
  };
  if (pdl == 0) griddep_launch();  // PDL: nll_kernel may queue up behind this grid
  if (TMA) {
    if (tid == 0) {
      for (int i = 0; i < nstage; ++i) mbar_init(&bars[i], 1);
      fence_mbar_init();
    }
    __syncthreads();
  }
  griddep_wait();  // utt / lattices / cmax come from the previous kernels
  if (TMA && tid == 0)
    for (int i = 0; i < nstage; ++i) issue_next();

  // a forward kernel of this call gave up waiting for another CTA (ctc_ws.cuh kWsErr*): the
  // lattices are incomplete, the gradient must not look plausible
  const bool ws_bad = __ldcg(err) != 0;
  uint32_t n = 0;
  RowCursor rc;
  rc.init(blockIdx.x, gridDim.x, Tfr);
  // the per-row constant is fetched one row ahead (its latency hides behind the current row)
  float c_nx = rc.g < nrows ? cmax[rc.g] : 0.f;
  for (; rc.g < nrows; rc.next()) {
    const int b = rc.b, t = rc.t;
    const int64_t g = rc.g;
    const float c2 = c_nx;
    if (g + rc.stride < nrows) c_nx = cmax[g + rc.stride];
    const UttInfo u = utt[b];
    T* gr = grad + g * V;
    // TMA instantiations are only launched with a 16-byte aligned gradient base (rows are then
    // aligned too: row_bytes % 16 == 0): the scalar store path does not exist in them
    const bool al16 = TMA ? true : (reinterpret_cast<uintptr_t>(gr) & 15) == 0;
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
    float qbase = -(float)u.lam;
    if (u.infinite || ws_bad) {  // inf loss kept (!zero_infinity): the reference's gradient is NaN
      gb = __int_as_float(0x7fc00000);
      qbase = gb;
    }
    const float* al = alpha + (size_t)g * Sp;
    const float* be = beta + (size_t)g * Sp;
    // frame offsets: one per block of kRenormEvery recursion steps (t / T_b-1-t)
    const int* oa = offA + ((size_t)b * NB + (t / kRenormEvery)) * NWp;
    const int* ob = offB + ((size_t)b * NB + ((u.T - 1 - t) / kRenormEvery)) * NWp;
    const int* lb = lab + (size_t)b * Lp;
    // log2 gamma_t(s) = alpha + beta - em + (FA_wA + FB_wB - aend) - lam   (ctc_ws.cuh)
    auto lattice_term = [&](int s) {
      const int sb = S - 1 - s;  // beta is stored in recursion order
      const int fo = oa[s >> wshift] + ob[sb >> wshift] - u.aend;
      return al[s] + be[sb] + ((float)fo + qbase);
    };
    constexpr int SM = 3;  // lattice positions per thread whose loads are issued before the wait
    float ab[SM];
    if (!WIDE) {
#pragma unroll
      for (int j = 0; j < SM; ++j) {
        const int s = tid + j * NT;
        ab[j] = s < S ? lattice_term(s) : kNegInf;
      }
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
    // ---- wide lattices (S > NT*SM, long-form utterances): the occupancies of the positions
    // beyond the prefetched ones are computed HERE, before the row occupies 40 registers,
    // with kW positions' alpha / beta / offset loads in flight per thread (this loop is
    // bound by their latency); no iterations at C2
    float blank_part = 0.f;
    {
      constexpr int kW = 8;
      const float* als = reinterpret_cast<const float*>(stg + lat_off);  // WIDE: staged copies
      const float* bes = als + Sp;
      const int* lbs = reinterpret_cast<const int*>(bes + Sp);
      for (int s0 = WIDE ? 0 : NT * SM; s0 < S; s0 += NT * kW) {
        float abw[kW];
#pragma unroll
        for (int j = 0; j < kW; ++j) {
          const int s = s0 + tid + j * NT;
          if (WIDE) {
            const int sc = s < S ? s : 0, sb = S - 1 - sc;
            const int fo = oa[sc >> wshift] + ob[sb >> wshift] - u.aend;
            abw[j] = s < S ? als[sc] + bes[sb] + ((float)fo + qbase) : kNegInf;
          } else {
            abw[j] = s < S ? lattice_term(s) : kNegInf;
          }
        }
#pragma unroll
        for (int j = 0; j < kW; ++j) {
          const int s = s0 + tid + j * NT;
          if (s < S) {
            const bool odd = s & 1;
            const int k = odd ? (s + 1) >> 1 : 0;
            const int v = odd ? (WIDE ? lbs[k] : lb[k]) : blank;
            B2_DEV_CHECK(v >= 0 && v < V);
            const float em = fmaf(to_f32<T>(xs[v]), kLog2e, -c2);
            const float p = ex2(abw[j] - em);
            if (odd) pbuf[k] = p; else blank_part += p;
          }
        }
      }
    }
    // ---- phase 0: whole row -> registers; thread-local max and sum of 2^(x2 - max)
    float x[PER][N];
    float m = kNegInf;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = tid + j * NT;
      if (i < nchunk) {
        unpack16<T>(reinterpret_cast<const uint4*>(stg)[i], x[j]);
#pragma unroll
        for (int e = 0; e < N; ++e) m = fmaxf(m, x[j][e]);
      }
    }
    float xt = kNegInf;
    if (tid < ntail) {
      xt = to_f32<T>(xs[nchunk * N + tid]);
      m = fmaxf(m, xt);
    }
    m = warp_max(m);  // warp-level maximum keeps the later combine short
    const float mw2 = ((m == kNegInf) ? 0.f : m) * kLog2e;
    float sum = 0.f;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = tid + j * NT;
      if (i < nchunk) {
#pragma unroll
        for (int e = 0; e < N; ++e) {
          x[j][e] = ex2(fmaf(x[j][e], kLog2e, -mw2));  // kept: reused for the output
          sum += x[j][e];
        }
      }
    }
    if (tid < ntail) {
      xt = ex2(fmaf(xt, kLog2e, -mw2));
      sum += xt;
    }
    sum = warp_sum(sum);
    if (lane == 0) wms[n & 1][wid] = make_float2(mw2, sum);
    // ---- phase 1a: occupancies gamma_t(s) (independent of the row's normaliser)
    if (!WIDE) {
#pragma unroll
      for (int j = 0; j < SM; ++j) {
        const int s = tid + j * NT;
        if (s < S) {
          const float abv = ab[j];
          const bool odd = s & 1;
          const int k = odd ? (s + 1) >> 1 : 0;
          const int v = odd ? lb[k] : blank;
          B2_DEV_CHECK(v >= 0 && v < V);
          const float em = fmaf(to_f32<T>(xs[v]), kLog2e, -c2);
          const float p = ex2(abv - em);
          if (odd) pbuf[k] = p; else blank_part += p;
        }
      }
    }
    blank_part = warp_sum(blank_part);
    if (lane == 0) wp[wid] = blank_part;
    __syncthreads();  // wms, pbuf, wp visible
    // this row's alpha / beta were read for the last time (every thread's loads completed
    // before the barrier): drop their lines from L2 instead of writing 77 MB back to DRAM.
    // (Sp is a multiple of 32 floats: rows do not share 128-byte lines.)
    if (!WIDE && (l2mode & 2)) {
      const int nl = (4 * Sp) >> 7;
      if (tid < nl) l2_discard_128(reinterpret_cast<const char*>(al) + 128 * tid);
      else if (tid < 2 * nl) l2_discard_128(reinterpret_cast<const char*>(be) + 128 * (tid - nl));
    }
    // ---- row normaliser: combine the per-warp (max, sum) pairs
    float M2 = kNegInf;
#pragma unroll
    for (int w2 = 0; w2 < NWARP; ++w2) M2 = fmaxf(M2, wms[n & 1][w2].x);
    float tot = 0.f;
#pragma unroll
    for (int w2 = 0; w2 < NWARP; ++w2) tot += wms[n & 1][w2].y * ex2(wms[n & 1][w2].x - M2);
    const float l2 = M2 + log2f(tot);
    if (tid == 0) lse2[g] = l2;
    const float scale = gb * ex2(mw2 - l2);  // g_b * 2^(warp max - lse2)
    // ---- phase 1b: dense part, g_b * softmax(x), straight from registers
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = tid + j * NT;
      if (i < nchunk) {
#pragma unroll
        for (int e = 0; e < N; ++e) x[j][e] *= scale;
        if (al16) store16<T>(gr + i * N, x[j]);
        else
#pragma unroll
          for (int e = 0; e < N; ++e) gr[i * N + e] = from_f32<T>(x[j][e]);
      }
    }
    if (tid < ntail) gr[nchunk * N + tid] = from_f32<T>(xt * scale);
    __syncthreads();  // dense stores are ordered before the column fix-ups below
    // ---- phase 2: the <= L+1 columns that carry occupancy are overwritten by
    // their single owner thread (deterministic for repeated labels)
    const int* db = WIDE ? reinterpret_cast<const int*>(stg + lat_off + 2 * lat_bytes + lab_bytes)
                         : dup + (size_t)b * Lp;
    const int* lb2 = WIDE ? reinterpret_cast<const int*>(stg + lat_off + 2 * lat_bytes) : lb;
    for (int k0 = tid + 1; k0 <= L; k0 += NT * 4) {  // 4 positions' codes / labels in flight
      int code[4], vv[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int k = k0 + j * NT;
        code[j] = k <= L ? db[k] : 0;
        vv[j] = k <= L ? lb2[k] : 0;
      }
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        if (code[j] & kDupFirst) {
          float acc = pbuf[k0 + j * NT];
          for (int nx = code[j] & kDupMask; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
          const float sm = ex2(fmaf(to_f32<T>(xs[vv[j]]), kLog2e, -l2));
          gr[vv[j]] = from_f32<T>(gb * (sm - acc));
        }
      }
    }
    if (tid == NT - 1) {
      float acc = 0.f;
      for (int w2 = 0; w2 < NWARP; ++w2) acc += wp[w2];
      for (int nx = u.blank_head; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
      const float sm = ex2(fmaf(to_f32<T>(xs[blank]), kLog2e, -l2));
      gr[blank] = from_f32<T>(gb * (sm - acc));
    }
    __syncthreads();  // stage, pbuf and wp are free again
    if (TMA && tid == 0) issue_next();
    ++n;
  }
  if (pdl == 1) griddep_launch();
}

#ifdef B200CTC_TUNING  // measured loser (profiles/README.md round 2): tuning builds only
// ---- DIRECT variant: one CTA per (b,t) row, not persistent, no TMA / mbarrier: the row goes
// global -> registers with plain 128-bit loads (all in flight at once) and is ALSO parked in
// shared memory, so that the ~450 label-column reads of the occupancy and fix-up phases are LDS
// (as scattered global loads they cost 32 L1 tag look-ups per warp instruction: +78 us, measured).
// The hardware block scheduler and the other resident CTAs of the SM are the pipeline.
// mode bit 0 (timing probe): dense part only.
template <typename T, int NT, int PER>
__global__ void __launch_bounds__(NT) grad_rows_direct_kernel(
    const T* __restrict__ logits, T* __restrict__ grad, int Tfr, int V, int blank,
    const UttInfo* __restrict__ utt, const int* __restrict__ lab, const int* __restrict__ dup,
    int Lp, int Sp, int NWp, int NB, int wshift, float* __restrict__ lse2, const float* __restrict__ cmax,
    const float* __restrict__ alpha, const float* __restrict__ beta,
    const int* __restrict__ offA, const int* __restrict__ offB,
    const float* __restrict__ gscale, int gstride, float gmul, const int* __restrict__ err, int mode) {
  constexpr int N = Elems<T>::N;
  constexpr int NWARP = NT / 32;
  extern __shared__ __align__(16) unsigned char dsm[];  // [row | pbuf (Lp floats)]
  __shared__ float wp[32];
  __shared__ float2 wms[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t g = blockIdx.x;
  const int b = (int)(g / Tfr), t = (int)(g - (int64_t)b * Tfr);
  const int nchunk = V / N;  // the launcher guarantees V % N == 0 and 16-byte aligned rows
  const T* xr = logits + g * V;
  T* gr = grad + g * V;
  uint4* srow = reinterpret_cast<uint4*>(dsm);
  const T* xs = reinterpret_cast<const T*>(dsm);
  float* pbuf = reinterpret_cast<float*>(dsm + (((size_t)V * sizeof(T) + 127) & ~(size_t)127));
  griddep_wait();
  const UttInfo u = utt[b];
  if (t >= u.T || u.zeroed) {  // padded frame or zeroed utterance: gradient is 0
    float z[N];
#pragma unroll
    for (int e = 0; e < N; ++e) z[e] = 0.f;
    for (int i = tid; i < nchunk; i += NT) store16<T>(gr + i * N, z);
    return;
  }
  // the row: all loads in flight first
  uint4 raw[PER];
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = tid + j * NT;
    if (i < nchunk) raw[j] = __ldg(reinterpret_cast<const uint4*>(xr) + i);
  }
  const bool ws_bad = __ldcg(err) != 0;
  const float c2 = cmax[g];
  const int L = (mode & 1) ? 0 : u.L, S = (mode & 1) ? 0 : 2 * L + 1;
  float gb = gmul * (gscale ? gscale[(size_t)b * gstride] : 1.f);
  float qbase = -(float)u.lam;
  if (u.infinite || ws_bad) {
    gb = __int_as_float(0x7fc00000);
    qbase = gb;
  }
  const float* al = alpha + (size_t)g * Sp;
  const float* be = beta + (size_t)g * Sp;
  const int* oa = offA + ((size_t)b * NB + (t / kRenormEvery)) * NWp;
  const int* ob = offB + ((size_t)b * NB + ((u.T - 1 - t) / kRenormEvery)) * NWp;
  const int* lb = lab + (size_t)b * Lp;
  constexpr int SM = NT >= 256 ? 2 : 3;  // lattice positions per thread with loads in flight early
  float ab[SM];
  int vcol[SM];
#pragma unroll
  for (int j = 0; j < SM; ++j) {
    const int s = tid + j * NT;
    const int sc = s < S ? s : 0, sb = S - 1 - sc;
    const int fo = oa[sc >> wshift] + ob[sb >> wshift] - u.aend;
    ab[j] = s < S ? al[sc] + be[sb] + ((float)fo + qbase) : kNegInf;
    vcol[j] = (sc & 1) ? lb[(sc + 1) >> 1] : blank;
  }
  // ---- phase 0: park the row in shared memory; thread-local max and sum of 2^(x2 - max)
  float x[PER][N];
  float m = kNegInf;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = tid + j * NT;
    if (i < nchunk) {
      srow[i] = raw[j];
      unpack16<T>(raw[j], x[j]);
#pragma unroll
      for (int e = 0; e < N; ++e) m = fmaxf(m, x[j][e]);
    }
  }
  m = warp_max(m);
  const float mw2 = ((m == kNegInf) ? 0.f : m) * kLog2e;
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = tid + j * NT;
    if (i < nchunk) {
#pragma unroll
      for (int e = 0; e < N; ++e) {
        x[j][e] = ex2(fmaf(x[j][e], kLog2e, -mw2));
        sum += x[j][e];
      }
    }
  }
  sum = warp_sum(sum);
  if (lane == 0) wms[wid] = make_float2(mw2, sum);
  __syncthreads();  // srow and wms visible
  // ---- phase 1a: occupancies gamma_t(s)
  float blank_part = 0.f;
#pragma unroll
  for (int j = 0; j < SM; ++j) {
    const int s = tid + j * NT;
    if (s < S) {
      const float em = fmaf(to_f32<T>(xs[vcol[j]]), kLog2e, -c2);
      const float p = ex2(ab[j] - em);
      if (s & 1) pbuf[(s + 1) >> 1] = p; else blank_part += p;
    }
  }
  for (int s0 = NT * SM; s0 < S; s0 += NT) {  // wider lattices
    const int s = s0 + tid;
    if (s < S) {
      const int sb = S - 1 - s;
      const int fo = oa[s >> wshift] + ob[sb >> wshift] - u.aend;
      const float abv = al[s] + be[sb] + ((float)fo + qbase);
      const bool odd = s & 1;
      const int k = odd ? (s + 1) >> 1 : 0;
      const int v = odd ? lb[k] : blank;
      const float em = fmaf(to_f32<T>(xs[v]), kLog2e, -c2);
      const float p = ex2(abv - em);
      if (odd) pbuf[k] = p; else blank_part += p;
    }
  }
  blank_part = warp_sum(blank_part);
  if (lane == 0) wp[wid] = blank_part;
  // ---- row normaliser and the dense part, straight from registers
  float M2 = kNegInf;
#pragma unroll
  for (int w2 = 0; w2 < NWARP; ++w2) M2 = fmaxf(M2, wms[w2].x);
  float tot = 0.f;
#pragma unroll
  for (int w2 = 0; w2 < NWARP; ++w2) tot += wms[w2].y * ex2(wms[w2].x - M2);
  const float l2 = M2 + log2f(tot);
  if (tid == 0) lse2[g] = l2;
  const float scale = gb * ex2(mw2 - l2);
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = tid + j * NT;
    if (i < nchunk) {
#pragma unroll
      for (int e = 0; e < N; ++e) x[j][e] *= scale;
      store16<T>(gr + i * N, x[j]);
    }
  }
  __syncthreads();  // pbuf / wp visible; dense stores ordered before the column fix-ups
  if (mode & 1) return;
  // ---- phase 2: label columns, single owner thread per vocabulary column
  const int* db = dup + (size_t)b * Lp;
  for (int k0 = tid + 1; k0 <= L; k0 += NT * 2) {
    int code[2], vv[2];
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      const int k = k0 + j * NT;
      code[j] = k <= L ? db[k] : 0;
      vv[j] = k <= L ? lb[k] : 0;
    }
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      if (code[j] & kDupFirst) {
        float acc = pbuf[k0 + j * NT];
        for (int nx = code[j] & kDupMask; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
        const float sm = ex2(fmaf(to_f32<T>(xs[vv[j]]), kLog2e, -l2));
        gr[vv[j]] = from_f32<T>(gb * (sm - acc));
      }
    }
  }
  if (tid == NT - 1) {
    float acc = 0.f;
    for (int w2 = 0; w2 < NWARP; ++w2) acc += wp[w2];
    for (int nx = u.blank_head; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
    const float sm = ex2(fmaf(to_f32<T>(xs[blank]), kLog2e, -l2));
    gr[blank] = from_f32<T>(gb * (sm - acc));
  }
}

#endif  // B200CTC_TUNING

// PERX: 16-byte chunks per thread (0: the default 40 elements per thread)
template <typename T, int NT, int PERX = 0>
static int launch_pass_b(const void* logits, void* grad, int B, int Tfr, int V, int blank,
                         const LossLayout& w, unsigned char* ws, const float* gscale,
                         int gstride, float gmul, bool last_use, cudaStream_t st) {
  constexpr int PER = PERX ? PERX : 40 / Elems<T>::N;
  const size_t extra = align_up(sizeof(float) * (size_t)w.Lp, 128);
  RowLaunchCfg c = row_launch_cfg(logits, (size_t)V * sizeof(T), extra, NT, "B200CTC_B");
  if ((reinterpret_cast<uintptr_t>(grad) & 15) != 0) {  // unaligned gradient: the generic kernel
    c.tma = false;
    c.nstage = 1;
  }
  // WIDE staging: alpha / beta / labels / duplicate codes of the row travel with the logits row
  // in one TMA transaction, the row loop then has no dependent global loads.
  //  * long lattices (S > 3 NT): needed - the occupancy loop was latency-bound without it (C4);
  //  * short lattices at 128 threads per row (C2): the instantiation needs 72 registers instead
  //    of 80, so SEVEN CTAs per SM fit (24 KB stages): row pass 238 -> 225 us at C2, measured.
  //    Taken only if 7 CTAs are really resident (occupancy query), else the 6-CTA kernel.
  static const int wide_on = tuning_int("B200CTC_B_WIDE", 1);
  const size_t lat_stage = align_up(4 * (2 * (size_t)w.Sp + 2 * (size_t)w.Lp), 128);
  const bool wide_needed = 2 * w.Lmax + 1 > 3 * NT;
  bool wide = wide_on && c.tma && (wide_needed || NT == 128 || (NT == 256 && PERX));
  if (wide) {
    const size_t row_al = align_up((size_t)V * sizeof(T), 128);
    // (256 threads per row - long lattices, see dispatch_pass_b: 61 registers, 4 single-stage
    //  CTAs = 32 warps per SM; measured at C4: 831 us vs 989 us at 128 threads, 947 us with 3 CTAs)
    const int want = wide_needed ? (NT == 256 ? 4 : 0) : 7;
    const RowLaunchCfg cw = row_launch_cfg(logits, row_al + lat_stage, extra, NT, "B200CTC_B", want,
                                           (wide_needed && NT == 256) ? 1 : 0);
    bool ok = cw.ctas_per_sm * ((size_t)cw.nstage * cw.stride + extra + 1024) <= 220 * 1024 &&
              cw.ctas_per_sm >= (wide_needed ? 3 : (NT == 128 ? 7 : 4));
    if (ok && !wide_needed) {  // all 7 must be resident at once: a persistent grid must not run a second wave
      static thread_local int occ_cached = -1;
      static thread_local size_t occ_smem = 0;
      const size_t smem_w = (size_t)cw.nstage * cw.stride + extra;
      if (occ_cached < 0 || occ_smem != smem_w) {
        auto kw = grad_rows_kernel<T, NT, PER, true, true>;
        cudaFuncSetAttribute(kw, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_w);
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kw, NT, smem_w) != cudaSuccess) occ = 0;
        cudaGetLastError();
        occ_cached = occ;
        occ_smem = smem_w;
      }
      ok = occ_cached >= cw.ctas_per_sm;
    }
    if (ok) {
      c.nstage = cw.nstage; c.ctas_per_sm = cw.ctas_per_sm; c.stride = cw.stride;
    } else {
      wide = false;
    }
  }
  B2_REQUIRE((size_t)c.nstage * c.stride + extra <= 226 * 1024,
             "ctc_loss grad: V=%d too large for the shared-memory row path", V);
  const size_t smem = (size_t)c.nstage * c.stride + extra;
  const int64_t rows = (int64_t)B * Tfr;
  int64_t grid = (int64_t)sm_count() * c.ctas_per_sm;
  if (grid > rows) grid = rows;
  int wshift = 0;
  while ((1 << wshift) < 32 * w.R) ++wshift;
  // bit 0: logits evict_first; bit 1: discard the row's alpha / beta lines after use
  static const int l2tune = tuning_int("B200CTC_ROWS_L2", 0);
  // (deferred mode keeps the lattices for a possible second backward: no discard there)
  const int l2mode = last_use ? l2tune : (l2tune & ~2);
#ifdef B200CTC_TUNING
  // experiment: the direct row kernel (one CTA per row)
  static const int direct_on = tuning_int("B200CTC_ROWS_DIRECT", 0);
  if (direct_on && !wide && c.tma && (reinterpret_cast<uintptr_t>(grad) & 15) == 0 && rows < (1ll << 31)) {
    constexpr int NT2 = NT * 2 > 1024 ? 1024 : NT * 2;   // twice the threads, half the row per thread
    constexpr int PER2 = NT2 == NT ? PER : PER / 2;
    auto k = grad_rows_direct_kernel<T, NT2, PER2>;
    const size_t dsmem = align_up((size_t)V * sizeof(T), 128) + extra;
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dsmem);
    launch_pdl(k, dim3((unsigned)rows), dim3(NT2), dsmem, st, 1, static_cast<const T*>(logits),
               static_cast<T*>(grad), Tfr, V, blank, reinterpret_cast<const UttInfo*>(ws + w.utt),
               reinterpret_cast<const int*>(ws + w.lab), reinterpret_cast<const int*>(ws + w.dup),
               w.Lp, w.Sp, w.NW, w.NB, wshift, reinterpret_cast<float*>(ws + w.lse2),
               reinterpret_cast<const float*>(ws + w.cmax), reinterpret_cast<const float*>(ws + w.alpha),
               reinterpret_cast<const float*>(ws + w.beta), reinterpret_cast<const int*>(ws + w.offA),
               reinterpret_cast<const int*>(ws + w.offB), gscale, gstride, gmul,
               reinterpret_cast<const int*>(ws + w.err), tuning_int("B200CTC_DIRECT_MODE", 0));
    count_launch();
    return check_launch("ctc grad_rows (direct)");
  }
#endif
  auto go = [&](auto k) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    launch_pdl(
        // (no PDL here: a persistent grid must not become resident piecemeal beside
        //  the still-running lattice - late CTAs would form a second wave)
        k, dim3((unsigned)grid), dim3(NT), smem, st, 1,
        static_cast<const T*>(logits), static_cast<T*>(grad), B, Tfr, V, blank,
        reinterpret_cast<const UttInfo*>(ws + w.utt), reinterpret_cast<const int*>(ws + w.lab),
        reinterpret_cast<const int*>(ws + w.dup), w.Lp, w.Sp, w.NW, w.NB, wshift,
        reinterpret_cast<float*>(ws + w.lse2), reinterpret_cast<const float*>(ws + w.cmax),
        reinterpret_cast<const float*>(ws + w.alpha), reinterpret_cast<const float*>(ws + w.beta),
        reinterpret_cast<const int*>(ws + w.offA), reinterpret_cast<const int*>(ws + w.offB),
        gscale, gstride, gmul, c.nstage, c.stride, pdl_mode(),
        reinterpret_cast<const int*>(ws + w.err), l2mode);
  };
  if (wide) go(grad_rows_kernel<T, NT, PER, true, true>);
  else if (c.tma) go(grad_rows_kernel<T, NT, PER, true, false>);
  else go(grad_rows_kernel<T, NT, PER, false, false>);
  count_launch();
  return check_launch("ctc grad_rows");
}

template <typename T>
static int dispatch_pass_b(const void* logits, void* grad, int B, int Tfr, int V, int blank,
                           const LossLayout& w, unsigned char* ws, const float* gscale,
                           int gstride, float gmul, bool last_use, cudaStream_t st) {
  const int need = (V + 39) / 40;
  // long lattices (C4: 19 lattice positions and ~10 label columns per thread at 128 threads):
  // 256 threads per row with 20 elements each halve the serial work per thread
  static const int wide256 = tuning_int("B200CTC_B_WIDE256", 1);  // (2: also short lattices, tuning builds)
  if (wide256 && need <= 128 && (2 * w.Lmax + 1 > 3 * 128 || wide256 >= 2))
    return launch_pass_b<T, 256, 20 / Elems<T>::N>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, last_use, st);
  if (need <= 128) return launch_pass_b<T, 128>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, last_use, st);
  if (need <= 256) return launch_pass_b<T, 256>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, last_use, st);
  if (need <= 512) return launch_pass_b<T, 512>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, last_use, st);
  if (need <= 1024) return launch_pass_b<T, 1024>(logits, grad, B, Tfr, V, blank, w, ws, gscale, gstride, gmul, last_use, st);
  set_error("ctc_loss grad: V=%d exceeds the supported row size", V);
  return B200CTC_ERR_ARG;
}

int launch_grad_rows(const void* logits, int dtype, void* grad, int B, int T, int V, int blank,
                     const LossLayout& w, unsigned char* ws, const float* gscale, int gstride,
                     float gmul, bool last_use, cudaStream_t st) {
  switch (dtype) {
    case B200CTC_F32:
      return dispatch_pass_b<float>(logits, grad, B, T, V, blank, w, ws, gscale, gstride, gmul, last_use, st);
    case B200CTC_F16:
      return dispatch_pass_b<__half>(logits, grad, B, T, V, blank, w, ws, gscale, gstride, gmul, last_use, st);
    default:
      return dispatch_pass_b<__nv_bfloat16>(logits, grad, B, T, V, blank, w, ws, gscale, gstride,
                                            gmul, last_use, st);
  }
}

// ---- lazy rescale of an eager gradient -------------------------------------
// The eager forward baked `hint` into the gradient; when the real upstream
// gradient equals the hint (loss.sum()/B with grad_output 1) every CTA reads
// one scalar and exits - no pass over the (B,T,V) tensor.
template <typename T>
__global__ void __launch_bounds__(256) rescale_kernel(T* __restrict__ grad, int64_t per_utt,
                                                      const float* __restrict__ gscale,
                                                      int gstride, float gmul, float hint0,
                                                      const float* __restrict__ hints) {
  constexpr int N = Elems<T>::N;
  const int b = blockIdx.y;
  griddep_wait();  // (programmatic launch: gscale / grad come from the previous kernels)
  const float gv = gscale[(size_t)b * gstride] * gmul;
  const float hint = hints ? hints[b] : hint0;
  if (gv == hint) return;
  const float ratio = gv / hint;
  T* base = grad + (size_t)b * per_utt;
  const int64_t nvec = per_utt / N;
  const bool al16 = (reinterpret_cast<uintptr_t>(base) & 15) == 0;
  if (al16) {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * 256) {
      float f[N];
      unpack16<T>(reinterpret_cast<const uint4*>(base)[i], f);
#pragma unroll
      for (int e = 0; e < N; ++e) f[e] *= ratio;
      store16<T>(base + i * N, f);
    }
    for (int64_t i = nvec * N + (int64_t)blockIdx.x * 256 + threadIdx.x; i < per_utt;
         i += (int64_t)gridDim.x * 256)
      base[i] = from_f32<T>(to_f32<T>(base[i]) * ratio);
  } else {
    for (int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x; i < per_utt; i += (int64_t)gridDim.x * 256)
      base[i] = from_f32<T>(to_f32<T>(base[i]) * ratio);
  }
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" int b200ctc_loss_bwd(const void* logits, int dtype, const int64_t* in_lens, int B, int T,
                                int V, int max_tgt_len, int blank, int zero_infinity,
                                const float* gscale, int gscale_stride, float gscale_mul,
                                void* grad, void* ws_, size_t ws_bytes,
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
  unsigned char* ws = static_cast<unsigned char*>(ws_);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  prof_mark(3, 0, st);
  const int rc = launch_grad_rows(logits, dtype, grad, B, T, V, blank, w, ws, gscale,
                                  gscale_stride, gscale_mul, false, st);
  prof_mark(3, 1, st);
  return rc;
}

static int grad_rescale_impl(void* grad, int dtype, const float* gscale, int gscale_stride,
                             float gscale_mul, float grad_hint, const float* utt_hints, int B, int T,
                             int V, b200ctc_stream_t stream) {
  B2_REQUIRE(B >= 0 && T >= 0 && V > 0, "grad_rescale: bad shape");
  if (B == 0 || T == 0) return B200CTC_OK;
  B2_REQUIRE(grad && gscale, "grad_rescale: null pointer");
  B2_REQUIRE(dtype >= 0 && dtype <= 2, "grad_rescale: unknown dtype %d", dtype);
  B2_REQUIRE(gscale_stride == 0 || gscale_stride == 1, "grad_rescale: gscale_stride must be 0 or 1");
  B2_REQUIRE(utt_hints || (grad_hint != 0.f && grad_hint == grad_hint),
             "grad_rescale: grad_hint must be non-zero");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int64_t per_utt = (int64_t)T * V;
  int gx = (2 * sm_count() + B - 1) / B;
  if (gx < 1) gx = 1;
  dim3 grid(gx, B);
  prof_mark(6, 0, st);
  switch (dtype) {
    case B200CTC_F32:
      launch_pdl(rescale_kernel<float>, grid, dim3(256), 0, st, 4, static_cast<float*>(grad), per_utt,
                 gscale, gscale_stride, gscale_mul, grad_hint, utt_hints);
      break;
    case B200CTC_F16:
      launch_pdl(rescale_kernel<__half>, grid, dim3(256), 0, st, 4, static_cast<__half*>(grad), per_utt,
                 gscale, gscale_stride, gscale_mul, grad_hint, utt_hints);
      break;
    default:
      launch_pdl(rescale_kernel<__nv_bfloat16>, grid, dim3(256), 0, st, 4,
                 static_cast<__nv_bfloat16*>(grad), per_utt, gscale, gscale_stride, gscale_mul, grad_hint,
                 utt_hints);
      break;
  }
  prof_mark(6, 1, st);
  count_launch();
  return check_launch("ctc grad_rescale");
}

extern "C" int b200ctc_grad_rescale(void* grad, int dtype, const float* gscale, int gscale_stride,
                                    float gscale_mul, float grad_hint, int B, int T, int V,
                                    b200ctc_stream_t stream) {
  return grad_rescale_impl(grad, dtype, gscale, gscale_stride, gscale_mul, grad_hint, nullptr, B, T, V,
                           stream);
}

extern "C" int b200ctc_grad_rescale_hints(void* grad, int dtype, const float* gscale,
                                          int gscale_stride, float gscale_mul,
                                          const float* utt_hints, int B, int T, int V,
                                          b200ctc_stream_t stream) {
  B2_REQUIRE(utt_hints || B == 0, "grad_rescale_hints: null utt_hints");
  return grad_rescale_impl(grad, dtype, gscale, gscale_stride, gscale_mul, 0.f, utt_hints, B, T, V,
                           stream);
}
