// CTC forward: (1) label prep, (2) fused row pass "A": log-sum-exp over the
// vocabulary + gather of the lattice emissions, streamed through shared memory
// by the TMA engine (1-D bulk copies, mbarrier ring), (3) alpha/beta lattice
// recursion (ctc_lattice.cuh), one CTA per utterance and direction.
#include "ctc_lattice.cuh"
#include "rows_common.cuh"

namespace b200ctc {

// ------------------------------------------------------------------ prep
// int64 labels -> per-utterance int32 rows (k = 0 is the blank), duplicate
// chains (so that repeated labels are accumulated deterministically by one
// owner thread in the gradient pass), clamped lengths.  Labels are staged in
// shared memory: the O(L^2) duplicate search then runs on broadcast LDS.
// Executed by a whole CTA; `publish` = also write lab/dup/utt for the later
// kernels (one CTA per utterance does, the others only need `slab`).
struct PrepArgs {
  const int64_t* targets;
  const int64_t* offs;  // may be NULL: prefix sum of tgt_lens
  const int64_t* in_lens;
  const int64_t* tgt_lens;
  int T, V, Lmax, Lp, blank;
  UttInfo* utt;
  int* lab;
  int* dup;
  int* counter;
  int* err;  // sticky error word, cleared here; NULL: the caller's memset already did
};

// `table` (optional, 2*V ints of shared memory: first occurrence and count per vocabulary entry):
// with it only the labels that really repeat pay for a scan - at C4 (L = 1200, V = 5000) ~20 % of
// the positions instead of two full scans by every position (55 -> 15 us).  Without it (the prep
// fused into the sparse gather, where L is small and shared memory is occupancy) both scans run.
__device__ __forceinline__ void prep_utterance(const PrepArgs& P, int b, int* slab, bool publish,
                                               int& Tb_out, int& L_out, int* table = nullptr) {
  __shared__ int s_bad, s_head;
  __shared__ long long s_base;
  const int NT = blockDim.x, tid = threadIdx.x;
  if (publish && b == 0 && tid == 0) {
    *P.counter = 0;  // nll_kernel's "last CTA" ticket
    if (P.err) *P.err = 0;
  }
  const long long L64 = P.tgt_lens[b], T64 = P.in_lens[b];
  const int L = (int)(L64 < 0 ? 0 : (L64 > P.Lmax ? P.Lmax : L64));
  const int Tb = (int)(T64 < 0 ? 0 : (T64 > P.T ? P.T : T64));
  Tb_out = Tb;
  L_out = L;
  if (tid == 0) { s_bad = 0; s_head = 0x7fffffff; s_base = 0; }
  __syncthreads();
  if (P.offs) {
    if (tid == 0) s_base = P.offs[b];
  } else {
    long long part = 0;
    for (int i = tid; i < b; i += NT) part += P.tgt_lens[i] < 0 ? 0 : P.tgt_lens[i];
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if ((tid & 31) == 0 && part)
      atomicAdd(reinterpret_cast<unsigned long long*>(&s_base), (unsigned long long)part);
  }
  __syncthreads();
  const int64_t base = s_base;
  int* lb = P.lab + (size_t)b * P.Lp;
  for (int k = tid; k < P.Lp; k += NT) {
    int v = P.blank;
    if (k >= 1 && k <= L) {
      long long t = P.targets[base + k - 1];
      if (t < 0 || t >= P.V) { s_bad = 1; t = t < 0 ? 0 : P.V - 1; }
      v = (int)t;
    }
    slab[k] = v;
    if (publish) lb[k] = v;
  }
  __syncthreads();
  if (!publish) return;
  int* tfirst = table;
  int* tcount = table ? table + P.V : nullptr;
  if (table) {
    for (int v = tid; v < P.V; v += NT) { tfirst[v] = 0x7fffffff; tcount[v] = 0; }
    __syncthreads();
    for (int k = tid + 1; k <= L; k += NT) {
      atomicMin(&tfirst[slab[k]], k);
      atomicAdd(&tcount[slab[k]], 1);
    }
    __syncthreads();
  }
  int* db = P.dup + (size_t)b * P.Lp;
  for (int k = tid; k < P.Lp; k += NT) {
    int code = 0;
    if (k >= 1 && k <= L) {
      const int v = slab[k];
      // branch-free scans: the shared-memory loads of a scan are independent and pipeline
      // (an early exit makes every iteration wait for its load: 95 us at L = 1200)
      bool seen = false;
      int next = 0x7fffffff;
      if (table) {
        seen = tfirst[v] != k;
        if (tcount[v] > 1) {
#pragma unroll 4
          for (int j = k + 1; j <= L; ++j) next = min(next, slab[j] == v ? j : 0x7fffffff);
        }
      } else {
#pragma unroll 4
        for (int j = 1; j < k; ++j) seen |= slab[j] == v;
#pragma unroll 4
        for (int j = k + 1; j <= L; ++j) next = min(next, slab[j] == v ? j : 0x7fffffff);
      }
      const bool first = (v != P.blank) && !seen;
      if (next == 0x7fffffff) next = 0;
      code = next | (first ? kDupFirst : 0);
      if (v == P.blank) atomicMin(&s_head, k);
    }
    db[k] = code;
  }
  __syncthreads();
  if (tid == 0) {
    UttInfo u;
    u.lam = 0; u.aend = 0;
    u.T = Tb; u.L = L; u.zeroed = 0; u.infinite = 0;
    u.blank_head = s_head == 0x7fffffff ? 0 : s_head;
    u.bad_label = s_bad; u.pad = 0;
    P.utt[b] = u;
  }
}

__global__ void __launch_bounds__(1024) prep_kernel(PrepArgs P, int use_table) {
  extern __shared__ int slab[];  // [Lp] (+ [2*V] first-occurrence / count table)
  int Tb, L;
  prep_utterance(P, blockIdx.x, slab, true, Tb, L, use_table ? slab + P.Lp : nullptr);
}

// ------------------------------------------------- emission gather (sparse)
// The lattice needs only the logits at the utterance's label columns (and the
// blank): (L+1) of the V entries of every row, i.e. ~1/4 of the row's 32-byte
// sectors at C2 instead of a full streaming read.  A warp handles kGatherRows
// (b,t) rows at once (all their loads in flight together);
// em[t][k] = logit[t][label_k]*log2(e) - c[t] with c[t] = max_k (ctc_ws.cuh).
// The label preparation is fused in: every CTA stages its utterance's labels
// itself, and the first CTA of each utterance publishes lab/dup/utt.
// one 4-byte (2-byte) read per 128-byte line: ask L2 for 64 bytes around it, not the line
#ifndef B200CTC_GATHER_LTC
#define B200CTC_GATHER_LTC 64
#endif
#ifndef B200CTC_GATHER_POL
#define B200CTC_GATHER_POL 0
#endif
template <typename T>
__device__ __forceinline__ float ldg_gather(const T* p) {
#if B200CTC_GATHER_LTC == 64 && B200CTC_GATHER_POL != 0
  // + L2 evict_first: the gather's lines are dead on arrival; do not let them push out the
  // previous kernel's dirty lines (their write-back then happens under the lattice, HBM idle)
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  if (sizeof(T) == 4) {
    float v;
    asm volatile("ld.global.nc.L2::cache_hint.L2::64B.f32 %0, [%1], %2;" : "=f"(v) : "l"(p), "l"(pol));
    return v;
  } else {
    unsigned short h;
    asm volatile("ld.global.nc.L2::cache_hint.L2::64B.u16 %0, [%1], %2;" : "=h"(h) : "l"(p), "l"(pol));
    return to_f32<T>(*reinterpret_cast<const T*>(&h));
  }
#elif B200CTC_GATHER_LTC == 64
  if (sizeof(T) == 4) {
    float v;
    asm volatile("ld.global.nc.L2::64B.f32 %0, [%1];" : "=f"(v) : "l"(p));
    return v;
  } else {
    unsigned short h;
    asm volatile("ld.global.nc.L2::64B.u16 %0, [%1];" : "=h"(h) : "l"(p));
    return to_f32<T>(*reinterpret_cast<const T*>(&h));
  }
#else
  return to_f32<T>(__ldg(p));
#endif
}
#ifndef B200CTC_GATHER_WARPS
#define B200CTC_GATHER_WARPS 8
#endif
#ifndef B200CTC_GATHER_ROWS
#define B200CTC_GATHER_ROWS 1
#endif
constexpr int kGatherWarps = B200CTC_GATHER_WARPS;
constexpr int kGatherRows = B200CTC_GATHER_ROWS;   // rows per warp in flight
constexpr int kGatherRegs = 8;   // values kept in registers per lane and row (L+1 <= 256)

template <typename T>
__global__ void __launch_bounds__(kGatherWarps * 32) gather_kernel(
    const T* __restrict__ logits, PrepArgs P, float* __restrict__ em, float* __restrict__ cmax,
    int pdl) {
  extern __shared__ int slab[];  // [Lp] label columns of this utterance
  // launched with programmatic stream serialisation (its launch latency overlaps the tail of the
  // previous kernel); everything it reads may come from that kernel: wait first
  griddep_wait();
  if (pdl == 0) griddep_launch();  // the lattice CTAs may become resident and set up meanwhile
  struct AtExit {  // every return path below fires the end-of-kernel trigger
    int pdl;
    __device__ ~AtExit() { if (pdl == 1) griddep_launch(); }
  } at_exit{pdl};
  const int b = blockIdx.y;
  int Tb, L;
  prep_utterance(P, b, slab, blockIdx.x == 0, Tb, L);
  const int Tfr = P.T, V = P.V, Lp = P.Lp;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  constexpr int kFramesPerCta = kGatherWarps * kGatherRows;
  const int t0 = blockIdx.x * kFramesPerCta + wid * kGatherRows;
  if (t0 >= Tb) return;
  const int nr = Tb - t0 < kGatherRows ? Tb - t0 : kGatherRows;
  const size_t g0 = (size_t)b * Tfr + t0;
  const bool in_regs = L + 1 <= 32 * kGatherRegs;
  float m[kGatherRows];
#pragma unroll
  for (int r = 0; r < kGatherRows; ++r) m[r] = kNegInf;
  if (in_regs) {
    float v[kGatherRows][kGatherRegs];
#pragma unroll
    for (int j = 0; j < kGatherRegs; ++j) {
      const int k = lane + 32 * j;
      const int col = slab[k <= L ? k : 0];
      B2_DEV_CHECK(col >= 0 && col < V);
#pragma unroll
      for (int r = 0; r < kGatherRows; ++r) {  // clamped row: always a valid address
        const T* row = logits + (g0 + (r < nr ? r : nr - 1)) * V;
        v[r][j] = (j == 0 || k <= L) ? ldg_gather<T>(row + col) * kLog2e : kNegInf;
      }
    }
#pragma unroll
    for (int r = 0; r < kGatherRows; ++r) {
#pragma unroll
      for (int j = 0; j < kGatherRegs; ++j) m[r] = fmaxf(m[r], (lane + 32 * j <= L) ? v[r][j] : kNegInf);
      m[r] = warp_max(m[r]);
      if (!(m[r] > kNegInf && m[r] < -kNegInf)) m[r] = 0.f;  // all -inf / NaN rows: no centring
    }
#pragma unroll
    for (int r = 0; r < kGatherRows; ++r) {
      if (r < nr) {
        float* er = em + (g0 + r) * Lp;
#pragma unroll
        for (int j = 0; j < kGatherRegs; ++j) {
          const int k = lane + 32 * j;
          if (k <= L) er[k == 0 ? 0 : kEmLabel0 + k - 1] = v[r][j] - m[r];
        }
        if (lane == 0) cmax[g0 + r] = m[r];
      }
    }
  } else {
    for (int r = 0; r < nr; ++r) {
      const T* row = logits + (g0 + r) * V;
      float mm = kNegInf;
      for (int k = lane; k <= L; k += 32) mm = fmaxf(mm, to_f32<T>(__ldg(row + slab[k])) * kLog2e);
      mm = warp_max(mm);
      if (!(mm > kNegInf && mm < -kNegInf)) mm = 0.f;
      float* er = em + (g0 + r) * Lp;
      for (int k = lane; k <= L; k += 32)
        er[k == 0 ? 0 : kEmLabel0 + k - 1] = to_f32<T>(__ldg(row + slab[k])) * kLog2e - mm;
      if (lane == 0) cmax[g0 + r] = mm;
    }
  }
}

template <typename T>
static int launch_gather(const void* logits, const PrepArgs& P, int B, const LossLayout& w,
                         unsigned char* ws, cudaStream_t st) {
  constexpr int fpc = kGatherWarps * kGatherRows;
  int gx = (P.T + fpc - 1) / fpc;
  if (gx < 1) gx = 1;  // the first CTA of every utterance also publishes the label prep
  dim3 grid(gx, B);
  const size_t smem = sizeof(int) * (size_t)w.Lp;
  auto k = gather_kernel<T>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  launch_pdl(k, grid, dim3(kGatherWarps * 32), smem, st, 3, static_cast<const T*>(logits), P,
             reinterpret_cast<float*>(ws + w.em), reinterpret_cast<float*>(ws + w.cmax), pdl_mode());
  count_launch();
  return check_launch("ctc gather");
}

}  // namespace b200ctc
#ifdef B200CTC_TUNING  // measured loser at C2 (profiles/README.md round 2): tuning builds only
#include "ctc_fused.cuh"
#endif
namespace b200ctc {

// --------------------------------------------------------- pass A (rows)
// Persistent CTAs; each loops over (b,t) rows.  TMA=true: rows arrive in an
// NSTAGE-deep shared-memory ring via cp.async.bulk; TMA=false (row bytes or
// base not 16-B aligned): the CTA copies the row itself.  Few threads per row
// (128 for V<=5120) keep the reduction overhead per element small; several
// CTAs per SM overlap each other's barriers.
template <typename T, int NT, int PER, bool TMA>
__global__ void __launch_bounds__(NT) lse_gather_kernel(
    const T* __restrict__ logits, int B, int Tfr, int V, const UttInfo* __restrict__ utt,
    const int* __restrict__ lab, int Lp, float* __restrict__ lse2, float* __restrict__ cmax,
    float* __restrict__ em, int nstage, int stage_stride) {
  constexpr int N = Elems<T>::N;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t bars[kMaxRowStages];
  __shared__ float red[2][32];
  unsigned char* stages = dsm;
  float* gbuf = reinterpret_cast<float*>(dsm + (size_t)nstage * stage_stride);  // [2][Lp]
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t nrows = (int64_t)B * Tfr;
  const uint32_t row_bytes = (uint32_t)V * sizeof(T);
  const int nchunk = V / N;
  const int ntail = V - nchunk * N;

  RowCursor pc;  // producer cursor (thread 0 only)
  pc.init(blockIdx.x, gridDim.x, Tfr);
  uint32_t np = 0;
  auto issue_next = [&]() {
    while (pc.g < nrows && pc.t >= utt[pc.b].T) pc.next();
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
  for (rc.init(blockIdx.x, gridDim.x, Tfr); rc.g < nrows; rc.next()) {
    const int b = rc.b;
    const int2 TL = *reinterpret_cast<const int2*>(&utt[b].T);  // {T, L}
    if (rc.t >= TL.x) continue;
    const int L = TL.y;
    const int64_t g = rc.g;
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
    // whole row -> registers
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
      xt = to_f32<T>(reinterpret_cast<const T*>(stg)[nchunk * N + tid]);
      m = fmaxf(m, xt);
    }
    // raw logits of the blank (k = 0) and the L labels -> scratch (double buffered per row)
    float* gb = gbuf + (n & 1) * Lp;
    const int* lb = lab + (size_t)b * Lp;
    for (int k = tid; k <= L; k += NT) gb[k] = to_f32<T>(reinterpret_cast<const T*>(stg)[lb[k]]);
    // block max
    m = warp_max(m);
    if (NT > 32) {
      if (lane == 0) red[0][wid] = m;
      __syncthreads();
      m = warp_max(lane < NT / 32 ? red[0][lane] : kNegInf);
    }
    // every thread holds its part of the row in registers / scratch: the stage is free
    if (TMA && tid == 0) issue_next();
    const float mm = (m == kNegInf) ? 0.f : m;
    const float m2 = mm * kLog2e;
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      const int i = tid + j * NT;
      if (i < nchunk) {
#pragma unroll
        for (int e = 0; e < N; ++e) s += ex2(fmaf(x[j][e], kLog2e, -m2));
      }
    }
    if (tid < ntail) s += ex2(fmaf(xt, kLog2e, -m2));
    s = warp_sum(s);
    if (NT > 32) {
      if (lane == 0) red[1][wid] = s;
      __syncthreads();
      s = warp_sum(lane < NT / 32 ? red[1][lane] : 0.f);
    }
    if (tid == 0) {
      lse2[g] = m2 + log2f(s);
      cmax[g] = m2;  // the per-frame constant of this mode is the row maximum
    }
    float* er = em + (size_t)g * Lp;
    for (int k = tid; k <= L; k += NT) er[k == 0 ? 0 : kEmLabel0 + k - 1] = fmaf(gb[k], kLog2e, -m2);
    ++n;
  }
}

// launch helper -------------------------------------------------------------
template <typename T, int NT>
static int launch_pass_a(const void* logits, int B, int Tfr, int V, const LossLayout& w,
                         unsigned char* ws, cudaStream_t st) {
  constexpr int N = Elems<T>::N;
  constexpr int PER = 40 / N;
  const size_t extra = align_up(sizeof(float) * 2 * (size_t)w.Lp, 128);
  const RowLaunchCfg c = row_launch_cfg(logits, (size_t)V * sizeof(T), extra, NT, "B200CTC_A");
  B2_REQUIRE((size_t)c.nstage * c.stride + extra <= 226 * 1024,
             "ctc_loss: V=%d too large for the shared-memory row path", V);
  const size_t smem = (size_t)c.nstage * c.stride + extra;
  const int64_t rows = (int64_t)B * Tfr;
  int64_t grid = (int64_t)sm_count() * c.ctas_per_sm;
  if (grid > rows) grid = rows;
  auto* utt = reinterpret_cast<const UttInfo*>(ws + w.utt);
  auto* lab = reinterpret_cast<const int*>(ws + w.lab);
  auto* lse2 = reinterpret_cast<float*>(ws + w.lse2);
  auto* cmax = reinterpret_cast<float*>(ws + w.cmax);
  auto* em = reinterpret_cast<float*>(ws + w.em);
  auto go = [&](auto k) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<(unsigned)grid, NT, smem, st>>>(static_cast<const T*>(logits), B, Tfr, V, utt, lab, w.Lp,
                                        lse2, cmax, em, c.nstage, c.stride);
  };
  if (c.tma) go(lse_gather_kernel<T, NT, PER, true>);
  else go(lse_gather_kernel<T, NT, PER, false>);
  count_launch();
  return check_launch("ctc lse_gather");
}

template <typename T>
static int dispatch_pass_a(const void* logits, int B, int Tfr, int V, const LossLayout& w,
                           unsigned char* ws, cudaStream_t st) {
  const int need = (V + 39) / 40;  // threads so that NT*PER*N >= V (PER*N = 40)
  if (need <= 128) return launch_pass_a<T, 128>(logits, B, Tfr, V, w, ws, st);
  if (need <= 256) return launch_pass_a<T, 256>(logits, B, Tfr, V, w, ws, st);
  if (need <= 512) return launch_pass_a<T, 512>(logits, B, Tfr, V, w, ws, st);
  if (need <= 1024) return launch_pass_a<T, 1024>(logits, B, Tfr, V, w, ws, st);
  set_error("ctc_loss: V=%d exceeds the supported row size", V);
  return B200CTC_ERR_ARG;
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" size_t b200ctc_loss_workspace_bytes(int B, int T, int V, int max_tgt_len) {
  if (B < 0 || T < 0 || V <= 0 || max_tgt_len < 0) return 0;
  return make_loss_layout(B, T, V, max_tgt_len).total;
}

namespace b200ctc {

static int check_loss_args(const char* who, const void* logits, int dtype, const int64_t* targets,
                           const int64_t* tgt_offsets, const int64_t* in_lens,
                           const int64_t* tgt_lens, int B, int T, int V, int max_tgt_len, int blank,
                           const float* nll, const void* ws_, size_t ws_bytes, LossLayout* out) {
  B2_REQUIRE(B >= 0 && T >= 0 && V > 0 && max_tgt_len >= 0, "%s: bad shape", who);
  if (B == 0) return B200CTC_OK;
  (void)tgt_offsets;  // may be NULL: prep_kernel then prefix-sums tgt_lens itself
  B2_REQUIRE(logits && in_lens && tgt_lens && nll && ws_, "%s: null pointer", who);
  B2_REQUIRE(targets || max_tgt_len == 0, "%s: null targets", who);
  B2_REQUIRE(blank >= 0 && blank < V, "%s: blank=%d outside [0,%d)", who, blank, V);
  B2_REQUIRE(dtype >= 0 && dtype <= 2, "%s: unknown dtype %d", who, dtype);
  B2_REQUIRE(2 * (int64_t)max_tgt_len + 1 <= kMaxLattice,
             "%s: max_tgt_len=%d exceeds the supported lattice width %d", who, max_tgt_len,
             kMaxLattice);
  *out = make_loss_layout(B, T, V, max_tgt_len);
  if (ws_bytes < out->total) {
    set_error("%s: workspace %zu < required %zu bytes", who, ws_bytes, out->total);
    return B200CTC_ERR_WORKSPACE;
  }
  B2_REQUIRE((reinterpret_cast<uintptr_t>(ws_) & 255) == 0, "%s: workspace must be 256-B aligned",
             who);
  return B200CTC_OK;
}

static PrepArgs make_prep_args(const LossLayout& w, unsigned char* ws, const int64_t* targets,
                               const int64_t* tgt_offsets, const int64_t* in_lens,
                               const int64_t* tgt_lens, int blank) {
  PrepArgs P;
  P.targets = targets; P.offs = tgt_offsets; P.in_lens = in_lens; P.tgt_lens = tgt_lens;
  P.T = w.T; P.V = w.V; P.Lmax = w.Lmax; P.Lp = w.Lp; P.blank = blank;
  P.utt = reinterpret_cast<UttInfo*>(ws + w.utt);
  P.lab = reinterpret_cast<int*>(ws + w.lab);
  P.dup = reinterpret_cast<int*>(ws + w.dup);
  P.counter = reinterpret_cast<int*>(ws + w.counter);
  P.err = reinterpret_cast<int*>(ws + w.err);
  return P;
}

static int launch_prep(const LossLayout& w, const PrepArgs& P, cudaStream_t st) {
  int nt = (w.Lmax + 1 + 31) / 32 * 32;
  if (nt < 64) nt = 64;
  if (nt > 1024) nt = 1024;
  prof_mark(0, 0, st);
  // the vocabulary table pays off for long targets and fits for V <= 12 K (96 KB)
  const int use_table = w.Lmax >= 256 && (size_t)w.V * 8 <= 96 * 1024;
  const size_t smem = sizeof(int) * ((size_t)w.Lp + (use_table ? 2 * (size_t)w.V : 0));
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(prep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  prep_kernel<<<w.B, nt, smem, st>>>(P, use_table);
  count_launch();
  prof_mark(0, 1, st);
  return check_launch("ctc prep");
}

// defined in ctc_bwd.cu
int launch_grad_rows(const void* logits, int dtype, void* grad, int B, int T, int V, int blank,
                     const LossLayout& w, unsigned char* ws, const float* gscale, int gstride,
                     float gmul, bool last_use, cudaStream_t st);

}  // namespace b200ctc

// forward (deferred-gradient mode): row pass A (lse + emissions) -> lattice -> nll
extern "C" int b200ctc_loss_fwd(const void* logits, int dtype, const int64_t* targets,
                                const int64_t* tgt_offsets, const int64_t* in_lens,
                                const int64_t* tgt_lens, int B, int T, int V, int max_tgt_len,
                                int blank, int zero_infinity, int need_grad, float* nll,
                                float* loss_out, float loss_scale, void* ws_, size_t ws_bytes,
                                b200ctc_stream_t stream) {
  LossLayout w;
  int rc = check_loss_args("loss_fwd", logits, dtype, targets, tgt_offsets, in_lens, tgt_lens, B, T,
                           V, max_tgt_len, blank, nll, ws_, ws_bytes, &w);
  if (rc || B == 0) return rc;
  unsigned char* ws = static_cast<unsigned char*>(ws_);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const PrepArgs P = make_prep_args(w, ws, targets, tgt_offsets, in_lens, tgt_lens, blank);
  rc = launch_prep(w, P, st);
  if (rc) return rc;
  if (T > 0) {
    prof_mark(1, 0, st);
    switch (dtype) {
      case B200CTC_F32: rc = dispatch_pass_a<float>(logits, B, T, V, w, ws, st); break;
      case B200CTC_F16: rc = dispatch_pass_a<__half>(logits, B, T, V, w, ws, st); break;
      default: rc = dispatch_pass_a<__nv_bfloat16>(logits, B, T, V, w, ws, st); break;
    }
    prof_mark(1, 1, st);
    if (rc) return rc;
  }
  prof_mark(2, 0, st);
  rc = launch_lattice(w, ws, T, zero_infinity, need_grad, st);
  prof_mark(2, 1, st);
  if (rc) return rc;
  return launch_nll(w, ws, T, zero_infinity, nll, loss_out, loss_scale, st);
}

// forward + gradient in one call (eager mode): the logits are read ONCE in
// full (row pass B); the lattice runs on a sparse gather of the label columns.
// peer != NULL: the batch-sharded variant - the kernel that finishes the loss also
// all-reduces {sum of losses, utterance count} over the ranks through peer memory.
static int loss_fwd_grad_impl(const void* logits, int dtype, const int64_t* targets,
                              const int64_t* tgt_offsets, const int64_t* in_lens,
                              const int64_t* tgt_lens, int B, int T, int V, int max_tgt_len,
                              int blank, int zero_infinity, float grad_hint, float* nll,
                              float* loss_out, float loss_scale, void* grad, void* ws_,
                              size_t ws_bytes, b200ctc_stream_t stream, b200ctc_peer* peer,
                              float* global_out, const float* utt_hints = nullptr) {
  PeerDev pd{nullptr, 0, 0, 0, 0};
  if (peer) {
    B2_REQUIRE(peer->connected, "loss_fwd_grad_sharded: peer handle is not connected");
    B2_REQUIRE(global_out, "loss_fwd_grad_sharded: null global_out");
    B2_REQUIRE(B > 0, "loss_fwd_grad_sharded: a rank without utterances must call b200ctc_peer_exchange");
    pd = PeerDev{peer->d_bufs, peer->world, peer->rank, peer->deferred, peer->timeout_ns};
  }
  LossLayout w;
  int rc = check_loss_args("loss_fwd_grad", logits, dtype, targets, tgt_offsets, in_lens, tgt_lens,
                           B, T, V, max_tgt_len, blank, nll, ws_, ws_bytes, &w);
  if (rc || B == 0) return rc;
  B2_REQUIRE(grad || T == 0, "loss_fwd_grad: null grad");
  B2_REQUIRE(utt_hints || (grad_hint != 0.f && grad_hint == grad_hint),
             "loss_fwd_grad: grad_hint must be non-zero");
  unsigned char* ws = static_cast<unsigned char*>(ws_);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  PrepArgs P = make_prep_args(w, ws, targets, tgt_offsets, in_lens, tgt_lens, blank);
  // Emissions: a sparse gather of the L+1 label columns moves one 64-byte DRAM atom per
  // column and row; once that exceeds the row itself (long targets: C4 has 1201 columns
  // = 77 KB of atoms against a 20 KB row) the full-row pass A (TMA row -> shared memory ->
  // columns) is the cheaper way to produce the same em / cmax arrays.
  static const int force_dense = tuning_int("B200CTC_DENSE_GATHER", -1);
  const bool dense_gather = T > 0 && (force_dense >= 0 ? force_dense != 0
      : (size_t)(w.Lmax + 1) * 64 > (size_t)V * (size_t)dtype_size(dtype));
  bool fused = false;
#ifdef B200CTC_TUNING
  // experiment (ctc_fused.cuh): prep + gather + lattice as ONE launch, SMs partitioned by role
  static const int fused_on = tuning_int("B200CTC_FUSED", 0);
  int fnring = 0;
  size_t fsmem = 0;
  fused = T > 0 && !dense_gather && fused_on && fused_fwd_fits(w, &fnring, &fsmem);
  if (fused) {
    P.err = nullptr;  // cleared by the launcher's memset (CTAs may set it before utterance 0's prep)
    prof_mark(2, 0, st);
    switch (dtype) {
      case B200CTC_F32: rc = launch_fused_fwd<float>(logits, P, w, ws, zero_infinity, fnring, fsmem, st); break;
      case B200CTC_F16: rc = launch_fused_fwd<__half>(logits, P, w, ws, zero_infinity, fnring, fsmem, st); break;
      default: rc = launch_fused_fwd<__nv_bfloat16>(logits, P, w, ws, zero_infinity, fnring, fsmem, st); break;
    }
    prof_mark(2, 1, st);
  }
#endif
  if (!fused) {
    if (dense_gather) {
      rc = launch_prep(w, P, st);
      if (!rc) {
        prof_mark(1, 0, st);
        switch (dtype) {
          case B200CTC_F32: rc = dispatch_pass_a<float>(logits, B, T, V, w, ws, st); break;
          case B200CTC_F16: rc = dispatch_pass_a<__half>(logits, B, T, V, w, ws, st); break;
          default: rc = dispatch_pass_a<__nv_bfloat16>(logits, B, T, V, w, ws, st); break;
        }
        prof_mark(1, 1, st);
      }
    } else {  // gather kernel with the label prep fused in
      prof_mark(1, 0, st);
      switch (dtype) {
        case B200CTC_F32: rc = launch_gather<float>(logits, P, B, w, ws, st); break;
        case B200CTC_F16: rc = launch_gather<__half>(logits, P, B, w, ws, st); break;
        default: rc = launch_gather<__nv_bfloat16>(logits, P, B, w, ws, st); break;
      }
      prof_mark(1, 1, st);
    }
    if (!rc) {
      prof_mark(2, 0, st);
      rc = launch_lattice(w, ws, T, zero_infinity, 1, st);
      prof_mark(2, 1, st);
    }
  }
  if (rc) return rc;
  if (T > 0) {
    prof_mark(3, 0, st);
    // the expected upstream gradient is baked into the gradient: one scalar, or one per utterance
    rc = utt_hints ? launch_grad_rows(logits, dtype, grad, B, T, V, blank, w, ws, utt_hints, 1, 1.f, true, st)
                   : launch_grad_rows(logits, dtype, grad, B, T, V, blank, w, ws, nullptr, 0, grad_hint, true, st);
    prof_mark(3, 1, st);
    if (rc) return rc;
  }
  return launch_nll(w, ws, T, zero_infinity, nll, loss_out, loss_scale, st, pd, global_out);
}

extern "C" int b200ctc_loss_fwd_grad(const void* logits, int dtype, const int64_t* targets,
                                     const int64_t* tgt_offsets, const int64_t* in_lens,
                                     const int64_t* tgt_lens, int B, int T, int V,
                                     int max_tgt_len, int blank, int zero_infinity,
                                     float grad_hint, float* nll, float* loss_out,
                                     float loss_scale, void* grad, void* ws, size_t ws_bytes,
                                     b200ctc_stream_t stream) {
  return loss_fwd_grad_impl(logits, dtype, targets, tgt_offsets, in_lens, tgt_lens, B, T, V,
                            max_tgt_len, blank, zero_infinity, grad_hint, nll, loss_out, loss_scale,
                            grad, ws, ws_bytes, stream, nullptr, nullptr);
}

extern "C" int b200ctc_loss_fwd_grad_sharded(const void* logits, int dtype, const int64_t* targets,
                                             const int64_t* tgt_offsets, const int64_t* in_lens,
                                             const int64_t* tgt_lens, int B, int T, int V,
                                             int max_tgt_len, int blank, int zero_infinity,
                                             float grad_hint, float* nll, float* loss_out,
                                             float loss_scale, void* grad, void* ws,
                                             size_t ws_bytes, b200ctc_peer* peer,
                                             float* global_out, b200ctc_stream_t stream) {
  B2_REQUIRE(peer, "loss_fwd_grad_sharded: null peer handle");
  return loss_fwd_grad_impl(logits, dtype, targets, tgt_offsets, in_lens, tgt_lens, B, T, V,
                            max_tgt_len, blank, zero_infinity, grad_hint, nll, loss_out, loss_scale,
                            grad, ws, ws_bytes, stream, peer, global_out);
}

// per-utterance hints (several CTC heads with different loss weights in ONE batched call)
extern "C" int b200ctc_loss_fwd_grad_hints(const void* logits, int dtype, const int64_t* targets,
                                           const int64_t* tgt_offsets, const int64_t* in_lens,
                                           const int64_t* tgt_lens, int B, int T, int V,
                                           int max_tgt_len, int blank, int zero_infinity,
                                           const float* utt_hints, float* nll, void* grad, void* ws,
                                           size_t ws_bytes, b200ctc_stream_t stream) {
  B2_REQUIRE(utt_hints || B == 0, "loss_fwd_grad_hints: null utt_hints");
  return loss_fwd_grad_impl(logits, dtype, targets, tgt_offsets, in_lens, tgt_lens, B, T, V,
                            max_tgt_len, blank, zero_infinity, 0.f, nll, nullptr, 0.f, grad, ws,
                            ws_bytes, stream, nullptr, nullptr, utt_hints);
}
