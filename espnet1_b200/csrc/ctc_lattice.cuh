// Alpha/beta lattice recursion (Graves 2006 eq. 6-11) in normalised log2 space.
//
// One CTA per (utterance, direction): blockIdx.y = 0 runs alpha forward in
// time, 1 runs beta backward in time as "alpha on the reversed lattice"
// (recursion position sp <-> lattice index s = S-1-sp).
//
// The CTA is a SYSTOLIC ARRAY OF WARPS with no CTA-wide barrier on the serial
// chain.  Thread i keeps R consecutive lattice positions in registers; inside a
// warp the one neighbour value it needs arrives by shuffle; between warps it
// travels through a shared-memory ring written by the last lane of warp w and
// read by warp w+1.  The dependency is one-directional (position sp needs
// sp-1, sp-2), so warp w+1 simply runs one BLOCK of kBlk steps behind warp w:
// all hand-shaking (ready / consumed counters, frame offsets) happens once per
// block, and inside a block the steps are straight-line code - one LDS, one
// STS and one SHFL per step beside the arithmetic.  Every warp re-centres its
// own slice on its own maximum every kRenormEvery steps with an INTEGER shift
// (exact to combine, see ctc_ws.cuh), so there is no reduction across warps
// either.  The emissions em[t][k] are streamed into a shared-memory ring by a
// dedicated producer warp with the TMA engine (1-D bulk copies of kBlk-frame
// chunks, full / empty mbarriers).
#pragma once
#include "ctc_ws.cuh"
#include "peer.cuh"

namespace b200ctc {

constexpr int kMaxLattice = 8192;  // 2*L+1 <= 8192  (R = 16, 16 warps + producer)
constexpr int kValRing = 64;       // boundary values in flight between adjacent warps (steps)
constexpr int kFRing = 2 * kValRing / kRenormEvery;  // frame offsets, one per renorm block
constexpr int kMaxEmRing = 20;     // chunks of emission frames in flight

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// The lattice uses a FINITE log-zero so that the serial chain needs no -inf /
// NaN guards: kLogZero absorbs every real increment (ulp(1e30) ~ 7e22), the
// difference of two log-zeros is 0 (not NaN), and 2^(kLogZero - x) underflows to 0.
constexpr float kLogZero = -1.0e30f;
#ifndef B200CTC_REF_OWN
#define B200CTC_REF_OWN 1
#endif
constexpr bool kRefOwn = B200CTC_REF_OWN != 0;
constexpr float kLogZeroTest = -1.0e29f;  // "is alive" threshold

// log2(2^a + 2^b) + e : 2 MUFU, chain FADD -> EX2 -> FADD -> LG2 -> FADD
__device__ __forceinline__ float lse2_fast(float a, float b, float e) {
  const float me = fmaxf(a, b) + e;
  return me + lg2(1.f + ex2(-fabsf(a - b)));
}
// log2(2^a + 2^b + 2^c) + e : 4 MUFU but the shortest dependent chain
__device__ __forceinline__ float lse3_fast(float a, float b, float c, float e) {
  const float m = fmaxf(fmaxf(a, b), c);
  const float s = ex2(a - m) + ex2(b - m) + ex2(c - m);
  return (m + e) + lg2(s);
}

// same with 3 MUFU (the largest term is exactly 1): for wide lattices, which are
// MUFU-throughput-bound rather than latency-bound
__device__ __forceinline__ float lse3_lean(float a, float b, float c, float e) {
  const float hi = fmaxf(a, b), lo = fminf(a, b);
  const float m = fmaxf(hi, c), t = fminf(hi, c);
  const float mid = fmaxf(lo, t), mn = fminf(lo, t);
  return (m + e) + lg2(1.f + ex2(mid - m) + ex2(mn - m));
}

// raw 32-bit shared-memory addressing for the serial loop (no generic->shared
// conversions or re-derived bases inside it)
__device__ __forceinline__ float4 lds_v4_u32(uint32_t addr) {
  float4 v;
  asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_f32_u32(uint32_t addr, float v) {
  asm volatile("st.volatile.shared.f32 [%0], %1;" ::"r"(addr), "f"(v) : "memory");
}
__device__ __forceinline__ int lds_i32_u32(uint32_t addr) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr));
  return v;
}
__device__ __forceinline__ void sts_i32_u32(uint32_t addr, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ int lds_acquire_i32(uint32_t addr) {
  int v;
  asm volatile("ld.acquire.cta.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void sts_release_i32(uint32_t addr, int v) {
  asm volatile("st.release.cta.shared.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ float lds_f32_u32(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
// emission ring reads: ordered after the mbarrier wait (volatile asm), plain ld.shared
__device__ __forceinline__ float lds_em(uint32_t addr) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ void mbar_arrive_u32(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// alpha / beta rows: scratch that the gradient row pass re-reads -> L2 policy operand
__device__ __forceinline__ void stg_v2(float* p, float x, float y, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v2.f32 [%0], {%1,%2}, %3;" ::"l"(p), "f"(x), "f"(y), "l"(pol)
               : "memory");
}
__device__ __forceinline__ void stg_v4(float* p, float x, float y, float z, float w, uint64_t pol) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "f"(x), "f"(y),
               "f"(z), "f"(w), "l"(pol)
               : "memory");
}
// Split lattices: boundary hand-off between CTAs through global memory (L2).  Every word
// validates itself: the arrays are filled with a sentinel before the launch, the producer
// overwrites a word exactly once (relaxed gpu-scope store), the consumer re-reads (relaxed
// gpu-scope loads, which can be issued a block ahead) until no sentinel is left.  No flag, no
// acquire/release fence on the serial chain.  Lattice values are never NaN by construction
// (finite log-zero), and arithmetic NaNs are canonical (0x7fffffff), not the sentinel.
constexpr uint32_t kBndValSentinel = 0xffffffffu;  // cudaMemset byte 0xff
constexpr int kBndFSentinel = (int)0x80808080u;    // cudaMemset byte 0x80 (|F| stays far below)
constexpr int kBndSpinLimit = 1 << 22;             // a lost producer: sticky error + NaN loss, not a hang
__device__ __forceinline__ void stg_relaxed_f32(float* p, float v) {
  asm volatile("st.relaxed.gpu.global.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}
__device__ __forceinline__ void stg_relaxed_i32(int* p, int v) {
  asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ float4 ldg_relaxed_v4(const float* p) {
  float4 v;
  asm volatile("ld.relaxed.gpu.global.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float ldg_relaxed_f32(const float* p) {
  float v;
  asm volatile("ld.relaxed.gpu.global.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ int ldg_relaxed_i32(const int* p) {
  int v;
  asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ bool bnd_written(float v) { return __float_as_uint(v) != kBndValSentinel; }
// make a loop-invariant value opaque so that it is kept in a register instead
// of being re-derived inside the serial loop
__device__ __forceinline__ uint32_t pin(uint32_t v) {
  asm volatile("mov.b32 %0, %0;" : "+r"(v));
  return v;
}

// alpha / beta are written by the lattice and read once by the gradient row pass, with the
// whole (B,T,V) logits + gradient streaming through L2 in between: evict_last keeps them there
static inline int lat_policy_host() {
  static const int v = tuning_int("B200CTC_LAT_POL", 2);
  return v;
}
struct LatticeArgs {
  UttInfo* utt;
  const int* lb;      // labels of this utterance (k = 0 blank)
  float* lat;         // alpha or beta rows of this utterance, (Tfr, Sp)
  int* off;           // offA or offB of this utterance, (NB, NWp): one entry per renorm block
  uint32_t ring;      // shared: em ring base
  uint32_t vals;      // shared: boundary values  [NW][kValRing] float
  uint32_t fring;     // shared: frame offsets    [NW][kFRing]  int
  uint32_t ready;     // shared: steps published  [NW] int
  uint32_t cons;      // shared: steps consumed   [NW] int
  uint32_t full;      // shared: full[kMaxEmRing] mbarriers, then empty[kMaxEmRing]
  int Tb, L, Lp, Sp, NWp, nw_live, zero_infinity, b, nring;
  // split lattices (SPLIT): global index of this CTA's warp 0, staged emission row (stride
  // Lps floats: [blank,-,-,-][label columns col0 ..]), hand-off arrays of the CTA boundary
  // in front of this CTA (in) and behind it (out): values [Tb1] (step j at j+1), frame
  // offsets [NB], steps published [1]
  int W0, Lps, col0, nwc;
  float* gv_in; int* gf_in; int* gr_in;
  float* gv_out; int* gf_out; int* gr_out;
  int* err;  // sticky error word of the workspace (ctc_ws.cuh)
  int lat_policy;  // L2 policy of the alpha / beta stores (common.cuh l2_policy)
};

// One compute warp of the systolic lattice (see the file header).  This is the
// serial chain of the whole loss: one warp per SM sub-partition with a single
// step in flight, so every instruction on the chain costs its full latency and
// everything else is moved off it:
//  * branch-free arithmetic (finite kLogZero; step 0 is an ordinary step started
//    from a virtual alpha_{-1}(0) = log 1; positions beyond the lattice compute
//    finite don't-care values and are masked only where they could be observed:
//    the re-centring maximum and the final read-out),
//  * compile-time direction, full blocks of K steps as straight-line code with the
//    block's emissions and upstream boundary values already in registers,
//  * all hand-shaking once per block, ring slots by increment (no div / mod).
template <int R, bool STORE, int DIR, int K, bool SPLIT>
__device__ __forceinline__ void lattice_warp(const LatticeArgs& A, int w, int lane) {
  static_assert(K == 4 || K == 8 || K == 16, "block of 4, 8 or 16 steps");
  static_assert(kRenormEvery % K == 0 && kValRing >= 4 * K, "block size vs renorm period / ring");
  constexpr bool kPre = R <= 4;  // whole block's emissions preloaded into registers
  const int Tb = A.Tb, L = A.L, S = 2 * L + 1, Lp = SPLIT ? A.Lps : A.Lp;
  const int gw = SPLIT ? A.W0 + w : w;  // warp index within the whole lattice
  const int sp0 = (gw * 32 + lane) * R;  // first recursion position of this thread
  // static description of this thread's positions: q even = blank, q odd = label
  uint32_t jlab[R / 2];  // byte offset of the label emission inside an em row
  bool skip[R / 2];      // may the path skip the blank in front of this label?
  {
    int la[R / 2], ln[R / 2];
    bool ok[R / 2];
#pragma unroll
    for (int h = 0; h < R / 2; ++h) {  // unconditional clamped loads: one global latency in all
      const int sp = sp0 + 2 * h + 1;
      const int s = DIR ? S - 1 - sp : sp;
      const int k = (s + 1) >> 1;  // 1..L when active
      const int kn = DIR ? k + 1 : k - 1;
      const bool valid = sp < S;
      ok[h] = valid && (DIR ? kn <= L : kn >= 1);
      la[h] = A.lb[valid ? k : 0];
      ln[h] = A.lb[ok[h] ? kn : 0];
      // (split: the staged row is [blank group][columns col0 ..]); dead: the blank column
      jlab[h] = pin(valid ? 4u * (uint32_t)(kEmLabel0 + k - 1 - (SPLIT ? A.col0 - 4 : 0)) : 0u);
    }
#pragma unroll
    for (int h = 0; h < R / 2; ++h) skip[h] = ok[h] && la[h] != ln[h];
  }
  const bool has_up = gw > 0, has_down = gw + 1 < A.nw_live;
  const bool lane0 = lane == 0;
  // neighbours in another CTA: hand-off through global memory
  const bool up_g = SPLIT && has_up && w == 0;
  const bool down_g = SPLIT && has_down && w == A.nwc - 1;
  const bool pub = has_down && !down_g && lane == 31;
  const bool pubg = down_g && lane == 31;
  const int wu = (has_up && !up_g) ? w - 1 : 0;
  const uint32_t vals_in = pin(A.vals + (uint32_t)wu * (kValRing * 4u));
  const uint32_t vals_out = pin(A.vals + (uint32_t)w * (kValRing * 4u));
  const uint32_t fr_in = pin(A.fring + (uint32_t)wu * (kFRing * 4u));
  const uint32_t fr_out = pin(A.fring + (uint32_t)w * (kFRing * 4u));
  const uint32_t ready_in = pin(A.ready + 4u * (uint32_t)wu);
  const uint32_t ready_out = pin(A.ready + 4u * (uint32_t)w);
  const uint32_t cons_mine = pin(A.cons + 4u * (uint32_t)w);
  const uint32_t cons_down = pin(A.cons + 4u * (uint32_t)((has_down && !down_g) ? w + 1 : w));
  const uint32_t row_bytes = pin(4u * (uint32_t)Lp);
  const uint32_t chunk_bytes = pin(4u * (uint32_t)(K * Lp));
  const uint32_t bar_end = pin(A.full + 8u * (uint32_t)A.nring);

  // walking pointers (frames ascend for alpha, descend for beta)
  const int t0 = DIR ? Tb - 1 : 0;
  float* lat_p = nullptr;
  int* off_p = nullptr;
  if (STORE) {
    lat_p = A.lat + (size_t)t0 * A.Sp + sp0;
    off_p = A.off + gw;  // indexed by the renorm block of the RECURSION step, both directions
  }
  const ptrdiff_t lat_step = DIR ? -(ptrdiff_t)A.Sp : (ptrdiff_t)A.Sp;
  const bool store_ok = STORE && sp0 < A.Sp;
  const uint64_t lat_pol = l2_policy(A.lat_policy);
  const bool store_off = STORE && lane0;

  float a[R];
#pragma unroll
  for (int q = 0; q < R; ++q) a[q] = kLogZero;
  if (sp0 == 0) a[0] = 0.f;  // virtual alpha_{-1}(0) = log 1
  int F = 0;       // integer frame offset of a[] (ctc_ws.cuh)
  int Dpend = 0;   // re-centring shift measured at the end of the previous renorm block
  int Fup = 0;     // upstream warp's frame offset in the current renorm block
  float adj = 0.f; // Fup - F: upstream frame -> my frame
  int ready_seen = 0, cons_seen = 0;
  uint32_t slot_addr = A.ring, full_addr = A.full, phase = 0;
  // split, upstream in another CTA: the next block's hand-off (a flag round trip and a data
  // round trip through L2) is fetched DURING the current block's steps
  static_assert(!SPLIT || K == kRenormEvery, "split lattices: one renorm block per block");
  float upn[SPLIT ? K : 1];
  int Fupn = 0;
  bool have_next = false;

  // one recursion step: emissions eb (blank) / el[] (labels), p1 = value of position
  // sp0-1 after the previous step in MY frame (one neighbour value per thread: R is
  // even, so sp0 is a blank and needs only sp0-1, and the label at sp0+1 takes its
  // skip transition from sp0-1 as well); the new a[R-1] goes to slot `out`.
  auto step = [&](float eb, const float (&el)[R / 2], float upv, uint32_t out, float* gout) {
    float p1 = __shfl_up_sync(0xffffffffu, a[R - 1], 1);
    p1 = lane0 ? upv : p1;
    float nw_[R];
#pragma unroll
    for (int q = 0; q < R; ++q) {
      const float x0 = a[q];
      const float x1 = q >= 1 ? a[q - 1] : p1;
      if (kRefOwn && (R == 2 || B200CTC_REF_OWN >= 2) && q == 1) {
        // The label position that takes the neighbour's value p1 (fresh from the SHFL): the
        // two own terms are summed relative to THEIR maximum before p1 is known (off the
        // chain), p1 joins with one FADD -> EX2 -> FADD -> LG2 -> FADD instead of passing a
        // 3-way maximum first.  d > 24: the own terms are below fp32 resolution of the result.
        const float mo = fmaxf(x0, x1);
        const float so = ex2(x0 - mo) + ex2(x1 - mo);
        const float d = p1 - (skip[0] ? mo : 1.0e30f);
        // (no clamp: for huge d, ex2 gives +inf, r = +inf - never NaN - and is not selected)
        const float r = (mo + el[0]) + lg2(so + ex2(d));
        nw_[q] = d > 24.f ? p1 + el[0] : r;
      } else if (q & 1) {
        const float x2r = q >= 2 ? a[q - 2] : p1;
        const float x2 = skip[q >> 1] ? x2r : kLogZero;
        nw_[q] = (R >= 8 || SPLIT) ? lse3_lean(x0, x1, x2, el[q >> 1]) : lse3_fast(x0, x1, x2, el[q >> 1]);
      } else {
        nw_[q] = lse2_fast(x0, x1, eb);
      }
    }
#pragma unroll
    for (int q = 0; q < R; ++q) a[q] = nw_[q];
    if (pub) sts_f32_u32(out, a[R - 1]);
    if (SPLIT && pubg) stg_relaxed_f32(gout, a[R - 1]);
    // (asm stores: they stay inside their step.  Left to the scheduler they sink to the end
    //  of the block and queue up behind one another on a single address register pair.)
    if (store_ok) {
      if (R == 2) {
        stg_v2(lat_p, a[0], a[1], lat_pol);
      } else {
#pragma unroll
        for (int q = 0; q < R; q += 4) stg_v4(lat_p + q, a[q], a[q + 1], a[q + 2], a[q + 3], lat_pol);
      }
    }
    lat_p += lat_step;
  };

  for (int i0 = 0; i0 < Tb; i0 += K) {
    const int n = (Tb - i0 < K) ? Tb - i0 : K;
    // ---------------- block prologue: all hand-shaking happens here, once per K steps
    const uint32_t vb = vals_in + ((uint32_t)i0 & (kValRing - 1)) * 4u;
    float up[K];
#pragma unroll
    for (int u = 0; u < K; ++u) up[u] = kLogZero;
    const bool apply = (i0 & (kRenormEvery - 1)) == 0;  // first block of a renorm block
    int Fprev = 0;
    // The block's emissions go to registers BEFORE the hand-shake with the upstream warp: the
    // chunk has normally landed long ago, and a trailing warp would otherwise pay the 2*K
    // shared-memory loads between "upstream has published" and its first step.
    // (Only where the emission ring has slack, i.e. the short lattices that run 16-step blocks
    //  in one CTA.  Where the ring is just deeper than the warp skew, waiting for the chunk
    //  before acknowledging the upstream values couples the two rings: split C4 0.71 -> 0.80
    //  ms, one-CTA S=601 / 1201 / 2401 lattices +2..5 %.)
    float ebv[kPre ? K : 1], elv[kPre ? K : 1][R / 2];
    auto preload = [&]() {
      mbar_wait_u32(full_addr, phase);
      if (kPre && n == K) {
        uint32_t erow = slot_addr + (DIR ? (uint32_t)(K - 1) * row_bytes : 0u);
#pragma unroll
      for (int u = 0; u < K; ++u) {
        ebv[u] = lds_em(erow);
#pragma unroll
        for (int h = 0; h < R / 2; ++h) elv[u][h] = lds_em(erow + jlab[h]);
        if (DIR) erow -= row_bytes; else erow += row_bytes;
      }
      // (The slot is released only after the steps below have CONSUMED these registers:
      //  an mbarrier arrive issued right behind the loads can overtake them - it is
      //  served by a different unit - and the producer's next TMA write then lands
      //  under loads that are still queued.  Seen on B200 with a busy co-resident CTA.)
      }
    };
    constexpr bool kEarlyPre = !SPLIT && K == 16;  // short lattices only: their ring has slack
    if (kEarlyPre) preload();
    if (SPLIT && up_g) {
      // steps i0-1 .. i0+K-2 of the upstream CTA's last warp sit at gv_in[i0 .. i0+K-1]
      // steps i0-1 .. i0+K-2 of the upstream CTA's last warp sit at gv_in[i0 .. i0+K-1]
      Fprev = Fup;  // the upstream frame of the previous renorm block (0 before the first)
      if (have_next) {  // fetched under the previous block's steps
#pragma unroll
        for (int j = 0; j < K; ++j) up[j] = upn[j];
        Fup = Fupn;
      }
      for (int spin = 0; spin < kBndSpinLimit; ++spin) {
        bool ok = have_next || spin > 0;
        if (ok) {
          ok = Fup != kBndFSentinel;
          if (n == K) {  // (a partial last block reads its values one by one below)
#pragma unroll
            for (int j = 0; j < K; ++j) ok = ok && (bnd_written(up[j]) || (i0 == 0 && j == 0));
          }
        }
        if (__all_sync(0xffffffffu, ok)) break;
#pragma unroll
        for (int j = 0; j < K; j += 4) {
          const float4 v = ldg_relaxed_v4(A.gv_in + i0 + j);
          up[j] = v.x; up[j + 1] = v.y; up[j + 2] = v.z; up[j + 3] = v.w;
        }
        Fup = ldg_relaxed_i32(A.gf_in + (i0 >> 4));
        if (spin == kBndSpinLimit - 1 && lane0) atomicExch(A.err, kWsErrHandoff);
      }
      have_next = false;
      if (i0 == 0) up[0] = kLogZero;  // "step -1": nothing arrived yet
    } else if (has_up) {
      // value of step j sits in slot (j+1) & 63: this block needs steps i0-1 .. i0+K-2.
      // The flag load and the data loads are issued back to back (shared-memory loads of
      // a warp execute in order): if the flag then says "published", the data is good and
      // the common case costs one shared-memory round trip; otherwise poll and reload.
      const int need = (i0 + K - 1 < Tb - 1) ? i0 + K - 1 : Tb - 1;
      bool first = true;
      do {
        if (!first || ready_seen < need) ready_seen = lds_acquire_i32(ready_in);
        first = false;
#pragma unroll
        for (int j = 0; j < K; j += 4) {
          const float4 v = lds_v4_u32(vb + 4u * j);
          up[j] = v.x; up[j + 1] = v.y; up[j + 2] = v.z; up[j + 3] = v.w;
        }
        if (apply) {
          Fprev = lds_i32_u32(fr_in + (((uint32_t)((i0 - 1) >> 4)) & (kFRing - 1)) * 4u);
          Fup = lds_i32_u32(fr_in + (((uint32_t)(i0 >> 4)) & (kFRing - 1)) * 4u);
        }
      } while (ready_seen < need);
    }
    if (has_down && !down_g) {
      // my slots of this block last held steps i0-64 ..: the consumer read them in the
      // block that ended at step i0 + 2K - 64
      const int needc = i0 + 2 * K - kValRing;
      while (cons_seen < needc) cons_seen = lds_i32_u32(cons_down);
    }
    float shift = 0.f, adj0 = adj;
    if (apply) {  // apply the shift measured at the end of the previous renorm block
      const int D = Dpend;
      Dpend = 0;
      shift = (float)D;
      if (has_up) {
        adj0 = (float)(Fprev - F);     // step i0 still adds to my old frame (then -shift)
        adj = (float)(Fup - (F + D));  // all later steps of the renorm block
      }
      F += D;
      if (pub) sts_i32_u32(fr_out + (((uint32_t)(i0 >> 4)) & (kFRing - 1)) * 4u, F);
      if (SPLIT && pubg) stg_relaxed_i32(A.gf_out + (i0 >> 4), F);
      if (STORE) {
        if (store_off) *off_p = F;
        off_p += A.NWp;
      }
    }
    if (has_up && !up_g && lane0) sts_i32_u32(cons_mine, i0 + K);  // up[] and the frame offsets are read
    float* const gob = SPLIT ? A.gv_out + i0 + 1 : nullptr;  // step i0+u goes to gv_out[i0+u+1]
    const uint32_t ob0 = vals_out + (((uint32_t)i0 & (kValRing - 1)) + 1u) * 4u;  // steps i0..i0+K-2
    const uint32_t ob1 = vals_out + ((uint32_t)(i0 + K) & (kValRing - 1)) * 4u;    // step i0+K-1
    if (!kEarlyPre) preload();
    if (n == K) {
      // ---------------- K straight-line steps
      uint32_t erow = slot_addr + (DIR ? (uint32_t)(K - 1) * row_bytes : 0u);
#pragma unroll
      for (int u = 0; u < K; ++u) {
        float eb, el[R / 2];
        if (kPre) {
          eb = ebv[u];
#pragma unroll
          for (int h = 0; h < R / 2; ++h) el[h] = elv[u][h];
        } else {
          eb = lds_em(erow);
#pragma unroll
          for (int h = 0; h < R / 2; ++h) el[h] = lds_em(erow + jlab[h]);
          if (DIR) erow -= row_bytes; else erow += row_bytes;
        }
        if (u == 0) {
          eb -= shift;
#pragma unroll
          for (int h = 0; h < R / 2; ++h) el[h] -= shift;
        }
        if (SPLIT && up_g && u == K / 2 && i0 + K < Tb) {
          // the next block's hand-off: loads in flight under the remaining steps of this one
          // (the upstream CTA is normally a block ahead; validated in the next prologue)
#pragma unroll
          for (int j = 0; j < K; j += 4) {
            const float4 v = ldg_relaxed_v4(A.gv_in + i0 + K + j);
            upn[j] = v.x; upn[j + 1] = v.y; upn[j + 2] = v.z; upn[j + 3] = v.w;
          }
          Fupn = ldg_relaxed_i32(A.gf_in + ((i0 + K) >> 4));
          have_next = true;
        }
        step(eb, el, up[u] + (u == 0 ? adj0 : adj), u == K - 1 ? ob1 : ob0 + 4u * u, gob + u);
      }
      __syncwarp();
      if (lane0) mbar_arrive_u32(full_addr + 8u * kMaxEmRing);  // empty[slot]
    } else {
      // ---------------- the last, partial block (at most once per utterance)
      uint32_t erow = slot_addr + (DIR ? (uint32_t)(n - 1) * row_bytes : 0u);
#pragma unroll 1
      for (int u = 0; u < n; ++u) {
        float eb = lds_em(erow), el[R / 2];
#pragma unroll
        for (int h = 0; h < R / 2; ++h) el[h] = lds_em(erow + jlab[h]);
        if (DIR) erow -= row_bytes; else erow += row_bytes;
        const float sh = u == 0 ? shift : 0.f;
        eb -= sh;
#pragma unroll
        for (int h = 0; h < R / 2; ++h) el[h] -= sh;
        float upr = kLogZero;
        if (SPLIT && up_g) {
          if (i0 + u) {
            upr = ldg_relaxed_f32(A.gv_in + i0 + u);
            for (int spin = 0; spin < kBndSpinLimit && !bnd_written(upr); ++spin) {
              upr = ldg_relaxed_f32(A.gv_in + i0 + u);
              if (spin == kBndSpinLimit - 1 && lane0) atomicExch(A.err, kWsErrHandoff);
            }
          }
        } else if (has_up) {
          upr = lds_f32_u32(vb + 4u * u);
        }
        const float upv = has_up ? upr + (u == 0 ? adj0 : adj) : kLogZero;
        step(eb, el, upv, ob0 + 4u * u, gob + u);  // u <= K-2: no wrap
      }
      __syncwarp();
      if (lane0) mbar_arrive_u32(full_addr + 8u * kMaxEmRing);
    }
    // ---------------- block epilogue
    if (pub) sts_release_i32(ready_out, i0 + n);
    if ((i0 & (kRenormEvery - 1)) == kRenormEvery - K && n == K) {
      float m = kLogZero;
#pragma unroll
      for (int q = 0; q < R; ++q) m = fmaxf(m, sp0 + q < S ? a[q] : kLogZero);
      // the shift is an integer anyway: round first, then ONE integer warp reduction
      // (REDUX) instead of a five-deep shuffle chain.  -2^30 stands for "nothing alive".
      constexpr float kDead = -1073741824.f;
      const int mi = __reduce_max_sync(0xffffffffu, __float2int_rn(fminf(fmaxf(m, kDead), 1048576.f)));
      // nothing alive yet: adopt the upstream warp's frame so that the first real
      // values to arrive are already centred.  (Clamped: a pathological -1e10
      // "logzero" logit must not overflow the integer offsets.)  Fup came from a
      // 4-byte broadcast LDS after the acquire, so it is warp-uniform.
      Dpend = (mi > (int)kDead) ? (mi < -1048576 ? -1048576 : mi) : (has_up ? Fup - F : 0);
    }
    slot_addr += chunk_bytes;
    full_addr += 8u;
    if (full_addr == bar_end) { slot_addr = A.ring; full_addr = A.full; phase ^= 1u; }
  }

  // ---- epilogue: log-likelihood from the last two positions (alpha only)
  if (DIR != 0) return;
  float p1 = __shfl_up_sync(0xffffffffu, a[R - 1], 1);
  if (SPLIT && up_g) {
    float v = ldg_relaxed_f32(A.gv_in + Tb);
    for (int spin = 0; spin < kBndSpinLimit && !bnd_written(v); ++spin) {
      v = ldg_relaxed_f32(A.gv_in + Tb);
      if (spin == kBndSpinLimit - 1 && lane0) atomicExch(A.err, kWsErrHandoff);
    }
    const int Fl = Fup;  // the upstream frame of the last renorm block, read in its prologue
    if (lane0) p1 = v + (float)(Fl - F);
  } else if (has_up) {
    while (ready_seen < Tb) ready_seen = lds_acquire_i32(ready_in);
    const float v = lds_f32_u32(vals_in + ((uint32_t)Tb & (kValRing - 1)) * 4u);
    const int Fl = lds_i32_u32(fr_in + (((uint32_t)((Tb - 1) >> 4)) & (kFRing - 1)) * 4u);
    if (lane0) p1 = v + (float)(Fl - F);
  } else if (lane0) {
    p1 = kLogZero;
  }
  const int q_own = (S - 1) % R;
  if (sp0 + q_own == S - 1) {
    float v1 = kLogZero, v2 = kLogZero;
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (q == q_own) {
        v1 = a[q];
        v2 = q >= 1 ? a[q - 1] : p1;
      }
    }
    if (S == 1) v2 = kLogZero;
    const double l1 = (double)v1, l2 = (double)v2;
    const double m = fmax(l1, l2);
    const bool inf = !(m > (double)kLogZeroTest);
    const double lam = inf ? (double)kNegInf : m + log2(exp2(l1 - m) + exp2(l2 - m));
    A.utt[A.b].lam = lam;
    A.utt[A.b].aend = F;
    A.utt[A.b].infinite = inf;
    A.utt[A.b].zeroed = inf && A.zero_infinity;
  }
}

struct SplitArgs {  // split lattices only (ctc_ws.cuh LossLayout::CL > 1)
  int CL, Lps, Tb1;
  float* bnd_val;
  int* bnd_f;
  int* bnd_ready;
};

template <int R, bool STORE, int K, bool SPLIT>
__global__ void __launch_bounds__(K == 16 ? 288 : 544, 1) lattice_kernel(
    UttInfo* __restrict__ utt, const int* __restrict__ lab, int Lp, const float* __restrict__ em,
    float* __restrict__ alpha, float* __restrict__ beta, int* __restrict__ offA,
    int* __restrict__ offB, int Tfr, int Sp, int NWp, int NB, int zero_infinity, int nring, int pdl,
    SplitArgs sa, int* __restrict__ err, int lat_pol, int* __restrict__ tickets) {
  static_assert(R % 2 == 0, "R must be even: even recursion positions are blanks");
  extern __shared__ __align__(128) float lsm[];  // em ring [nring][K*Lp] | vals | fring | ready | cons
  __shared__ __align__(8) uint64_t bars[2 * kMaxEmRing];  // full[], empty[]
  uint64_t* const full = bars;
  uint64_t* const empty = bars + kMaxEmRing;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int NW = (blockDim.x >> 5) - 1;  // compute warps; warp NW is the TMA producer
  // split: CTA cr of CL handles warps [cr*NW, (cr+1)*NW) of utterance b's lattice.  A downstream
  // CTA spin-waits on its upstream neighbour, so the logical position comes from a TICKET taken
  // in dispatch order (one counter per direction, zeroed before the launch), not from
  // blockIdx: whatever order the hardware starts CTAs in, a CTA only ever waits for CTAs that
  // took an earlier ticket, i.e. that are already running or done (decoupled look-back's rule).
  const int dir = blockIdx.y;
  int bx = blockIdx.x;
  if (SPLIT) {
    __shared__ int s_ticket;
    if (tid == 0) s_ticket = atomicAdd(tickets + dir, 1);
    __syncthreads();
    bx = s_ticket;
  }
  const int b = SPLIT ? bx / sa.CL : bx;
  const int cr = SPLIT ? bx - b * sa.CL : 0;
  if (pdl == 0) griddep_launch();  // PDL: let the gradient row pass become resident behind us
  griddep_wait();                  // utt / lab / em come from the previous kernel
  const int Tb = utt[b].T, L = utt[b].L, S = 2 * L + 1;
  if (Tb == 0) {
    if (dir == 0 && tid == 0 && cr == 0) {
      const bool inf = L > 0;
      utt[b].lam = inf ? (double)kNegInf : 0.0;
      utt[b].aend = 0;
      utt[b].infinite = inf;
      utt[b].zeroed = inf && zero_infinity;
    }
    return;
  }
  const int nw_live = (S + 32 * R - 1) / (32 * R);  // over the whole lattice
  const int W0 = cr * NW;
  const int nw_here = SPLIT ? (nw_live - W0 < NW ? nw_live - W0 : NW) : nw_live;  // live warps of this CTA
  if (SPLIT && nw_here <= 0) return;  // this utterance's lattice ends in an earlier CTA
  // split: the label columns this CTA's positions [P0, P1) need, as a 16-byte aligned range
  int col0 = 0, ncols = 0;
  if (SPLIT) {
    const int P0 = W0 * 32 * R;
    const int P1 = (P0 + NW * 32 * R < S) ? P0 + NW * 32 * R : S;
    const int s_lo = dir ? S - P1 : P0, s_hi = dir ? S - 1 - P0 : P1 - 1;  // lattice indices
    const int k_min = (s_lo + 2) >> 1, k_max = (s_hi + 1) >> 1;           // odd s <-> label (s+1)/2
    if (k_max >= k_min) {
      col0 = (kEmLabel0 + k_min - 1) & ~3;
      ncols = ((kEmLabel0 + k_max - 1 + 1 - col0) + 3) & ~3;
    }
  }
  const int row_elems = SPLIT ? sa.Lps : Lp;  // staged emission row
  const int chunk_elems = K * row_elems;
  float* ring = lsm;
  float* vals = lsm + (size_t)nring * chunk_elems;     // [NW][kValRing]
  int* fring = reinterpret_cast<int*>(vals + (size_t)NW * kValRing);  // [NW][kFRing]
  int* ready = fring + (size_t)NW * kFRing;            // [NW]
  int* cons = ready + NW;                              // [NW]
  const float* emb = em + (size_t)b * Tfr * Lp;

  if (tid == 0) {
    for (int i = 0; i < nring; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], nw_here); }
    fence_mbar_init();
  }
  for (int i = tid; i < NW * kValRing; i += blockDim.x) vals[i] = kLogZero;
  for (int i = tid; i < NW * (kFRing + 2); i += blockDim.x) fring[i] = 0;  // fring, ready, cons
  __syncthreads();

  // ------------------------------------------------------------ TMA producer warp
  if (SPLIT && wid == NW) {
    // partial rows: lane l copies frame l of the chunk (blank group + this CTA's columns)
    const int nchunks = (Tb + K - 1) / K;
    for (int c = 0; c < nchunks; ++c) {
      const int slot = c % nring, round = c / nring;
      if (round > 0 && lane == 0) mbar_wait(&empty[slot], (round - 1) & 1);
      const int i0 = c * K;
      const int n = (Tb - i0 < K) ? Tb - i0 : K;
      const int f0 = dir ? Tb - i0 - n : i0;
      if (lane == 0) mbar_expect_tx(&full[slot], (uint32_t)n * (16u + 4u * (uint32_t)ncols));
      __syncwarp();
      if (lane < n) {
        float* dst = ring + (size_t)slot * chunk_elems + (size_t)lane * row_elems;
        const float* src = emb + (size_t)(f0 + lane) * Lp;
        bulk_g2s(dst, src, 16u, &full[slot]);
        if (ncols) bulk_g2s(dst + 4, src + col0, 4u * (uint32_t)ncols, &full[slot]);
      }
    }
    return;
  }
  if (wid == NW) {
    if (lane == 0) {
      const int nchunks = (Tb + K - 1) / K;
      for (int c = 0; c < nchunks; ++c) {
        const int slot = c % nring, round = c / nring;
        if (round > 0) mbar_wait(&empty[slot], (round - 1) & 1);
        const int i0 = c * K;
        const int n = (Tb - i0 < K) ? Tb - i0 : K;
        const int f0 = dir ? Tb - i0 - n : i0;  // first frame of the (contiguous) block
        const uint32_t bytes = (uint32_t)n * Lp * 4u;
        mbar_expect_tx(&full[slot], bytes);
        bulk_g2s(ring + (size_t)slot * chunk_elems, emb + (size_t)f0 * Lp, bytes, &full[slot]);
      }
    }
    return;
  }
  if (wid >= nw_here) return;

  LatticeArgs A;
  A.utt = utt;
  A.lb = lab + (size_t)b * Lp;
  A.lat = (dir ? beta : alpha) + (size_t)b * Tfr * Sp;
  A.off = (dir ? offB : offA) + (size_t)b * NB * NWp;
  A.ring = smem_u32(ring);
  A.vals = smem_u32(vals);
  A.fring = smem_u32(fring);
  A.ready = smem_u32(ready);
  A.cons = smem_u32(cons);
  A.full = smem_u32(full);
  A.Tb = Tb; A.L = L; A.Lp = Lp; A.Sp = Sp; A.NWp = NWp;
  A.nw_live = nw_live; A.zero_infinity = zero_infinity; A.b = b; A.nring = nring;
  A.W0 = W0; A.Lps = row_elems; A.col0 = col0; A.nwc = NW;
  A.gv_in = A.gv_out = nullptr; A.gf_in = A.gf_out = nullptr; A.gr_in = A.gr_out = nullptr;
  A.err = err;
  A.lat_policy = lat_pol;
  if (SPLIT) {
    // boundary j (between CTA j and j+1) of (b, dir): index ((b*2 + dir)*(CL-1) + j)
    const size_t base = ((size_t)b * 2 + dir) * (size_t)(sa.CL - 1);
    if (cr > 0) {
      A.gv_in = sa.bnd_val + (base + cr - 1) * sa.Tb1;
      A.gf_in = sa.bnd_f + (base + cr - 1) * NB;
      A.gr_in = sa.bnd_ready + (base + cr - 1) * kBndReadyStride;
    }
    if (cr + 1 < sa.CL) {
      A.gv_out = sa.bnd_val + (base + cr) * sa.Tb1;
      A.gf_out = sa.bnd_f + (base + cr) * NB;
      A.gr_out = sa.bnd_ready + (base + cr) * kBndReadyStride;
    }
  }
  if (dir == 0) lattice_warp<R, STORE, 0, K, SPLIT>(A, wid, lane);
  else lattice_warp<R, STORE, 1, K, SPLIT>(A, wid, lane);
  if (pdl == 1 && wid == nw_here - 1) griddep_launch();  // the last warp of the pipeline is done
}

// nll[b] = -ln2 * (lam + aend + sum_t (c[t] - lse2[t]))   (ctc_ws.cuh)
// Optionally the CTA that finishes last also writes the batch reduction
// loss_out[0] = loss_scale * sum_b nll[b] (fixed summation order: deterministic),
// i.e. espnet2/asr/ctc.py:56-60 `loss.sum() / size` without further launches.
__global__ void __launch_bounds__(256) nll_kernel(const UttInfo* __restrict__ utt,
                                                  const float* __restrict__ cmax,
                                                  const float* __restrict__ lse2, int Tfr,
                                                  int zero_infinity, float* __restrict__ nll,
                                                  float* __restrict__ loss_out, float loss_scale,
                                                  int* __restrict__ counter,
                                                  const int* __restrict__ err, PeerDev peer,
                                                  float* __restrict__ global_out) {
  __shared__ double dred[8];
  __shared__ int s_last;
  const int b = blockIdx.x, tid = threadIdx.x;
  griddep_wait();
  const UttInfo u = utt[b];
  double part = 0.0;
  for (int t = tid; t < u.T; t += 256) {
    const size_t g = (size_t)b * Tfr + t;
    part += (double)cmax[g] - (double)lse2[g];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if ((tid & 31) == 0) dred[tid >> 5] = part;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
    for (int i = 0; i < 8; ++i) tot += dred[i];
    float v;
    if (u.infinite) v = zero_infinity ? 0.f : -kNegInf;
    else v = (float)(-(u.lam + (double)u.aend + tot) * 0.6931471805599453);
    // a kernel of this call gave up waiting for another CTA (ctc_ws.cuh kWsErr*): the
    // lattices are incomplete - fail loudly
    if (__ldcg(err) != 0) v = __int_as_float(0x7fc00000);
    nll[b] = v;
    s_last = 0;
    if (loss_out || peer.bufs) {
      __threadfence();
      s_last = atomicAdd(counter, 1) == (int)gridDim.x - 1;
    }
  }
  if (!loss_out && !peer.bufs) return;
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  double acc = 0.0;
  for (int i = tid; i < (int)gridDim.x; i += 256) acc += (double)__ldcg(nll + i);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((tid & 31) == 0) dred[tid >> 5] = acc;
  __syncthreads();
  if (tid == 0) {
    double tot = 0.0;
    for (int i = 0; i < 8; ++i) tot += dred[i];
    if (loss_out) loss_out[0] = (float)(tot * (double)loss_scale);
    dred[0] = tot;
    *counter = 0;  // ready for the next call on this workspace
  }
  if (!peer.bufs) return;
  // sharded training: the all-reduce of {sum of losses, utterance count} over the ranks
  // is the tail of this kernel (peer.cuh) - pushed to the peers through NVLink, no launch
  __syncthreads();
  if (tid < 32) peer_exchange_warp(peer, (float)dred[0], (float)gridDim.x, global_out);
}

template <int R, int K>
static int launch_lattice_rk(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                             int both_dirs, int nring, size_t smem, cudaStream_t st) {
  const int nt = (w.NW + 1) * 32;  // + the TMA producer warp
  dim3 grid(w.B, both_dirs ? 2 : 1);
  const SplitArgs sa{1, 0, 0, nullptr, nullptr, nullptr};
  auto go = [&](auto k) {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    launch_pdl(k, grid, dim3(nt), smem, st, 0, reinterpret_cast<UttInfo*>(ws + w.utt),
               reinterpret_cast<const int*>(ws + w.lab), w.Lp,
               reinterpret_cast<const float*>(ws + w.em), reinterpret_cast<float*>(ws + w.alpha),
               reinterpret_cast<float*>(ws + w.beta), reinterpret_cast<int*>(ws + w.offA),
               reinterpret_cast<int*>(ws + w.offB), Tfr, w.Sp, w.NW, w.NB, zero_infinity, nring,
               pdl_mode(), sa, reinterpret_cast<int*>(ws + w.err), lat_policy_host(),
               static_cast<int*>(nullptr));
  };
  if (both_dirs) go(lattice_kernel<R, true, K, false>);
  else go(lattice_kernel<R, false, K, false>);
  count_launch();
  return check_launch("ctc lattice");
}

// Split lattice (LossLayout::CL > 1): R = 4, kSplitWarps compute warps + the producer per CTA,
// 16-step blocks, CL CTAs per (utterance, direction).  blockIdx.x = b*CL + cr: an upstream CTA
// always has the smaller index, is dispatched first and never waits on a downstream CTA (the
// hand-off arrays hold all T steps: no back-pressure), so the wait of a downstream CTA on its
// upstream neighbour cannot deadlock whatever the number of resident CTAs.
static int launch_lattice_split(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                                int both_dirs, cudaStream_t st) {
  constexpr int R = 4, K = 16;
  const int lps = 4 + (int)align_up((size_t)w.NWc * 16 * R + 6, 4);  // blank group + columns (+ alignment)
  const int nring = w.NWc + 2;
  const size_t fixed = (size_t)w.NWc * (kValRing * 4 + kFRing * 4 + 8) + 256;
  const size_t smem = (size_t)nring * K * lps * 4 + fixed;
  B2_REQUIRE(smem <= kLatticeSmemBudget, "ctc lattice (split): shared memory %zu", smem);
  const size_t nb = (size_t)w.B * 2 * (size_t)(w.CL - 1);
  // sentinel fill of the hand-off arrays (bnd_val and bnd_f are adjacent sections)
  if (cudaMemsetAsync(ws + w.bnd_val, 0xff, 4 * nb * w.Tb1, st) != cudaSuccess ||
      cudaMemsetAsync(ws + w.bnd_f, 0x80, 4 * nb * w.NB, st) != cudaSuccess ||
      cudaMemsetAsync(ws + w.split_tickets, 0, 2 * sizeof(int), st) != cudaSuccess) {
    set_error("ctc lattice (split): memset failed");
    return B200CTC_ERR_DEVICE;
  }
  const SplitArgs sa{w.CL, lps, w.Tb1, reinterpret_cast<float*>(ws + w.bnd_val),
                     reinterpret_cast<int*>(ws + w.bnd_f), reinterpret_cast<int*>(ws + w.bnd_ready)};
  const int nt = (w.NWc + 1) * 32;
  dim3 grid(w.B * w.CL, both_dirs ? 2 : 1);
  auto go = [&](auto k) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    launch_pdl(k, grid, dim3(nt), smem, st, 0, reinterpret_cast<UttInfo*>(ws + w.utt),
               reinterpret_cast<const int*>(ws + w.lab), w.Lp,
               reinterpret_cast<const float*>(ws + w.em), reinterpret_cast<float*>(ws + w.alpha),
               reinterpret_cast<float*>(ws + w.beta), reinterpret_cast<int*>(ws + w.offA),
               reinterpret_cast<int*>(ws + w.offB), Tfr, w.Sp, w.NW, w.NB, zero_infinity, nring,
               pdl_mode(), sa, reinterpret_cast<int*>(ws + w.err), lat_policy_host(),
               reinterpret_cast<int*>(ws + w.split_tickets));
  };
  if (both_dirs) go(lattice_kernel<R, true, K, true>);
  else go(lattice_kernel<R, false, K, true>);
  count_launch();
  return check_launch("ctc lattice (split)");
}

template <int R>
static int launch_lattice_r(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                            int both_dirs, cudaStream_t st) {
  // shared memory: em ring of nring chunks of K frames + the per-warp hand-off rings.
  // Warp w+1 runs one block behind warp w, so the ring must span NW+1 blocks or the
  // first warp stalls on the last one (make_loss_layout picks R accordingly).
  B2_REQUIRE(w.NW <= 16, "ctc lattice: target too long (more than 16 warps of 32 positions/thread)");
  const size_t fixed = (size_t)w.NW * (kValRing * 4 + kFRing * 4 + 8) + 256;
  const size_t budget = kLatticeSmemBudget - fixed;
  const size_t frame = (size_t)w.Lp * 4;
  int K = 8;
  if ((size_t)(w.NW + 1) * 8 * frame > budget) K = 4;
  // short lattices: 16-step blocks halve the per-block hand-shaking (needs the registers
  // of a <= 8-warp CTA; the pipeline skew between the warps grows to 16 steps)
  static const int k16 = tuning_int("B200CTC_LAT_K16", 1);
  if (k16 && R <= 4 && w.NW <= 8 && (size_t)(w.NW + 3) * 16 * frame <= budget) K = 16;
  B2_REQUIRE(2 * (size_t)K * frame <= budget, "ctc lattice: target too long for shared memory");
  int nring = (int)(budget / (K * frame));
  if (nring > kMaxEmRing) nring = kMaxEmRing;
  // keep small cases small: ~32 KB of emissions in flight is plenty
  static const int slack = tuning_int("B200CTC_LAT_SLACK", 2);
  while (nring > w.NW + slack && nring > 3 && (size_t)nring * K * frame > 32 * 1024) --nring;
  const size_t smem = (size_t)nring * K * frame + fixed;
  if (K == 16) {
    if constexpr (R <= 4)
      return launch_lattice_rk<R, 16>(w, ws, Tfr, zero_infinity, both_dirs, nring, smem, st);
  }
  if (K == 8) return launch_lattice_rk<R, 8>(w, ws, Tfr, zero_infinity, both_dirs, nring, smem, st);
  return launch_lattice_rk<R, 4>(w, ws, Tfr, zero_infinity, both_dirs, nring, smem, st);
}

static inline int launch_lattice(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                                 int both_dirs, cudaStream_t st) {
  if (w.CL > 1) return launch_lattice_split(w, ws, Tfr, zero_infinity, both_dirs, st);
  switch (w.R) {
    case 2: return launch_lattice_r<2>(w, ws, Tfr, zero_infinity, both_dirs, st);
    case 4: return launch_lattice_r<4>(w, ws, Tfr, zero_infinity, both_dirs, st);
    case 8: return launch_lattice_r<8>(w, ws, Tfr, zero_infinity, both_dirs, st);
    case 16: return launch_lattice_r<16>(w, ws, Tfr, zero_infinity, both_dirs, st);
    default: return launch_lattice_r<32>(w, ws, Tfr, zero_infinity, both_dirs, st);
  }
}

static inline int launch_nll(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                             float* nll, float* loss_out, float loss_scale, cudaStream_t st,
                             PeerDev peer = PeerDev{nullptr, 0, 0, 0, 0}, float* global_out = nullptr) {
  launch_pdl(nll_kernel, dim3(w.B), dim3(256), 0, st, 2,
             reinterpret_cast<const UttInfo*>(ws + w.utt),
             reinterpret_cast<const float*>(ws + w.cmax),
             reinterpret_cast<const float*>(ws + w.lse2), Tfr, zero_infinity, nll, loss_out,
             loss_scale, reinterpret_cast<int*>(ws + w.counter),
             reinterpret_cast<const int*>(ws + w.err), peer, global_out);
  count_launch();
  return check_launch("ctc nll");
}

}  // namespace b200ctc
