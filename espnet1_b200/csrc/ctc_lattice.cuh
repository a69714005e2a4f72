// Alpha/beta lattice recursion (Graves 2006 eq. 6-11) in normalised log2 space.
//
// One CTA per (utterance, direction): blockIdx.y = 0 runs alpha forward in
// time, 1 runs beta backward in time as "alpha on the reversed lattice"
// (recursion position sp <-> lattice index s = S-1-sp).
//
// The CTA is a SYSTOLIC ARRAY OF WARPS with no CTA-wide barrier on the serial
// chain.  Thread i keeps R consecutive lattice positions in registers; inside a
// warp the neighbour value it needs arrives by shuffle; between warps they
// travel through a small shared-memory ring of 16-byte packets {-, a[-1], step
// tag, frame offset} written by the last lane of warp w and read by warp
// w+1 one step later.  The dependency is one-directional (position sp needs
// sp-1, sp-2), so warp w+1 simply trails warp w by a step or two; a consumed
// counter gives back-pressure on the ring.  Every warp re-centres its own slice
// on its own maximum every kRenormEvery frames with an INTEGER shift (exact to
// combine, see ctc_ws.cuh), so there is no reduction across warps either.
// The emissions em[t][k] are streamed into a shared-memory ring by a dedicated
// producer warp with the TMA engine (1-D bulk copies of frame chunks, full /
// empty mbarriers).
#pragma once
#include "ctc_ws.cuh"

namespace b200ctc {

constexpr int kMaxLattice = 8192;  // 2*L+1 <= 8192  (R = 16, 16 warps + producer)
constexpr int kRenormEvery = 16;
constexpr int kEmRing = 4;    // chunks of emission frames in flight
constexpr int kPktRing = 16;  // boundary packets in flight between adjacent warps

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ float4 lds_volatile_v4(const float4* p) {
  float4 v;
  asm volatile("ld.volatile.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ void sts_volatile_v4(float4* p, float4 v) {
  asm volatile("st.volatile.shared.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(smem_u32(p)), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}
__device__ __forceinline__ int lds_volatile_i32(const int* p) {
  int v;
  asm volatile("ld.volatile.shared.s32 %0, [%1];" : "=r"(v) : "r"(smem_u32(p)));
  return v;
}
__device__ __forceinline__ void sts_volatile_i32(int* p, int v) {
  asm volatile("st.volatile.shared.s32 [%0], %1;" ::"r"(smem_u32(p)), "r"(v) : "memory");
}

// log2(2^a + 2^b), 2 MUFU; -inf safe
__device__ __forceinline__ float lse2_log2(float a, float b) {
  const float m = fmaxf(a, b), n = fminf(a, b);
  const float ms = (m == kNegInf) ? 0.f : m;
  return m + lg2(1.f + ex2(n - ms));
}
// log2(2^a + 2^b + 2^c), 3 MUFU (the largest term is exactly 1); -inf safe
__device__ __forceinline__ float lse3_log2(float a, float b, float c) {
  const float hi = fmaxf(a, b), lo = fminf(a, b);
  const float m = fmaxf(hi, c), t = fminf(hi, c);
  const float mid = fmaxf(lo, t), mn = fminf(lo, t);
  const float ms = (m == kNegInf) ? 0.f : m;
  return m + lg2(1.f + ex2(mid - ms) + ex2(mn - ms));
}

template <int R, bool STORE>
__global__ void __launch_bounds__(R == 2 ? 544 : (R <= 8 ? 1024 : 576)) lattice_kernel(
    UttInfo* __restrict__ utt, const int* __restrict__ lab, int Lp, const float* __restrict__ em,
    float* __restrict__ alpha, float* __restrict__ beta, int* __restrict__ offA,
    int* __restrict__ offB, int Tfr, int Sp, int NWp, int zero_infinity, int chunk) {
  static_assert(R % 2 == 0, "R must be even: even recursion positions are blanks");
  extern __shared__ __align__(128) float lsm[];  // em ring [kEmRing][chunk*Lp], packet ring
  __shared__ __align__(8) uint64_t full[kEmRing], empty[kEmRing];
  __shared__ int consumed[32];
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int NW = (blockDim.x >> 5) - 1;  // compute warps; warp NW is the TMA producer
  const int b = blockIdx.x, dir = blockIdx.y;
  const int Tb = utt[b].T, L = utt[b].L, S = 2 * L + 1;
  if (Tb == 0) {
    if (dir == 0 && tid == 0) {
      const bool inf = L > 0;
      utt[b].lam = inf ? (double)kNegInf : 0.0;
      utt[b].aend = 0;
      utt[b].infinite = inf;
      utt[b].zeroed = inf && zero_infinity;
    }
    return;
  }
  const int nw_live = (S + 32 * R - 1) / (32 * R);
  const int chunk_elems = chunk * Lp;
  float* ring = lsm;
  float4* pkt = reinterpret_cast<float4*>(lsm + (size_t)kEmRing * chunk_elems);  // [NW][kPktRing]
  const float* emb = em + (size_t)b * Tfr * Lp;
  const int nchunks = (Tb + chunk - 1) / chunk;

  if (tid == 0) {
    for (int i = 0; i < kEmRing; ++i) { mbar_init(&full[i], 1); mbar_init(&empty[i], nw_live); }
    fence_mbar_init();
  }
  for (int i = tid; i < NW * kPktRing; i += blockDim.x)
    pkt[i] = make_float4(kNegInf, kNegInf, __int_as_float(-1), __int_as_float(0));
  if (tid < 32) consumed[tid] = 0;
  __syncthreads();

  // ------------------------------------------------------------ TMA producer warp
  if (wid == NW) {
    if (lane == 0) {
      for (int c = 0; c < nchunks; ++c) {
        const int slot = c % kEmRing, round = c / kEmRing;
        if (round > 0) mbar_wait(&empty[slot], (round - 1) & 1);
        const int i0 = c * chunk;
        const int n = (Tb - i0 < chunk) ? Tb - i0 : chunk;
        const int f0 = dir ? Tb - i0 - n : i0;  // first frame of the (contiguous) block
        const uint32_t bytes = (uint32_t)n * Lp * 4u;
        mbar_expect_tx(&full[slot], bytes);
        bulk_g2s(ring + (size_t)slot * chunk_elems, emb + (size_t)f0 * Lp, bytes, &full[slot]);
      }
    }
    return;
  }
  if (wid >= nw_live) return;

  // ------------------------------------------------------------ compute warps
  const int w = wid;
  const int sp0 = (w * 32 + lane) * R;  // first recursion position of this thread
  const int* lb = lab + (size_t)b * Lp;
  // static description of this thread's positions: q even = blank, q odd = label
  int jlab[R / 2];     // index into the em row of the label emission
  bool skip[R / 2];    // may the path skip the blank in front of this label?
  bool act[R];
#pragma unroll
  for (int q = 0; q < R; ++q) act[q] = sp0 + q < S;
#pragma unroll
  for (int h = 0; h < R / 2; ++h) {
    const int sp = sp0 + 2 * h + 1;
    const int s = dir ? S - 1 - sp : sp;
    const int k = (s + 1) >> 1;  // 1..L when active
    bool sk = false;
    if (sp < S) {
      if (dir == 0) sk = (k >= 2) && (lb[k] != lb[k - 1]);
      else sk = (k + 1 <= L) && (lb[k] != lb[k + 1]);
    }
    skip[h] = sk;
    jlab[h] = kEmLabel0 + ((sp < S) ? k - 1 : 0);
  }
  float* lat = nullptr;
  int* off = nullptr;
  if (STORE) {
    lat = (dir ? beta : alpha) + (size_t)b * Tfr * Sp + sp0;
    off = (dir ? offB : offA) + (size_t)b * Tfr * NWp + w;
  }
  const bool store_ok = sp0 < Sp;
  const float4* pk_in = pkt + (size_t)(w > 0 ? w - 1 : 0) * kPktRing;
  float4* pk_out = pkt + (size_t)w * kPktRing;
  const bool has_up = w > 0, has_down = w + 1 < nw_live;

  float a[R];
#pragma unroll
  for (int q = 0; q < R; ++q) a[q] = kNegInf;
  int F = 0;       // integer frame offset of a[] (ctc_ws.cuh)
  int Dpend = 0;   // shift to apply at the next step
  int upF = 0;     // frame offset carried by the last upstream packet
  int cons_seen = 0;  // downstream warp's consumed count (cached)

  // fetch the value of position sp0-1 at step `i-1`, expressed in MY frame.
  // (R is even, so sp0 is a blank and needs only sp0-1; the label at sp0+1 takes
  // its skip transition from sp0-1 as well: ONE neighbour value per thread.)
  auto fetch = [&](int i) {
    float p1 = __shfl_up_sync(0xffffffffu, a[R - 1], 1);
    if (has_up) {
      float4 pk;
      do {
        pk = lds_volatile_v4(&pk_in[(i - 1) % kPktRing]);
      } while (__float_as_int(pk.z) != i - 1);
      upF = __float_as_int(pk.w);
      if (lane == 0) {
        p1 = pk.y + (float)(upF - F);
        sts_volatile_i32(&consumed[w], i);
      }
    } else if (lane == 0) {
      p1 = kNegInf;
    }
    return p1;
  };

  for (int c = 0; c < nchunks; ++c) {
    const int i0 = c * chunk;
    const int n = (Tb - i0 < chunk) ? Tb - i0 : chunk;
    const int slot = c % kEmRing;
    const float* ech = ring + (size_t)slot * chunk_elems;
    mbar_wait(&full[slot], (c / kEmRing) & 1);
    for (int u = 0; u < n; ++u) {
      const int i = i0 + u;
      const int t = dir ? Tb - 1 - i : i;
      const float* erow = ech + (size_t)(dir ? n - 1 - u : u) * Lp;
      const float shift = (float)Dpend;
      const float eb = erow[0] - shift;
      float el[R / 2];
#pragma unroll
      for (int h = 0; h < R / 2; ++h) el[h] = erow[jlab[h]] - shift;
      float nw_[R];
      if (i == 0) {
#pragma unroll
        for (int q = 0; q < R; ++q) {
          const float e = (q & 1) ? el[q >> 1] : eb;
          nw_[q] = (sp0 + q < 2 && act[q]) ? e : kNegInf;
        }
      } else {
        const float p1 = fetch(i);
#pragma unroll
        for (int q = 0; q < R; ++q) {
          const float x0 = a[q];
          const float x1 = q >= 1 ? a[q - 1] : p1;
          float v;
          if (q & 1) {
            const float x2r = q >= 2 ? a[q - 2] : p1;
            const float x2 = skip[q >> 1] ? x2r : kNegInf;
            v = lse3_log2(x0, x1, x2) + el[q >> 1];
          } else {
            v = lse2_log2(x0, x1) + eb;
          }
          nw_[q] = act[q] ? v : kNegInf;
        }
      }
      F += Dpend;
      Dpend = 0;
#pragma unroll
      for (int q = 0; q < R; ++q) a[q] = nw_[q];
      // hand the top two positions to the next warp
      if (has_down) {
        const int need = i - kPktRing + 1;  // the slot's previous packet must be consumed
        if (need > cons_seen) {
          do { cons_seen = lds_volatile_i32(&consumed[w + 1]); } while (cons_seen < need);
        }
        if (lane == 31)
          sts_volatile_v4(&pk_out[i % kPktRing],
                          make_float4(0.f, a[R - 1], __int_as_float(i), __int_as_float(F)));
      }
      if (STORE) {
        if (store_ok) {
          float* dst = lat + (size_t)t * Sp;
          if (R == 2) {
            *reinterpret_cast<float2*>(dst) = make_float2(a[0], a[1]);
          } else {
#pragma unroll
            for (int q = 0; q < R; q += 4)
              *reinterpret_cast<float4*>(dst + q) = make_float4(a[q], a[q + 1], a[q + 2], a[q + 3]);
          }
        }
        if (lane == 0) off[(size_t)t * NWp] = F;
      }
      if ((i % kRenormEvery) == kRenormEvery - 1) {
        float m = a[0];
#pragma unroll
        for (int q = 1; q < R; ++q) m = fmaxf(m, a[q]);
        m = warp_max(m);
        // all -inf: adopt the upstream warp's frame so that the first finite
        // values to arrive are already centred
        // (clamped: a pathological -1e10 "logzero" logit must not overflow the offsets)
        Dpend = (m > kNegInf) ? (int)rintf(fminf(fmaxf(m, -1048576.f), 1048576.f))
                              : (has_up ? upF - F : 0);
      }
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&empty[slot]);
  }

  // ---- epilogue: log-likelihood from the last two positions (alpha only)
  if (dir != 0) return;
  const float p1 = fetch(Tb);
  const int q_own = (S - 1) % R;
  if (sp0 + q_own == S - 1) {
    float v1 = kNegInf, v2 = kNegInf;
#pragma unroll
    for (int q = 0; q < R; ++q) {
      if (q == q_own) {
        v1 = a[q];
        v2 = q >= 1 ? a[q - 1] : p1;
      }
    }
    if (S == 1) v2 = kNegInf;
    const double l1 = (double)v1, l2 = (double)v2;
    const double m = fmax(l1, l2);
    double lam;
    if (m == (double)kNegInf) lam = m;
    else lam = m + log2(exp2(l1 - m) + exp2(l2 - m));
    const bool inf = (lam == (double)kNegInf);
    utt[b].lam = lam;
    utt[b].aend = F;
    utt[b].infinite = inf;
    utt[b].zeroed = inf && zero_infinity;
  }
}

// nll[b] = -ln2 * (lam + aend + sum_t (c[t] - lse2[t]))   (ctc_ws.cuh)
__global__ void __launch_bounds__(256) nll_kernel(const UttInfo* __restrict__ utt,
                                                  const float* __restrict__ cmax,
                                                  const float* __restrict__ lse2, int Tfr,
                                                  int zero_infinity, float* __restrict__ nll) {
  __shared__ double dred[8];
  const int b = blockIdx.x, tid = threadIdx.x;
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
    nll[b] = v;
  }
}

template <int R>
static int launch_lattice_r(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                            int both_dirs, cudaStream_t st) {
  const int nt = (w.NW + 1) * 32;
  int chunk = 8192 / (w.Lp * 4);
  if (chunk < 1) chunk = 1;
  if (chunk > 32) chunk = 32;
  const size_t smem = sizeof(float) * (size_t)kEmRing * chunk * w.Lp +
                      sizeof(float4) * (size_t)w.NW * kPktRing;
  dim3 grid(w.B, both_dirs ? 2 : 1);
  auto go = [&](auto k) {
    if (smem > 48 * 1024)
      cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<grid, nt, smem, st>>>(reinterpret_cast<UttInfo*>(ws + w.utt),
                              reinterpret_cast<const int*>(ws + w.lab), w.Lp,
                              reinterpret_cast<const float*>(ws + w.em),
                              reinterpret_cast<float*>(ws + w.alpha),
                              reinterpret_cast<float*>(ws + w.beta),
                              reinterpret_cast<int*>(ws + w.offA),
                              reinterpret_cast<int*>(ws + w.offB), Tfr, w.Sp, w.NW, zero_infinity,
                              chunk);
  };
  if (both_dirs) go(lattice_kernel<R, true>);
  else go(lattice_kernel<R, false>);
  count_launch();
  return check_launch("ctc lattice");
}

static inline int launch_lattice(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                                 int both_dirs, cudaStream_t st) {
  switch (w.R) {
    case 2: return launch_lattice_r<2>(w, ws, Tfr, zero_infinity, both_dirs, st);
    case 4: return launch_lattice_r<4>(w, ws, Tfr, zero_infinity, both_dirs, st);
    case 8: return launch_lattice_r<8>(w, ws, Tfr, zero_infinity, both_dirs, st);
    default: return launch_lattice_r<16>(w, ws, Tfr, zero_infinity, both_dirs, st);
  }
}

static inline int launch_nll(const LossLayout& w, const unsigned char* ws, int Tfr,
                             int zero_infinity, float* nll, cudaStream_t st) {
  nll_kernel<<<w.B, 256, 0, st>>>(reinterpret_cast<const UttInfo*>(ws + w.utt),
                                  reinterpret_cast<const float*>(ws + w.cmax),
                                  reinterpret_cast<const float*>(ws + w.lse2), Tfr, zero_infinity,
                                  nll);
  count_launch();
  return check_launch("ctc nll");
}

}  // namespace b200ctc
