// Fused forward of the eager CTC step: label prep + sparse emission gather + alpha/beta
// lattice in ONE launch, with the SMs partitioned by role.
//
// Before: gather (HBM-bound on 64-byte DRAM atoms, 74 us at C2) -> lattice (latency-bound,
// 58 us, 22 MB of DRAM traffic) ran back to back: the memory system idles under the lattice
// and 128 SMs wait for the gather.  Co-residency does not fix that: a lattice warp that shares
// its SM's load/store queue with gather loads runs 2x slower (profiles/README.md, "wave
// pipeline").  Here the two run side by side on DISJOINT SMs:
//
//   * lattice CTAs (one per utterance): alpha AND beta of the utterance in one CTA - two
//     warp-systolic lattices (ctc_lattice.cuh) with their own rings, barriers and TMA
//     producer warp each; the CTA also does the utterance's label prep.  B CTAs, one per SM
//     (shared memory), so B SMs instead of 2B.
//   * gather CTAs (the remaining SMs): persistent; warps walk (frame group, utterance) items
//     OUTSIDE-IN in time - group 0 and the last group of every utterance first - because alpha
//     consumes frames ascending and beta descending; every finished quarter group (4 frames)
//     bumps a per-(utterance, group) counter (release), the lattice's TMA producer polls it
//     (acquire + proxy fence) before it copies the chunk into its ring.
//
// Roles are taken by an atomic ticket in dispatch order - the first G CTAs to start become
// gather CTAs, which wait for nobody - so a lattice CTA only ever waits for CTAs that are
// already running or finished: forward progress does not depend on co-residency.
#pragma once
#include "ctc_lattice.cuh"

namespace b200ctc {

constexpr int kFusedK = 16;           // lattice block = emission chunk = flag group (frames)
constexpr int kFusedQuarter = 4;      // frames per gather item (one warp, all loads in flight)
constexpr int kFusedRegs = 6;         // label columns per lane: L + 1 <= 192
constexpr int kFusedMaxNW = 6;        // compute warps per direction (R = 2: S <= 384)
constexpr int kFusedMaxThreads = 2 * (kFusedMaxNW + 1) * 32;
constexpr unsigned long long kFusedTimeoutNs = 4ull * 1000 * 1000 * 1000;

__device__ __forceinline__ int ld_acquire_gpu_i32(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_gpu_add(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() {
  asm volatile("fence.proxy.async;" ::: "memory");
}

struct FusedArgs {
  PrepArgs P;
  float* em;
  float* cmax;
  float* alpha;
  float* beta;
  int* offA;
  int* offB;
  int* flags;    // [B][NG] quarter groups finished (zeroed before the launch)
  int* ticket;   // role ticket (zeroed before the launch)
  int* err;      // sticky error word of the workspace
  int B, Sp, NW, NB, NG, G, zero_infinity, nring;
  int lat_policy;  // L2 policy of the alpha / beta stores
  int mode;  // tuning builds: 1 = gather role only, 2 = lattice role only (emissions of an earlier call)
};

// floats of one direction's shared-memory block (rings stay 128-byte aligned)
__host__ __device__ __forceinline__ size_t fused_per_dir_floats(int nring, int chunk_elems, int NW) {
  const size_t n = (size_t)nring * chunk_elems + (size_t)NW * (kValRing + kFRing + 2);
  return (n + 31) / 32 * 32;
}

// quarters (gather items) of frame group g of an utterance with Tb frames
__device__ __forceinline__ int fused_group_quarters(int Tb, int g) {
  int n = Tb - g * kFusedK;
  n = n > kFusedK ? kFusedK : n;
  return (n + kFusedQuarter - 1) / kFusedQuarter;
}

// ------------------------------------------------------------------ gather role
template <typename T>
__device__ __forceinline__ void fused_gather_role(const T* __restrict__ logits, const FusedArgs& A,
                                                  int cta, long long* sbase) {
  const PrepArgs& P = A.P;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, nwarp = blockDim.x >> 5;
  const int B = A.B, V = P.V, Lp = P.Lp;
  // target offsets of all utterances (exclusive prefix sum of the clamped lengths)
  if (wid == 0) {
    long long carry = 0;
    for (int b0 = 0; b0 < B; b0 += 32) {
      const int b = b0 + lane;
      long long len = 0;
      if (b < B) {
        if (P.offs) len = P.offs[b];
        else { len = P.tgt_lens[b]; len = len < 0 ? 0 : len; }
      }
      if (P.offs) {
        if (b < B) sbase[b] = len;
      } else {
        long long inc = len;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
          const long long v = __shfl_up_sync(0xffffffffu, inc, o);
          if (lane >= o) inc += v;
        }
        if (b < B) sbase[b] = carry + inc - len;
        carry += __shfl_sync(0xffffffffu, inc, 31);
      }
    }
  }
  __syncthreads();
  const int NG = A.NG;
  const int chalf = (NG + 1) >> 1;
  constexpr int kQ = kFusedK / kFusedQuarter;
  const long long nitems = (long long)chalf * B * 2 * kQ;
  const long long stride = (long long)A.G * nwarp;
  // consecutive items go to different CTAs: the early (outermost) groups are spread over all
  // gather SMs
  for (long long it = (long long)wid * A.G + cta; it < nitems; it += stride) {
    const int q = (int)(it % kQ);
    long long r = it / kQ;
    const int side = (int)(r & 1);
    r >>= 1;
    const int b = (int)(r % B);
    const int c = (int)(r / B);
    const long long T64 = P.in_lens[b];
    const int Tb = (int)(T64 < 0 ? 0 : (T64 > P.T ? P.T : T64));
    const int ngb = (Tb + kFusedK - 1) / kFusedK;
    // outside-in: side 0 walks the groups up from 0, side 1 down from the last one
    if (side == 0 ? 2 * c > ngb - 1 : 2 * c >= ngb - 1) continue;
    const int g = side ? ngb - 1 - c : c;
    const int t0 = g * kFusedK + q * kFusedQuarter;
    if (t0 >= Tb) continue;
    const int nr = Tb - t0 < kFusedQuarter ? Tb - t0 : kFusedQuarter;
    const long long L64 = P.tgt_lens[b];
    const int L = (int)(L64 < 0 ? 0 : (L64 > P.Lmax ? P.Lmax : L64));
    const long long base = sbase[b];
    int col[kFusedRegs];
#pragma unroll
    for (int j = 0; j < kFusedRegs; ++j) {
      const int k = lane + 32 * j;
      long long t = P.blank;
      if (k >= 1 && k <= L) {
        t = P.targets[base + k - 1];
        t = t < 0 ? 0 : (t >= V ? V - 1 : t);
      }
      col[j] = (int)t;
    }
    const size_t g0 = (size_t)b * P.T + t0;
    float v[kFusedQuarter][kFusedRegs];
#pragma unroll
    for (int j = 0; j < kFusedRegs; ++j) {
      const int k = lane + 32 * j;
#pragma unroll
      for (int rr = 0; rr < kFusedQuarter; ++rr) {  // clamped row: always a valid address
        const T* row = logits + (g0 + (rr < nr ? rr : nr - 1)) * V;
        v[rr][j] = (k <= L) ? ldg_gather<T>(row + col[j]) * kLog2e : kNegInf;
      }
    }
#pragma unroll
    for (int rr = 0; rr < kFusedQuarter; ++rr) {
      float m = kNegInf;
#pragma unroll
      for (int j = 0; j < kFusedRegs; ++j) m = fmaxf(m, v[rr][j]);
      m = warp_max(m);
      if (!(m > kNegInf && m < -kNegInf)) m = 0.f;  // all -inf / NaN rows: no centring
      if (rr < nr) {
        float* er = A.em + (g0 + rr) * Lp;
#pragma unroll
        for (int j = 0; j < kFusedRegs; ++j) {
          const int k = lane + 32 * j;
          if (k <= L) er[k == 0 ? 0 : kEmLabel0 + k - 1] = v[rr][j] - m;
        }
        if (lane == 0) A.cmax[g0 + rr] = m;
      }
    }
    __syncwarp();
    if (lane == 0) {
      __threadfence();
      red_release_gpu_add(A.flags + (size_t)b * NG + g, 1);
    }
  }
}

// ------------------------------------------------------------------ the kernel
template <typename T>
__global__ void __launch_bounds__(kFusedMaxThreads, 1) fused_fwd_kernel(const T* __restrict__ logits,
                                                                        FusedArgs A) {
  constexpr int K = kFusedK, R = 2;
  extern __shared__ __align__(128) float lsm[];
  __shared__ __align__(8) uint64_t bars[2][2 * kMaxEmRing];  // per direction: full[], empty[]
  __shared__ int s_role;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) s_role = atomicAdd(A.ticket, 1);
  __syncthreads();
  const int role = s_role;
  if (role < A.G) {
    if (A.mode != 2) fused_gather_role<T>(logits, A, role, reinterpret_cast<long long*>(lsm));
    return;
  }
  const int b = role - A.G;
  if (b >= A.B || A.mode == 1) return;
  const int NW = A.NW, Lp = A.P.Lp;
  // ---- label prep of this utterance (lab / dup / utt for the later kernels)
  int Tb, L;
  prep_utterance(A.P, b, reinterpret_cast<int*>(lsm), true, Tb, L);
  __syncthreads();  // the staging area becomes the emission rings
  UttInfo* utt = A.P.utt;
  const int S = 2 * L + 1;
  if (Tb == 0) {
    if (tid == 0) {
      const bool inf = L > 0;
      utt[b].lam = inf ? (double)kNegInf : 0.0;
      utt[b].aend = 0;
      utt[b].infinite = inf;
      utt[b].zeroed = inf && A.zero_infinity;
    }
    return;
  }
  const int nw_live = (S + 32 * R - 1) / (32 * R);
  const int dir = wid / (NW + 1), wl = wid - dir * (NW + 1);
  if (dir > 1) return;
  const int chunk_elems = K * Lp;
  const size_t per_dir = fused_per_dir_floats(A.nring, chunk_elems, NW);
  float* ring0 = lsm;
  float* ring = ring0 + dir * per_dir;
  float* vals = ring + (size_t)A.nring * chunk_elems;                  // [NW][kValRing]
  int* fring = reinterpret_cast<int*>(vals + (size_t)NW * kValRing);   // [NW][kFRing]
  int* ready = fring + (size_t)NW * kFRing;                            // [NW]
  int* cons = ready + NW;                                              // [NW]
  uint64_t* const full = bars[dir];
  uint64_t* const empty = bars[dir] + kMaxEmRing;
  if (tid == 0) {
    for (int d = 0; d < 2; ++d)
      for (int i = 0; i < A.nring; ++i) {
        mbar_init(&bars[d][i], 1);
        mbar_init(&bars[d][kMaxEmRing + i], nw_live);
      }
    fence_mbar_init();
  }
  for (int d = 0; d < 2; ++d) {
    float* vd = ring0 + d * per_dir + (size_t)A.nring * chunk_elems;
    for (int i = tid; i < NW * kValRing; i += blockDim.x) vd[i] = kLogZero;
    int* fd = reinterpret_cast<int*>(vd + (size_t)NW * kValRing);
    for (int i = tid; i < NW * (kFRing + 2); i += blockDim.x) fd[i] = 0;  // fring, ready, cons
  }
  __syncthreads();

  // ---- TMA producer warp of this direction: chunk c = recursion steps [c*K, c*K+n)
  if (wl == NW) {
    if (lane == 0) {
      const float* emb = A.em + (size_t)b * A.P.T * Lp;
      const int* fl = A.flags + (size_t)b * A.NG;
      const int nchunks = (Tb + K - 1) / K;
      for (int c = 0; c < nchunks; ++c) {
        const int slot = c % A.nring, round = c / A.nring;
        if (round > 0) mbar_wait(&empty[slot], (round - 1) & 1);
        const int i0 = c * K;
        const int n = (Tb - i0 < K) ? Tb - i0 : K;
        const int f0 = dir ? Tb - i0 - n : i0;  // first frame of the (contiguous) block
        // the gather CTAs must have finished every frame group the chunk overlaps
        const int g_lo = f0 / K, g_hi = (f0 + n - 1) / K;
        for (int g = g_lo; g <= g_hi; ++g) {
          const int need = fused_group_quarters(Tb, g);
          unsigned ns = 20;
          unsigned long long t0 = 0;
          while (ld_acquire_gpu_i32(fl + g) < need) {
            __nanosleep(ns);
            if (ns < 200) {
              ns += 20;
            } else {  // a gather CTA that never shows up: sticky error + NaN loss, not a hang
              const unsigned long long now = globaltimer_ns();
              if (t0 == 0) t0 = now;
              if (now - t0 > kFusedTimeoutNs) { atomicExch(A.err, kWsErrGather); break; }
            }
          }
        }
        fence_proxy_async_all();  // generic-proxy writes (other SMs) -> the async-proxy read below
        const uint32_t bytes = (uint32_t)n * Lp * 4u;
        mbar_expect_tx(&full[slot], bytes);
        bulk_g2s(ring + (size_t)slot * chunk_elems, emb + (size_t)f0 * Lp, bytes, &full[slot]);
      }
    }
    return;
  }
  if (wl >= nw_live) return;

  LatticeArgs LA;
  LA.utt = utt;
  LA.lb = A.P.lab + (size_t)b * Lp;
  LA.lat = (dir ? A.beta : A.alpha) + (size_t)b * A.P.T * A.Sp;
  LA.off = (dir ? A.offB : A.offA) + (size_t)b * A.NB * NW;
  LA.ring = smem_u32(ring);
  LA.vals = smem_u32(vals);
  LA.fring = smem_u32(fring);
  LA.ready = smem_u32(ready);
  LA.cons = smem_u32(cons);
  LA.full = smem_u32(full);
  LA.Tb = Tb; LA.L = L; LA.Lp = Lp; LA.Sp = A.Sp; LA.NWp = NW;
  LA.nw_live = nw_live; LA.zero_infinity = A.zero_infinity; LA.b = b; LA.nring = A.nring;
  LA.W0 = 0; LA.Lps = Lp; LA.col0 = 0; LA.nwc = NW;
  LA.gv_in = LA.gv_out = nullptr; LA.gf_in = LA.gf_out = nullptr; LA.gr_in = LA.gr_out = nullptr;
  LA.err = A.err;
  LA.lat_policy = A.lat_policy;
  if (dir == 0) lattice_warp<R, true, 0, K, false>(LA, wl, lane);
  else lattice_warp<R, true, 1, K, false>(LA, wl, lane);
}

// Is the fused forward applicable?  One CTA holds both directions of an utterance: R = 2
// lattices of at most kFusedMaxNW warps whose two emission rings fit one SM.
static inline bool fused_fwd_fits(const LossLayout& w, int* nring_out, size_t* smem_out) {
  if (w.CL != 1 || w.R != 2 || w.NW > kFusedMaxNW || w.Lmax + 1 > 32 * kFusedRegs) return false;
  const int nring = w.NW + 2;  // warp w+1 trails warp w by one block (+ one chunk of slack)
  if (nring > kMaxEmRing) return false;
  size_t smem = 2 * 4 * fused_per_dir_floats(nring, kFusedK * w.Lp, w.NW);
  if (smem < (size_t)w.B * 8 + 64) smem = (size_t)w.B * 8 + 64;  // gather role: target offsets
  if (smem < (size_t)w.Lp * 4) smem = (size_t)w.Lp * 4;          // label staging of the prep
  if (smem > kLatticeSmemBudget) return false;
  // (at least half of the shared memory: one CTA per SM, so that roles never share an SM)
  if (smem < 116 * 1024) smem = 116 * 1024;
  *nring_out = nring;
  *smem_out = smem;
  return true;
}

template <typename T>
static int launch_fused_fwd(const void* logits, const PrepArgs& P, const LossLayout& w,
                            unsigned char* ws, int zero_infinity, int nring, size_t smem,
                            cudaStream_t st) {
  FusedArgs A;
  A.P = P;
  A.em = reinterpret_cast<float*>(ws + w.em);
  A.cmax = reinterpret_cast<float*>(ws + w.cmax);
  A.alpha = reinterpret_cast<float*>(ws + w.alpha);
  A.beta = reinterpret_cast<float*>(ws + w.beta);
  A.offA = reinterpret_cast<int*>(ws + w.offA);
  A.offB = reinterpret_cast<int*>(ws + w.offB);
  A.flags = reinterpret_cast<int*>(ws + w.flags);
  A.ticket = reinterpret_cast<int*>(ws + w.ticket);
  A.err = reinterpret_cast<int*>(ws + w.err);
  A.B = w.B; A.Sp = w.Sp; A.NW = w.NW; A.NB = w.NB; A.NG = w.NG;
  A.zero_infinity = zero_infinity; A.nring = nring;
  const int sms = sm_count();
  int G = sms - w.B;
  if (G < sms / 4) G = sms / 4;
  A.G = tuning_int("B200CTC_FUSED_G", G);
  A.mode = tuning_int("B200CTC_FUSED_MODE", 0);
  A.lat_policy = lat_policy_host();
  // ticket + flags are adjacent sections: one memset
  if (A.mode == 2) cudaMemsetAsync(ws + w.flags, 0x01, 4 * (size_t)w.B * w.NG, st);
  if (cudaMemsetAsync(ws + w.ticket, 0, (w.flags - w.ticket) + (A.mode == 2 ? 0 : 4 * (size_t)w.B * w.NG), st) != cudaSuccess) {
    set_error("ctc fused forward: memset failed");
    return B200CTC_ERR_DEVICE;
  }
  auto k = fused_fwd_kernel<T>;
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  const int nt = 2 * (w.NW + 1) * 32;
  k<<<A.G + w.B, nt, smem, st>>>(static_cast<const T*>(logits), A);
  count_launch();
  return check_launch("ctc fused forward");
}

}  // namespace b200ctc
