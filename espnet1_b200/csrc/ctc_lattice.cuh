// Alpha/beta lattice recursion (Graves 2006 eq. 6-11) in normalised log2 space.
// One CTA per (utterance, direction): blockIdx.y = 0 runs alpha forward in
// time, 1 runs beta backward in time as "alpha on the reversed lattice".
// The lattice row lives in shared memory (double buffered, one barrier per
// frame).  The emissions em[t][k] of the next frames are streamed into a
// shared-memory ring by the TMA engine (1-D bulk copies of whole frame chunks,
// completion on mbarriers) so the dependent chain never touches global memory;
// every K frames the row is re-centred on its maximum (shift recorded in fp64).
#pragma once
#include "ctc_ws.cuh"

namespace b200ctc {

constexpr int kMaxLattice = 8192;  // 2*L+1 <= 8192  (R <= 8 positions per thread)
constexpr int kRenormEvery = 16;
constexpr int kEmRing = 4;         // chunks of emission frames in flight
constexpr int kHalo = 4;           // -inf cells in front of the row (keeps rows 16-B aligned)

__device__ __forceinline__ float lse3_log2(float a0, float a1, float a2) {
  const float m = fmaxf(fmaxf(a0, a1), a2);
  const float ms = (m == kNegInf) ? 0.f : m;
  return ms + lg2(ex2(a0 - ms) + ex2(a1 - ms) + ex2(a2 - ms));
}

template <int R>
__global__ void __launch_bounds__(1024) lattice_kernel(
    UttInfo* __restrict__ utt, const int* __restrict__ lab, int Lp, const float* __restrict__ em,
    const float* __restrict__ mt, float* __restrict__ alpha, float* __restrict__ beta,
    double* __restrict__ offA, double* __restrict__ offB, int Tfr, int Sp, int zero_infinity,
    float* __restrict__ nll, int chunk) {
  extern __shared__ __align__(128) float lsm[];
  __shared__ __align__(8) uint64_t bars[kEmRing];
  __shared__ double dred[32];
  __shared__ float wmax[32];
  const int NT = blockDim.x;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5, NW = NT >> 5;
  const int b = blockIdx.x, dir = blockIdx.y;
  const int Tb = utt[b].T, L = utt[b].L, S = 2 * L + 1;
  if (Tb == 0) {
    if (dir == 0 && tid == 0) {
      const bool inf = L > 0;
      utt[b].infinite = inf;
      utt[b].zeroed = inf && zero_infinity;
      nll[b] = inf ? (zero_infinity ? 0.f : -kNegInf) : 0.f;
    }
    return;
  }
  const int rowlen = Sp + kHalo;
  float* buf0 = lsm;
  float* buf1 = lsm + rowlen;
  float* ring = lsm + 2 * rowlen;  // [kEmRing][chunk * Lp]
  const int chunk_elems = chunk * Lp;
  float* lat = (dir ? beta : alpha) + (size_t)b * Tfr * Sp;
  double* off = (dir ? offB : offA) + (size_t)b * Tfr;
  const float* emb = em + (size_t)b * Tfr * Lp;
  const int* lb = lab + (size_t)b * Lp;
  const int nchunks = (Tb + chunk - 1) / chunk;

  // chunk c holds recursion steps [c*chunk, (c+1)*chunk): frames ascending for
  // alpha, the mirrored (still contiguous) frame block for beta
  auto issue_chunk = [&](int c) {
    const int i0 = c * chunk;
    const int n = (Tb - i0 < chunk) ? Tb - i0 : chunk;
    const int f0 = dir ? Tb - i0 - n : i0;  // first frame of the block
    const int slot = c % kEmRing;
    const uint32_t bytes = (uint32_t)n * Lp * 4u;
    mbar_expect_tx(&bars[slot], bytes);
    bulk_g2s(ring + (size_t)slot * chunk_elems, emb + (size_t)f0 * Lp, bytes, &bars[slot]);
  };
  if (tid == 0) {
    for (int i = 0; i < kEmRing; ++i) mbar_init(&bars[i], 1);
    fence_mbar_init();
  }
  for (int i = tid; i < 2 * rowlen; i += NT) lsm[i] = kNegInf;
  __syncthreads();
  if (tid == 0)
    for (int c = 0; c < kEmRing && c < nchunks; ++c) issue_chunk(c);

  // static description of this thread's lattice positions
  int s_of[R], k_of[R];
  bool act[R], skip[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int sp = tid + j * NT;          // position in recursion order
    act[j] = sp < S;
    const int s = dir ? S - 1 - sp : sp;  // true lattice index
    s_of[j] = s;
    const bool odd = s & 1;
    const int k = odd ? (s + 1) >> 1 : 0;
    k_of[j] = k;
    bool sk = false;
    if (act[j] && odd) {
      if (dir == 0) sk = (s >= 3) && (lb[k] != lb[k - 1]);
      else sk = (s + 2 <= S - 1) && (lb[k] != lb[k + 1]);
    }
    skip[j] = sk;
  }

  double a_run = 0.0;
  for (int c = 0; c < nchunks; ++c) {
    const int i0 = c * chunk;
    const int n = (Tb - i0 < chunk) ? Tb - i0 : chunk;
    const int slot = c % kEmRing;
    const float* ech = ring + (size_t)slot * chunk_elems;
    mbar_wait(&bars[slot], (c / kEmRing) & 1);
    for (int u = 0; u < n; ++u) {
      const int i = i0 + u;
      const int t = dir ? Tb - 1 - i : i;
      const float* erow = ech + (size_t)(dir ? n - 1 - u : u) * Lp;
      const float* cur = (i & 1) ? buf1 : buf0;
      float* nxt = (i & 1) ? buf0 : buf1;
      float cshift = 0.f;
      if (i > 0 && (i % kRenormEvery) == 0) {  // apply the shift measured last frame
        cshift = warp_max(lane < NW ? wmax[lane] : kNegInf);
        if (!(cshift > kNegInf && cshift < -kNegInf)) cshift = 0.f;
        a_run += (double)cshift;
      }
      float vmax = kNegInf;
#pragma unroll
      for (int j = 0; j < R; ++j) {
        if (act[j]) {
          const int sp = tid + j * NT;
          const float e = erow[k_of[j]];
          float v;
          if (i == 0) {
            v = sp < 2 ? e : kNegInf;
          } else {
            const float a0 = cur[sp + kHalo], a1 = cur[sp + kHalo - 1];
            const float a2 = skip[j] ? cur[sp + kHalo - 2] : kNegInf;
            v = lse3_log2(a0, a1, a2) + (e - cshift);
          }
          nxt[sp + kHalo] = v;
          lat[(size_t)t * Sp + s_of[j]] = v;
          vmax = fmaxf(vmax, v);
        }
      }
      if ((i % kRenormEvery) == kRenormEvery - 1) {  // measure for the next frame
        vmax = warp_max(vmax);
        if (lane == 0) wmax[wid] = vmax;
      }
      if (tid == 0) off[t] = a_run;
      __syncthreads();
    }
    // every thread is past its last read of this ring slot: refill it
    if (tid == 0 && c + kEmRing < nchunks) issue_chunk(c + kEmRing);
  }
  if (dir != 0) return;
  // ---- alpha epilogue: log-likelihood ------------------------------------
  const float* mtb = mt + (size_t)b * Tfr;
  double part = 0.0;
  for (int t = tid; t < Tb; t += NT) part += (double)mtb[t];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (lane == 0) dred[wid] = part;
  __syncthreads();
  if (tid == 0) {
    double msum = 0.0;
    for (int w2 = 0; w2 < NW; ++w2) msum += dred[w2];
    const float* fin = ((Tb - 1) & 1) ? buf0 : buf1;  // buffer written by the last frame
    const double l1 = (double)fin[(S - 1) + kHalo];
    const double l2 = S > 1 ? (double)fin[(S - 2) + kHalo] : (double)kNegInf;
    const double m = fmax(l1, l2);
    double lam;
    if (m == (double)kNegInf) lam = m;
    else lam = m + log2(exp2(l1 - m) + exp2(l2 - m));
    const bool inf = (lam == (double)kNegInf);
    utt[b].lam = lam;
    utt[b].aend = a_run;
    utt[b].msum = msum;
    utt[b].infinite = inf;
    utt[b].zeroed = inf && zero_infinity;
    const double nll2 = -(lam + msum + a_run);
    nll[b] = inf ? (zero_infinity ? 0.f : -kNegInf) : (float)(nll2 * 0.6931471805599453);
  }
}

// qt[b,t] = mt + offA + offB - aend - lam : the per-frame scalar of the occupancy
// exponent (ctc_ws.cuh), combined in fp64 once so the gradient pass reads one float
__global__ void __launch_bounds__(256) finalize_kernel(const UttInfo* __restrict__ utt,
                                                       const float* __restrict__ mt,
                                                       const double* __restrict__ offA,
                                                       const double* __restrict__ offB,
                                                       float* __restrict__ qt, int Tfr) {
  const int b = blockIdx.y;
  const int t = blockIdx.x * 256 + threadIdx.x;
  if (t >= utt[b].T) return;
  const size_t g = (size_t)b * Tfr + t;
  qt[g] = (float)((double)mt[g] + offA[g] + offB[g] - utt[b].aend - utt[b].lam);
}

template <int R>
static int launch_lattice_r(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                            int need_grad, float* nll, cudaStream_t st) {
  const int Smax = 2 * w.Lmax + 1;
  int nt = (Smax + R - 1) / R;
  nt = (nt + 31) / 32 * 32;
  if (nt < 32) nt = 32;
  int chunk = 8192 / (w.Lp * 4);
  if (chunk < 1) chunk = 1;
  if (chunk > 32) chunk = 32;
  const size_t smem =
      sizeof(float) * (2 * (size_t)(w.Sp + kHalo) + (size_t)kEmRing * chunk * w.Lp) + 128;
  auto k = lattice_kernel<R>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  dim3 grid(w.B, need_grad ? 2 : 1);
  k<<<grid, nt, smem, st>>>(reinterpret_cast<UttInfo*>(ws + w.utt),
                            reinterpret_cast<const int*>(ws + w.lab), w.Lp,
                            reinterpret_cast<const float*>(ws + w.em),
                            reinterpret_cast<const float*>(ws + w.mt),
                            reinterpret_cast<float*>(ws + w.alpha),
                            reinterpret_cast<float*>(ws + w.beta),
                            reinterpret_cast<double*>(ws + w.offA),
                            reinterpret_cast<double*>(ws + w.offB), Tfr, w.Sp, zero_infinity, nll,
                            chunk);
  count_launch();
  if (need_grad && Tfr > 0) {
    finalize_kernel<<<dim3((Tfr + 255) / 256, w.B), 256, 0, st>>>(
        reinterpret_cast<const UttInfo*>(ws + w.utt), reinterpret_cast<const float*>(ws + w.mt),
        reinterpret_cast<const double*>(ws + w.offA), reinterpret_cast<const double*>(ws + w.offB),
        reinterpret_cast<float*>(ws + w.qt), Tfr);
    count_launch();
  }
  return check_launch("ctc lattice");
}

static inline int launch_lattice(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                                 int need_grad, float* nll, cudaStream_t st) {
  const int Smax = 2 * w.Lmax + 1;
  const int r = (Smax + 1023) / 1024;
  if (r <= 1) return launch_lattice_r<1>(w, ws, Tfr, zero_infinity, need_grad, nll, st);
  if (r <= 2) return launch_lattice_r<2>(w, ws, Tfr, zero_infinity, need_grad, nll, st);
  if (r <= 4) return launch_lattice_r<4>(w, ws, Tfr, zero_infinity, need_grad, nll, st);
  return launch_lattice_r<8>(w, ws, Tfr, zero_infinity, need_grad, nll, st);
}

}  // namespace b200ctc
