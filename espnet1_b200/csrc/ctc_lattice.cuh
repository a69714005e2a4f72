// Alpha/beta lattice recursion (Graves 2006 eq. 6-11) in normalised log2 space.
// One CTA per (utterance, direction): blockIdx.y = 0 runs alpha forward in
// time, 1 runs beta backward in time as "alpha on the reversed lattice".
// The lattice row lives in shared memory (double buffered, one barrier per
// frame); emissions are prefetched PF frames ahead into registers; every K
// frames the row is re-centred on its maximum (shift recorded in fp64).
#pragma once
#include "ctc_ws.cuh"

namespace b200ctc {

constexpr int kMaxLattice = 8192;  // 2*L+1 <= 8192  (R <= 8 positions per thread)
constexpr int kRenormEvery = 16;

__device__ __forceinline__ float lse3_log2(float a0, float a1, float a2) {
  const float m = fmaxf(fmaxf(a0, a1), a2);
  const float ms = (m == kNegInf) ? 0.f : m;
  return ms + lg2(ex2(a0 - ms) + ex2(a1 - ms) + ex2(a2 - ms));
}

template <int R, int PF>
__global__ void __launch_bounds__(1024) lattice_kernel(
    UttInfo* __restrict__ utt, const int* __restrict__ lab, int Lp, const float* __restrict__ em,
    const float* __restrict__ mt, float* __restrict__ alpha, float* __restrict__ beta,
    double* __restrict__ offA, double* __restrict__ offB, int Tfr, int Sp, int zero_infinity,
    float* __restrict__ nll) {
  extern __shared__ float lsm[];
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
  float* buf0 = lsm;                // [Sp + 4], cells 0,1 are the -inf halo
  float* buf1 = lsm + Sp + 4;
  float* wmax = lsm + 2 * (Sp + 4);  // [32]
  __shared__ double dred[32];
  float* lat = (dir ? beta : alpha) + (size_t)b * Tfr * Sp;
  double* off = (dir ? offB : offA) + (size_t)b * Tfr;
  const float* emb = em + (size_t)b * Tfr * Lp;
  const int* lb = lab + (size_t)b * Lp;

  for (int i = tid; i < 2 * (Sp + 4); i += NT) lsm[i] = kNegInf;

  // static description of this thread's lattice positions
  int s_of[R], k_of[R];
  bool act[R], skip[R];
#pragma unroll
  for (int j = 0; j < R; ++j) {
    const int sp = tid + j * NT;       // position in recursion order
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
  // emission prefetch ring
  float pf[R][PF];
#pragma unroll
  for (int u = 0; u < PF; ++u) {
    const int t = dir ? Tb - 1 - u : u;
#pragma unroll
    for (int j = 0; j < R; ++j)
      pf[j][u] = (u < Tb && act[j]) ? emb[(size_t)t * Lp + k_of[j]] : 0.f;
  }
  __syncthreads();

  double a_run = 0.0;
  for (int i0 = 0; i0 < Tb; i0 += PF) {
#pragma unroll
    for (int u = 0; u < PF; ++u) {
      const int i = i0 + u;
      if (i < Tb) {
        const int t = dir ? Tb - 1 - i : i;
        const float* cur = (i & 1) ? buf1 : buf0;
        float* nxt = (i & 1) ? buf0 : buf1;
        float c = 0.f;
        if (i > 0 && (i % kRenormEvery) == 0) {  // apply the shift measured last frame
          c = warp_max(lane < NW ? wmax[lane] : kNegInf);
          if (!(c > kNegInf && c < -kNegInf)) c = 0.f;
          a_run += (double)c;
        }
        float vmax = kNegInf;
#pragma unroll
        for (int j = 0; j < R; ++j) {
          const float e = pf[j][u];
          const int ip = i + PF;
          if (ip < Tb && act[j]) {
            const int tp = dir ? Tb - 1 - ip : ip;
            pf[j][u] = emb[(size_t)tp * Lp + k_of[j]];
          }
          if (act[j]) {
            const int sp = tid + j * NT;
            float v;
            if (i == 0) {
              v = sp < 2 ? e : kNegInf;
            } else {
              const float a0 = cur[sp + 2], a1 = cur[sp + 1];
              const float a2 = skip[j] ? cur[sp] : kNegInf;
              v = lse3_log2(a0, a1, a2) + e - c;
            }
            nxt[sp + 2] = v;
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
    }
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
    const double l1 = (double)fin[(S - 1) + 2];
    const double l2 = S > 1 ? (double)fin[(S - 2) + 2] : (double)kNegInf;
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

template <int R>
static int launch_lattice_r(const LossLayout& w, unsigned char* ws, int Tfr, int zero_infinity,
                            int need_grad, float* nll, cudaStream_t st) {
  constexpr int PF = R <= 2 ? 8 : (R <= 4 ? 4 : 2);
  const int Smax = 2 * w.Lmax + 1;
  int nt = (Smax + R - 1) / R;
  nt = (nt + 31) / 32 * 32;
  if (nt < 32) nt = 32;
  const size_t smem = sizeof(float) * (2 * (size_t)(w.Sp + 4) + 32);
  auto k = lattice_kernel<R, PF>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  dim3 grid(w.B, need_grad ? 2 : 1);
  k<<<grid, nt, smem, st>>>(reinterpret_cast<UttInfo*>(ws + w.utt),
                            reinterpret_cast<const int*>(ws + w.lab), w.Lp,
                            reinterpret_cast<const float*>(ws + w.em),
                            reinterpret_cast<const float*>(ws + w.mt),
                            reinterpret_cast<float*>(ws + w.alpha),
                            reinterpret_cast<float*>(ws + w.beta),
                            reinterpret_cast<double*>(ws + w.offA),
                            reinterpret_cast<double*>(ws + w.offB), Tfr, w.Sp, zero_infinity, nll);
  count_launch();
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
