// Eager-gradient mode, split so that the lattice's serial latency is hidden:
//
//   stream A (caller's):  prep -> emission gather -> lattice ............ -> fix-up -> nll
//   stream B (internal):  dense rows (read logits once, write g*softmax) ....^
//
// dense_rows_kernel does not depend on the lattice at all: it is the one full
// read of the (B,T,V) logits and the one full write of the gradient
// (g * softmax, plus the row's lse2), HBM-bound, and runs CONCURRENTLY with the
// latency-bound gather + lattice chain (whose CTAs fit beside it on every SM).
// fixup_kernel then corrects only the <= L+1 label columns of every row:
//     grad[b,t,v] = g * ( softmax(x)[v] - sum_{s: l_s = v} gamma_t(s) )
// using the emissions kept by the gather (no second look at the logits).
#include "rows_common.cuh"

namespace b200ctc {

// ------------------------------------------------------------ dense rows
template <typename T, int NT, int PER, bool TMA>
__global__ void __launch_bounds__(NT) dense_rows_kernel(
    const T* __restrict__ logits, T* __restrict__ grad, int B, int Tfr, int V,
    const int64_t* __restrict__ in_lens, float* __restrict__ lse2, float gmul, int nstage,
    int stage_stride) {
  constexpr int N = Elems<T>::N;
  constexpr int NWARP = NT / 32;
  extern __shared__ __align__(128) unsigned char dsm[];
  __shared__ __align__(8) uint64_t bars[kMaxRowStages];
  __shared__ float2 wms[2][32];  // per-warp (max, sum) of the row, double buffered
  unsigned char* stages = dsm;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int64_t nrows = (int64_t)B * Tfr;
  const uint32_t row_bytes = (uint32_t)V * sizeof(T);
  const int nchunk = V / N;
  const int ntail = V - nchunk * N;
  auto len_of = [&](int b) {
    const long long l = in_lens[b];
    return (int)(l < 0 ? 0 : (l > Tfr ? Tfr : l));
  };

  RowCursor pc;
  pc.init(blockIdx.x, gridDim.x, Tfr);
  uint32_t np = 0;
  auto issue_next = [&]() {
    while (pc.g < nrows && pc.t >= len_of(pc.b)) pc.next();
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
    const int64_t g = rc.g;
    T* gr = grad + g * V;
    const bool al16 = (reinterpret_cast<uintptr_t>(gr) & 15) == 0;
    if (rc.t >= len_of(rc.b)) {  // padded frame: gradient is 0
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
    const unsigned char* stg;
    if (TMA) {
      const int st = n % nstage;
      stg = stages + (size_t)st * stage_stride;
      mbar_wait(&bars[st], (n / nstage) & 1);
    } else {
      stg = stages;
      const T* src = logits + g * V;
      T* dst = reinterpret_cast<T*>(stages);
      __syncthreads();  // previous row's readers are done with the stage
      for (int i = tid; i < V; i += NT) dst[i] = src[i];
      __syncthreads();
    }
    const T* xs = reinterpret_cast<const T*>(stg);
    // whole row -> registers; warp-level max, then 2^(x2 - warp max) kept for the output
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
    if (tid < ntail) {
      xt = ex2(fmaf(xt, kLog2e, -mw2));
      sum += xt;
    }
    sum = warp_sum(sum);
    if (lane == 0) wms[n & 1][wid] = make_float2(mw2, sum);
    __syncthreads();  // every thread holds its part of the row in registers: the stage is free
    if (TMA && tid == 0) issue_next();
    float M2 = kNegInf;
#pragma unroll
    for (int w2 = 0; w2 < NWARP; ++w2) M2 = fmaxf(M2, wms[n & 1][w2].x);
    float tot = 0.f;
#pragma unroll
    for (int w2 = 0; w2 < NWARP; ++w2) tot += wms[n & 1][w2].y * ex2(wms[n & 1][w2].x - M2);
    const float l2 = M2 + log2f(tot);
    if (tid == 0) lse2[g] = l2;
    const float scale = gmul * ex2(mw2 - l2);  // g * 2^(warp max - lse2)
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
    ++n;
  }
}

// ------------------------------------------------------------ label-column fix-up
// One warp per (b,t) row.  Occupancies gamma_t(s) = 2^(alpha + beta - em + frame
// offsets - lam) (ctc_ws.cuh) are combined per vocabulary column through the
// duplicate chains of prep_kernel (deterministic, no atomics); the soft-max of the
// touched columns is rebuilt from the gathered emissions: 2^(em + c - lse2).
constexpr int kFixWarps = 4;

template <typename T>
__global__ void __launch_bounds__(kFixWarps * 32) fixup_kernel(
    T* __restrict__ grad, int B, int Tfr, int V, int blank, const UttInfo* __restrict__ utt,
    const int* __restrict__ lab, const int* __restrict__ dup, int Lp, int Sp, int NWp, int NB, int wshift,
    const float* __restrict__ lse2, const float* __restrict__ cmax, const float* __restrict__ em,
    const float* __restrict__ alpha, const float* __restrict__ beta,
    const int* __restrict__ offA, const int* __restrict__ offB, float gmul) {
  extern __shared__ float pbuf_all[];  // [kFixWarps][Lp]
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  float* pbuf = pbuf_all + (size_t)wid * Lp;
  const int64_t nrows = (int64_t)B * Tfr;
  const int64_t stride = (int64_t)gridDim.x * kFixWarps;
  for (int64_t g = (int64_t)blockIdx.x * kFixWarps + wid; g < nrows; g += stride) {
    const int b = (int)(g / Tfr), t = (int)(g - (int64_t)b * Tfr);
    const UttInfo u = utt[b];
    if (t >= u.T) continue;  // padded frame: dense_rows_kernel wrote zeros
    T* gr = grad + g * V;
    if (u.zeroed || u.infinite) {  // whole-utterance 0 (zero_infinity) or NaN (inf loss kept)
      const T fill = from_f32<T>(u.zeroed ? 0.f : __int_as_float(0x7fc00000));
      for (int i = lane; i < V; i += 32) gr[i] = fill;
      continue;
    }
    const int L = u.L, S = 2 * L + 1;
    const float* al = alpha + (size_t)g * Sp;
    const float* be = beta + (size_t)g * Sp;
    // frame offsets: one per block of kRenormEvery recursion steps (t / T_b-1-t)
    const int* oa = offA + ((size_t)b * NB + (t / kRenormEvery)) * NWp;
    const int* ob = offB + ((size_t)b * NB + ((u.T - 1 - t) / kRenormEvery)) * NWp;
    const float* er = em + (size_t)g * Lp;
    const int* lb = lab + (size_t)b * Lp;
    const int* db = dup + (size_t)b * Lp;
    const float qbase = -(float)u.lam;
    const float cl = cmax[g] - lse2[g];  // log2-softmax of a column = em + cl
    float blank_part = 0.f;
    for (int s = lane; s < S; s += 32) {
      const int sb = S - 1 - s;  // beta is stored in recursion order
      const int fo = oa[s >> wshift] + ob[sb >> wshift] - u.aend;
      const bool odd = s & 1;
      const int k = (s + 1) >> 1;
      const float e = er[odd ? kEmLabel0 + k - 1 : 0];
      const float p = ex2(al[s] + be[sb] + ((float)fo + qbase) - e);
      if (odd) pbuf[k] = p; else blank_part += p;
    }
    blank_part = warp_sum(blank_part);
    __syncwarp();
    for (int k = lane + 1; k <= L; k += 32) {
      const int code = db[k];
      if (code & kDupFirst) {
        float acc = pbuf[k];
        for (int nx = code & kDupMask; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
        const float sm = ex2(er[kEmLabel0 + k - 1] + cl);
        gr[lb[k]] = from_f32<T>(gmul * (sm - acc));
      }
    }
    if (lane == 0) {
      float acc = blank_part;
      for (int nx = u.blank_head; nx != 0; nx = db[nx] & kDupMask) acc += pbuf[nx];
      const float sm = ex2(er[0] + cl);
      gr[blank] = from_f32<T>(gmul * (sm - acc));
    }
    __syncwarp();  // pbuf is reused by the next row
  }
}

// ------------------------------------------------------------ launchers
template <typename T, int NT>
static int launch_dense_nt(const void* logits, void* grad, int B, int Tfr, int V,
                           const int64_t* in_lens, float* lse2, float gmul, cudaStream_t st) {
  constexpr int PER = 40 / Elems<T>::N;
  // one CTA per SM fewer than the fused row pass: leaves shared memory for a lattice CTA
  RowLaunchCfg c = row_launch_cfg(logits, (size_t)V * sizeof(T), 0, NT, "B200CTC_D");
  if (!getenv("B200CTC_D_CTAS") && c.ctas_per_sm > 4) c.ctas_per_sm = 4;
  B2_REQUIRE((size_t)c.nstage * c.stride <= 226 * 1024,
             "ctc_loss: V=%d too large for the shared-memory row path", V);
  const size_t smem = (size_t)c.nstage * c.stride;
  const int64_t rows = (int64_t)B * Tfr;
  int64_t grid = (int64_t)sm_count() * c.ctas_per_sm;
  if (grid > rows) grid = rows;
  auto go = [&](auto k) {
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    k<<<(unsigned)grid, NT, smem, st>>>(static_cast<const T*>(logits), static_cast<T*>(grad), B,
                                        Tfr, V, in_lens, lse2, gmul, c.nstage, c.stride);
  };
  if (c.tma) go(dense_rows_kernel<T, NT, PER, true>);
  else go(dense_rows_kernel<T, NT, PER, false>);
  count_launch();
  return check_launch("ctc dense_rows");
}

template <typename T>
static int launch_dense_t(const void* logits, void* grad, int B, int Tfr, int V,
                          const int64_t* in_lens, float* lse2, float gmul, cudaStream_t st) {
  const int need = (V + 39) / 40;
  if (need <= 128) return launch_dense_nt<T, 128>(logits, grad, B, Tfr, V, in_lens, lse2, gmul, st);
  if (need <= 256) return launch_dense_nt<T, 256>(logits, grad, B, Tfr, V, in_lens, lse2, gmul, st);
  if (need <= 512) return launch_dense_nt<T, 512>(logits, grad, B, Tfr, V, in_lens, lse2, gmul, st);
  if (need <= 1024) return launch_dense_nt<T, 1024>(logits, grad, B, Tfr, V, in_lens, lse2, gmul, st);
  set_error("ctc_loss: V=%d exceeds the supported row size", V);
  return B200CTC_ERR_ARG;
}

int launch_dense_rows(const void* logits, int dtype, void* grad, int B, int T, int V,
                      const int64_t* in_lens, const LossLayout& w, unsigned char* ws, float gmul,
                      cudaStream_t st) {
  float* lse2 = reinterpret_cast<float*>(ws + w.lse2);
  switch (dtype) {
    case B200CTC_F32: return launch_dense_t<float>(logits, grad, B, T, V, in_lens, lse2, gmul, st);
    case B200CTC_F16: return launch_dense_t<__half>(logits, grad, B, T, V, in_lens, lse2, gmul, st);
    default: return launch_dense_t<__nv_bfloat16>(logits, grad, B, T, V, in_lens, lse2, gmul, st);
  }
}

template <typename T>
static int launch_fixup_t(void* grad, int B, int Tfr, int V, int blank, const LossLayout& w,
                          const unsigned char* ws, float gmul, cudaStream_t st) {
  const int64_t rows = (int64_t)B * Tfr;
  int64_t grid = (rows + kFixWarps - 1) / kFixWarps;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (grid > cap) grid = cap;
  int wshift = 0;
  while ((1 << wshift) < 32 * w.R) ++wshift;
  const size_t smem = sizeof(float) * (size_t)kFixWarps * w.Lp;
  auto k = fixup_kernel<T>;
  if (smem > 48 * 1024) cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  k<<<(unsigned)grid, kFixWarps * 32, smem, st>>>(
      static_cast<T*>(grad), B, Tfr, V, blank, reinterpret_cast<const UttInfo*>(ws + w.utt),
      reinterpret_cast<const int*>(ws + w.lab), reinterpret_cast<const int*>(ws + w.dup), w.Lp,
      w.Sp, w.NW, w.NB, wshift, reinterpret_cast<const float*>(ws + w.lse2),
      reinterpret_cast<const float*>(ws + w.cmax), reinterpret_cast<const float*>(ws + w.em),
      reinterpret_cast<const float*>(ws + w.alpha), reinterpret_cast<const float*>(ws + w.beta),
      reinterpret_cast<const int*>(ws + w.offA), reinterpret_cast<const int*>(ws + w.offB), gmul);
  count_launch();
  return check_launch("ctc fixup");
}

int launch_fixup(void* grad, int dtype, int B, int T, int V, int blank, const LossLayout& w,
                 const unsigned char* ws, float gmul, cudaStream_t st) {
  switch (dtype) {
    case B200CTC_F32: return launch_fixup_t<float>(grad, B, T, V, blank, w, ws, gmul, st);
    case B200CTC_F16: return launch_fixup_t<__half>(grad, B, T, V, blank, w, ws, gmul, st);
    default: return launch_fixup_t<__nv_bfloat16>(grad, B, T, V, blank, w, ws, gmul, st);
  }
}

}  // namespace b200ctc
