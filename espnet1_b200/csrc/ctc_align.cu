// Forced alignment (espnet/nets/pytorch_backend/ctc.py:163-226): Viterbi over the
// blank-interleaved label lattice with back-pointers, then back-tracking.
// The reference keeps the lattice in fp64 (numpy default) with a finite
// log-zero of -1e11 and lets index -1 wrap to the last state for s = 0; both
// are reproduced so that the integer alignment is identical.
#include <mutex>

#include "common.cuh"

namespace b200ctc {

// Back-pointers take one of three values (stay, previous state - with numpy's wrap for s = 0 -,
// skip), so a state packs the 16 frames of a group into one 32-bit word of its own: T*S/4 bytes
// instead of T*S*4.  That fits shared memory for ordinary utterances (T=500, 150 labels: 38 KB),
// and the back-tracking - T dependent reads by one thread - then walks shared memory instead of
// global memory (it was half of the 0.41 ms a T=500 alignment took; the other half: one exposed
// global-memory round trip per frame for the emission, now fetched 8 frames ahead).
constexpr int kAlignPF = 8;      // frames of emissions in flight (one state per thread)

constexpr int kAlignGroup = 16;  // frames per packed word

__device__ __forceinline__ int align_prev(int s, int code, int S) {
  return code == 0 ? s : code == 1 ? (s >= 1 ? s - 1 : S - 1) : s - 2;
}

// Back along the pointers from state `s_end` (one thread), the visited states going to shared
// memory; then all threads turn states into labels and write the alignment.  Nothing on the
// T-long dependent chain touches global memory when the packed pointers are in shared memory
// (with `out[t] = label(s)` inside the chain every step waited for a global load).
__device__ __forceinline__ void align_backtrack(int s_end, int T, int S, int blank,
                                                const int64_t* __restrict__ target,
                                                const unsigned* back, int* seq, int64_t* __restrict__ out) {
  if (threadIdx.x == 0) {
    int s = s_end;
    for (int t = T - 1; t >= 0; --t) {
      seq[t] = s;
      if (t > 0) {
        B2_DEV_CHECK(s >= 0 && s < S);
        const unsigned w = back[(size_t)(t / kAlignGroup) * S + s];
        B2_DEV_CHECK(((w >> (2 * (t % kAlignGroup))) & 3u) < 3u);
        s = align_prev(s, (w >> (2 * (t % kAlignGroup))) & 3u, S);
      }
    }
  }
  __syncthreads();
  for (int t = threadIdx.x; t < T; t += blockDim.x) {
    const int s = seq[t];
    out[t] = (s & 1) ? target[(s - 1) >> 1] : (int64_t)blank;
  }
}

__global__ void __launch_bounds__(1024) align_kernel(const float* __restrict__ lp,
                                                     const int64_t* __restrict__ target, int T,
                                                     int V, int L, int blank,
                                                     int64_t* __restrict__ out,
                                                     int* __restrict__ back_global, int back_in_smem) {
  extern __shared__ double dl[];  // [2][S], then (back_in_smem) the packed back-pointers
  const int S = 2 * L + 1;
  const int NT = blockDim.x;
  int* seq = reinterpret_cast<int*>(dl + 2 * (size_t)S);  // T visited states (back-tracking)
  unsigned* back = back_in_smem ? reinterpret_cast<unsigned*>(seq + T)
                                : reinterpret_cast<unsigned*>(back_global);
  const double logzero = -100000000000.0;
  auto label = [&](int s) -> int { return (s & 1) ? (int)target[(s - 1) >> 1] : blank; };
  for (int s = threadIdx.x; s < S; s += NT) {
    dl[s] = (s < 2) ? (double)lp[label(s)] : logzero;  // delta[0,s] = lp[0][y_int[s]], s<2
  }
  __syncthreads();
  // one frame of one state; `em` = lp[t][label(s)]
  auto step = [&](int t, int s, int ls, bool can_skip, float em) {
    const double* cur = dl + ((t - 1) & 1) * S;
    double* nxt = dl + (t & 1) * S;
    const int s1 = s >= 1 ? s - 1 : S - 1;  // numpy's [-1] wrap for s = 0
    double best = cur[s];
    unsigned code = 0;
    if (cur[s1] > best) { best = cur[s1]; code = 1; }
    if (can_skip && cur[s - 2] > best) { best = cur[s - 2]; code = 2; }
    nxt[s] = best + (double)em;
    unsigned* w = back + (size_t)(t / kAlignGroup) * S + s;
    const unsigned bits = code << (2 * (t % kAlignGroup));
    *w = (t % kAlignGroup == 0 || t == 1) ? bits : (*w | bits);  // the word's first frame assigns
  };
  if (S <= NT) {
    // one state per thread: emissions kAlignPF frames ahead in registers
    const int s = threadIdx.x;
    const bool mine = s < S;
    const int ls = mine ? label(s) : blank;
    const bool can_skip = mine && !(ls == blank || s < 2 || ls == label(s - 2));
    const float* col = lp + ls;
    float em[kAlignPF];
#pragma unroll
    for (int j = 0; j < kAlignPF; ++j) em[j] = (1 + j < T) ? __ldg(col + (size_t)(1 + j) * V) : 0.f;
    for (int t0 = 1; t0 < T; t0 += kAlignPF) {
#pragma unroll
      for (int j = 0; j < kAlignPF; ++j) {
        const int t = t0 + j;
        if (t < T) {  // uniform over the CTA
          const float e = em[j];
          em[j] = (t + kAlignPF < T) ? __ldg(col + (size_t)(t + kAlignPF) * V) : 0.f;
          if (mine) step(t, s, ls, can_skip, e);
          __syncthreads();
        }
      }
    }
  } else {
    for (int t = 1; t < T; ++t) {
      for (int s = threadIdx.x; s < S; s += NT) {
        const int ls = label(s);
        step(t, s, ls, !(ls == blank || s < 2 || ls == label(s - 2)), lp[(size_t)t * V + ls]);
      }
      __syncthreads();
    }
  }
  const double* fin = dl + ((T - 1) & 1) * S;
  align_backtrack((S >= 2 && fin[S - 2] > fin[S - 1]) ? S - 2 : S - 1, T, S, blank, target, back, seq, out);
}

// Up to 512 states: FOUR warps, R = 2..4 consecutive states per thread in registers.  The CTA version above
// pays a block barrier and three shared-memory round trips per frame (0.47 us per frame at
// S = 301); here a frame is R independent compare/add chains per lane, the two values a lane
// needs from its left neighbour arrive by shuffle and numpy's wrap for s = 0 by one more.
// The emissions come through a shared-memory ring of 4-frame chunks filled by 1-D TMA copies
// (mbarrier completion) 28 frames ahead: a first version that fetched lp[t][label] itself, four
// frames ahead in registers, spent 44 % of its time waiting for those loads; and four warps
// (one per scheduler) share the states, exchanging a warp's last two values and the wrap
// value through shared memory with one block barrier per frame - a single warp was bound by
// its own instruction stream (R = 12..16 fp64 chains per lane).  What is left is the fp64
// dependent chain and the barrier (profiles/r02_align_raw.csv).  For that a small pre-pass gathers the emissions of the
// utterance's states into the workspace, one row of 32*R floats per frame in the order the
// lanes read them (word i*32 + lane = state lane*R + i: conflict-free, rows 16-byte multiples).
// Same fp64 values, same candidate order (stay, previous, skip: the first maximum wins), same
// packed back-pointers as the CTA version.
constexpr int kAlignCH = 4;   // frames per TMA chunk
constexpr int kAlignNCH = 8;  // chunks in the ring

__global__ void __launch_bounds__(512) align_gather_kernel(const float* __restrict__ lp,
                                                           const int64_t* __restrict__ target, int V, int L,
                                                           int blank, int R, float* __restrict__ em) {
  const int t = blockIdx.x, p = threadIdx.x;  // blockDim.x = NT * R, word p = i * NT + tid
  const int NT = blockDim.x / R;
  const int S = 2 * L + 1, s = (p % NT) * R + p / NT;
  float v = 0.f;
  if (s < S) v = lp[(size_t)t * V + ((s & 1) ? (int)target[(s - 1) >> 1] : blank)];
  em[(size_t)t * blockDim.x + p] = v;
}

constexpr int kAlignNW = 4;  // warps of the register-resident kernel (one per scheduler)

template <int R>
__global__ void __launch_bounds__(32 * kAlignNW) align_warp_kernel(const float* __restrict__ lp,
                                                                   const float* __restrict__ em_g,
                                                                   const int64_t* __restrict__ target, int T,
                                                                   int V, int L, int blank,
                                                                   int64_t* __restrict__ out) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  constexpr int NT = 32 * kAlignNW;
  constexpr int ROW = NT * R;                      // floats per frame
  constexpr int CHUNK = kAlignCH * ROW;            // floats per chunk
  float* ring = reinterpret_cast<float*>(smem_raw);
  uint64_t* bars = reinterpret_cast<uint64_t*>(ring + kAlignNCH * CHUNK);
  double* bnd = reinterpret_cast<double*>(bars + kAlignNCH);  // [2][kAlignNW][2] + [2] wrap value
  int* seq = reinterpret_cast<int*>(bnd + 2 * kAlignNW * 2 + 2);  // T visited states (back-tracking)
  unsigned* back = reinterpret_cast<unsigned*>(seq + T);          // packed back-pointers
  const int S = 2 * L + 1, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  const int n_chunks = (T + kAlignCH - 1) / kAlignCH;
  auto issue = [&](int ck) {  // thread 0: chunk ck -> its ring slot
    const int frames = min(kAlignCH, T - ck * kAlignCH);
    const uint32_t bytes = (uint32_t)frames * ROW * sizeof(float);
    uint64_t* bar = bars + ck % kAlignNCH;
    mbar_expect_tx(bar, bytes);
    bulk_g2s(ring + (ck % kAlignNCH) * CHUNK, em_g + (size_t)ck * CHUNK, bytes, bar);
  };
  if (tid == 0) {
    for (int i = 0; i < kAlignNCH; ++i) mbar_init(bars + i, 1);
    fence_mbar_init();
    for (int ck = 0; ck < kAlignNCH && ck < n_chunks; ++ck) issue(ck);
  }
  const double logzero = -100000000000.0;
  auto label = [&](int s) -> int { return (s & 1) ? (int)target[(s - 1) >> 1] : blank; };
  const int s0 = tid * R;
  unsigned skip = 0;  // bit i: state s0 + i has the skip transition
  double d[R];
#pragma unroll
  for (int i = 0; i < R; ++i) {
    const int s = s0 + i;
    const int l = s < S ? label(s) : blank;
    if (s < S && !(l == blank || s < 2 || l == label(s - 2))) skip |= 1u << i;
    d[i] = (s < 2 && s < S) ? (double)lp[l] : logzero;
  }
  const int own_last = (S - 1) / R, idx_last = (S - 1) - own_last * R;  // thread / slot of state S-1
  // frame 0's boundary values: a warp's last two states, and state S-1 (numpy's wrap for s = 0)
  double* wrapv = bnd + 2 * kAlignNW * 2;
  if (lane == 31) { bnd[(0 * kAlignNW + wid) * 2] = d[R - 1]; bnd[(0 * kAlignNW + wid) * 2 + 1] = d[R - 2]; }
  if (tid == own_last) {
    double v = logzero;
#pragma unroll
    for (int i = 0; i < R; ++i) v = (i == idx_last) ? d[i] : v;
    wrapv[0] = v;
  }
  unsigned bk[R];
#pragma unroll
  for (int i = 0; i < R; ++i) bk[i] = 0;
  __syncthreads();  // barriers initialised, frame 0's boundary values in place
  for (int t = 1; t < T; ++t) {
    const int ck = t / kAlignCH, slot = ck % kAlignNCH, pb = (t - 1) & 1, nb = t & 1;
    if (t % kAlignCH == 0 || t == 1) mbar_wait(bars + slot, (uint32_t)(ck / kAlignNCH) & 1u);
    const float* row = ring + slot * CHUNK + (t % kAlignCH) * ROW + tid;
    float em[R];
#pragma unroll
    for (int i = 0; i < R; ++i) em[i] = row[i * NT];
    double up1 = __shfl_up_sync(0xffffffffu, d[R - 1], 1);
    double up2 = __shfl_up_sync(0xffffffffu, d[R - 2], 1);
    if (lane == 0) {
      if (wid == 0) { up1 = wrapv[pb]; up2 = logzero; }
      else { up1 = bnd[(pb * kAlignNW + wid - 1) * 2]; up2 = bnd[(pb * kAlignNW + wid - 1) * 2 + 1]; }
    }
    double n[R];
#pragma unroll
    for (int i = 0; i < R; ++i) {
      const double c1 = i >= 1 ? d[i - 1] : up1;
      const double c2 = i >= 2 ? d[i - 2] : (i == 1 ? up1 : up2);
      // stay, then previous, then skip; the first maximum wins.  The three comparisons are
      // independent (one fp64 compare latency on the chain instead of two in a row)
      const bool p01 = c1 > d[i], p02 = c2 > d[i], p12 = c2 > c1;
      const bool take2 = ((skip >> i) & 1u) && (p01 ? p12 : p02);
      const double best = take2 ? c2 : (p01 ? c1 : d[i]);
      const unsigned code = take2 ? 2u : (p01 ? 1u : 0u);
      n[i] = best + (double)em[i];
      bk[i] |= code << (2 * (t % kAlignGroup));
    }
#pragma unroll
    for (int i = 0; i < R; ++i) d[i] = (s0 + i < S) ? n[i] : logzero;
    if (lane == 31) { bnd[(nb * kAlignNW + wid) * 2] = d[R - 1]; bnd[(nb * kAlignNW + wid) * 2 + 1] = d[R - 2]; }
    if (tid == own_last) {
      double v = logzero;
#pragma unroll
      for (int i = 0; i < R; ++i) v = (i == idx_last) ? d[i] : v;
      wrapv[nb] = v;
    }
    if (t % kAlignGroup == kAlignGroup - 1 || t == T - 1) {
#pragma unroll
      for (int i = 0; i < R; ++i) {
        if (s0 + i < S) back[(size_t)(t / kAlignGroup) * S + s0 + i] = bk[i];
        bk[i] = 0;
      }
    }
    __syncthreads();  // boundary values of frame t visible; frame t's emissions consumed by all
    if (tid == 0 && (t % kAlignCH == kAlignCH - 1) && ck + kAlignNCH < n_chunks) issue(ck + kAlignNCH);
  }
  // the better of the last two states (S-1 and S-2), then back along the pointers
  double* fin = bnd;  // reuse: [0] = delta[T-1][S-1], [1] = delta[T-1][S-2]
  __syncthreads();
#pragma unroll
  for (int i = 0; i < R; ++i) {
    if (s0 + i == S - 1) fin[0] = d[i];
    if (s0 + i == S - 2) fin[1] = d[i];
  }
  __syncthreads();
  align_backtrack((S >= 2 && fin[1] > fin[0]) ? S - 2 : S - 1, T, S, blank, target, back, seq, out);
}

// shared memory of the register-resident kernel: emission ring + barriers + boundary values +
// visited states + packed pointers
static size_t align_warp_smem(int R, int T, int S) {
  return (size_t)kAlignNCH * kAlignCH * 32 * kAlignNW * R * sizeof(float) + kAlignNCH * sizeof(uint64_t) +
         (2 * kAlignNW * 2 + 2) * sizeof(double) + (size_t)T * sizeof(int) +
         (size_t)((T + kAlignGroup - 1) / kAlignGroup) * S * sizeof(unsigned);
}
static int align_warp_R(int S) { return S <= 256 ? 2 : S <= 384 ? 3 : 4; }
// bytes of workspace the register-resident path needs for the gathered emissions (0: not taken)
static size_t align_warp_ws(int T, int S) {
  if (S > 512) return 0;
  const int R = align_warp_R(S);
  if (align_warp_smem(R, T, S) > 200 * 1024) return 0;
  return (size_t)T * 32 * kAlignNW * R * sizeof(float);
}

template <int R>
static void launch_align_warp(const float* lp, const float* em, const int64_t* target, int T, int V,
                              int tgt_len, int blank, int64_t* out_ids, size_t smem, cudaStream_t st) {
  if (smem > 48 * 1024)
    cudaFuncSetAttribute(align_warp_kernel<R>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  align_warp_kernel<R><<<1, 32 * kAlignNW, smem, st>>>(lp, em, target, T, V, tgt_len, blank, out_ids);
}

// launch: the packed back-pointers go to shared memory when they fit beside the lattice
static int launch_align(const float* lp, const int64_t* target, int T, int V, int tgt_len, int blank,
                        int64_t* out_ids, void* ws, size_t ws_bytes, cudaStream_t st, const char* what) {
  const int S = 2 * tgt_len + 1;
  const size_t need_em = align_warp_ws(T, S);
  if (need_em && need_em <= ws_bytes && (reinterpret_cast<uintptr_t>(ws) & 15) == 0) {
    // one warp, states in registers, emissions gathered into the workspace first
    const int R = align_warp_R(S);
    float* em = static_cast<float*>(ws);
    align_gather_kernel<<<T, 32 * kAlignNW * R, 0, st>>>(lp, target, V, tgt_len, blank, R, em);
    const size_t smem_w = align_warp_smem(R, T, S);
    if (R == 2) launch_align_warp<2>(lp, em, target, T, V, tgt_len, blank, out_ids, smem_w, st);
    else if (R == 3) launch_align_warp<3>(lp, em, target, T, V, tgt_len, blank, out_ids, smem_w, st);
    else launch_align_warp<4>(lp, em, target, T, V, tgt_len, blank, out_ids, smem_w, st);
    count_launch(2);
    return check_launch(what);
  }
  const size_t lattice = 2 * (size_t)S * sizeof(double);
  const size_t packed = (size_t)((T + kAlignGroup - 1) / kAlignGroup) * S * sizeof(unsigned);
  B2_REQUIRE(lattice <= 200 * 1024, "%s: target too long (%d)", what, tgt_len);
  const size_t seq_b = (size_t)T * sizeof(int);
  const int in_smem = lattice + seq_b + packed <= 200 * 1024;
  const size_t smem = lattice + seq_b + (in_smem ? packed : 0);
  B2_REQUIRE(smem <= 200 * 1024, "%s: input too long for shared memory (T=%d, %d labels)", what, T, tgt_len);
  B2_REQUIRE(in_smem || packed <= ws_bytes, "%s: workspace too small for the back-pointers", what);
  int nt = (S + 31) / 32 * 32;
  if (nt > 1024) nt = 1024;
  if (smem > 48 * 1024 &&
      cudaFuncSetAttribute(align_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem) != cudaSuccess) {
    set_error("%s: cannot reserve %zu bytes of shared memory", what, smem);
    cudaGetLastError();
    return B200CTC_ERR_DEVICE;
  }
  align_kernel<<<1, nt, smem, st>>>(lp, target, T, V, tgt_len, blank, out_ids, static_cast<int*>(ws), in_smem);
  count_launch();
  return check_launch(what);
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" size_t b200ctc_align_workspace_bytes(int T, int tgt_len) {
  if (T <= 0 || tgt_len < 0) return 0;
  const size_t base = (size_t)T * (2 * (size_t)tgt_len + 1) * sizeof(int);
  const size_t em = align_warp_ws(T, 2 * tgt_len + 1);  // gathered emissions of the one-warp path
  return (em > base ? em : base) + 64;
}

extern "C" int b200ctc_forced_align(const float* lp, const int64_t* target, int T, int V,
                                    int tgt_len, int blank, int64_t* out_ids, void* ws,
                                    size_t ws_bytes, b200ctc_stream_t stream) {
  B2_REQUIRE(lp && target && out_ids && ws, "forced_align: null pointer");
  B2_REQUIRE(T > 0 && V > 0 && tgt_len >= 1 && blank >= 0 && blank < V, "forced_align: bad shape");
  if (ws_bytes < b200ctc_align_workspace_bytes(T, tgt_len)) {
    set_error("forced_align: workspace too small");
    return B200CTC_ERR_WORKSPACE;
  }
  return launch_align(lp, target, T, V, tgt_len, blank, out_ids, ws, ws_bytes, static_cast<cudaStream_t>(stream),
                      "forced_align");
}

// ---- allocation-free variant ------------------------------------------------------------
// b200ctc_forced_align_static: the same kernel with its back-pointer scratch in a file-scope
// device array instead of a caller-provided workspace, for callers that cannot allocate (code
// under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  No
// counterpart in the reference.
//  * the array costs B200CTC_ALIGN_STATIC_MAX_BYTES on every GPU that loads the library,
//    until the process exits;
//  * in HOST code `g_align_scratch` is only a stand-in for the symbol, not a device address:
//    the address is resolved once per device (with that device current) by
//    cudaGetSymbolAddress below and that pointer is what the kernel receives;
//  * all calls on a device share the array and it keeps the previous call's contents.  Nothing
//    has to start at zero: the kernel writes back[t][s] for every t in [1,T), s in [0,S) before
//    thread 0 back-tracks through exactly those words, so stale contents are never read;
//  * calls that overlap on one GPU (two streams, two host threads) RACE on the array: the
//    caller must serialise them (one stream, or events).
static __device__ __align__(256) unsigned char g_align_scratch[B200CTC_ALIGN_STATIC_MAX_BYTES];

namespace b200ctc {
constexpr int kAlignMaxDevices = 64;
static std::mutex g_align_mu;
static void* g_align_ptr[kAlignMaxDevices] = {};

// device address of g_align_scratch on the CURRENT device (resolved once per device)
static int align_scratch_ptr(int** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= kAlignMaxDevices) {
    set_error("forced_align_static: no usable CUDA device (%s)",
              e != cudaSuccess ? cudaGetErrorString(e) : "device index out of range");
    return B200CTC_ERR_DEVICE;
  }
  std::lock_guard<std::mutex> lk(g_align_mu);
  if (!g_align_ptr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_align_scratch);
    if (e != cudaSuccess || !p) {
      set_error("forced_align_static: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
      return B200CTC_ERR_DEVICE;
    }
    g_align_ptr[dev] = p;
  }
  *out = static_cast<int*>(g_align_ptr[dev]);
  return B200CTC_OK;
}
}  // namespace b200ctc

extern "C" int b200ctc_forced_align_static(const float* lp, const int64_t* target, int T, int V,
                                           int tgt_len, int blank, int64_t* out_ids,
                                           b200ctc_stream_t stream) {
  B2_REQUIRE(lp && target && out_ids, "forced_align_static: null pointer");
  B2_REQUIRE(T > 0 && V > 0 && tgt_len >= 1 && blank >= 0 && blank < V,
             "forced_align_static: bad shape");
  const int S = 2 * tgt_len + 1;
  const size_t need = (size_t)T * (size_t)S * sizeof(int);
  if (need > (size_t)B200CTC_ALIGN_STATIC_MAX_BYTES) {
    set_error("forced_align_static: T=%d x (2*tgt_len+1)=%d back-pointers need %zu bytes, the static "
              "scratch holds %zu (B200CTC_ALIGN_STATIC_MAX_BYTES); use b200ctc_forced_align with a "
              "caller workspace", T, S, need, (size_t)B200CTC_ALIGN_STATIC_MAX_BYTES);
    return B200CTC_ERR_WORKSPACE;
  }
  int* back = nullptr;
  const int rc = align_scratch_ptr(&back);
  if (rc) return rc;
  return launch_align(lp, target, T, V, tgt_len, blank, out_ids, back, (size_t)B200CTC_ALIGN_STATIC_MAX_BYTES,
                      static_cast<cudaStream_t>(stream), "forced_align_static");
}
