// argmax over the vocabulary (first maximal index, NaN = maximum) and the
// greedy best-path decode built on it (collapse repeats, drop blanks).
#include "common.cuh"

namespace b200ctc {

// order-preserving key: larger float -> larger uint; NaN -> largest; -0 == +0
__device__ __forceinline__ uint32_t order_key(float f) {
  if (f != f) return 0xFFFFFFFFu;
  f += 0.0f;  // -0.0 -> +0.0
  const uint32_t u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
// pack (key, index) so that max() picks the largest key and, among equal keys,
// the smallest index
__device__ __forceinline__ unsigned long long pack(uint32_t key, uint32_t idx) {
  return ((unsigned long long)key << 32) | (unsigned long long)(0xFFFFFFFFu - idx);
}
__device__ __forceinline__ unsigned long long warp_max_u64(unsigned long long v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const unsigned long long w = __shfl_xor_sync(0xffffffffu, v, o);
    v = w > v ? w : v;
  }
  return v;
}

template <typename T, int VEC>
__device__ __forceinline__ unsigned long long row_argmax(const T* __restrict__ xr, int V,
                                                         unsigned long long* red,
                                                         const T* __restrict__ bias) {
  unsigned long long best = 0;
  if (VEC > 1) {
    const int nchunk = V / VEC;
    for (int i = threadIdx.x; i < nchunk; i += blockDim.x) {
      if (VEC == 4) {
        const float4 t = __ldcs(reinterpret_cast<const float4*>(xr) + i);
        float f[4] = {t.x, t.y, t.z, t.w};
        if (bias) {
          const float4 b = __ldg(reinterpret_cast<const float4*>(bias) + i);
          f[0] += b.x; f[1] += b.y; f[2] += b.z; f[3] += b.w;
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const unsigned long long p = pack(order_key(f[e]), i * 4 + e);
          best = p > best ? p : best;
        }
      } else {
        const uint4 t = __ldcs(reinterpret_cast<const uint4*>(xr) + i);
        const T* h = reinterpret_cast<const T*>(&t);
        uint4 tb = make_uint4(0, 0, 0, 0);
        if (bias) tb = __ldg(reinterpret_cast<const uint4*>(bias) + i);
        const T* hb = reinterpret_cast<const T*>(&tb);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float v = to_f32<T>(h[e]) + (bias ? to_f32<T>(hb[e]) : 0.f);
          const unsigned long long p = pack(order_key(v), i * 8 + e);
          best = p > best ? p : best;
        }
      }
    }
  } else {
    for (int i = threadIdx.x; i < V; i += blockDim.x) {
      const float v = to_f32<T>(xr[i]) + (bias ? to_f32<T>(bias[i]) : 0.f);
      const unsigned long long p = pack(order_key(v), i);
      best = p > best ? p : best;
    }
  }
  best = warp_max_u64(best);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = best;
  __syncthreads();
  best = (l < (blockDim.x >> 5)) ? red[l] : 0ull;
  best = warp_max_u64(best);
  __syncthreads();
  return best;
}

// Rows of up to 256*kArgChunks*VEC elements: the thread's share goes to registers with all
// its loads in flight, the maximum is a plain FMNMX chain (2 instructions per element instead
// of the 64-bit key/compare/select per element above), and only the thread's winner is packed
// for the block reduction.  NaN = maximum, first index among equals - as torch.argmax.
constexpr int kArgChunks = 5;

template <typename T, int VEC>
__device__ __forceinline__ unsigned long long row_argmax_regs(const T* __restrict__ xr, int nchunk,
                                                              unsigned long long* red,
                                                              const T* __restrict__ bias) {
  float v[kArgChunks][VEC];
#pragma unroll
  for (int j = 0; j < kArgChunks; ++j) {
    const int i = threadIdx.x + j * 256;
    if (i < nchunk) {
      if (VEC == 4) {
        const float4 t = __ldcs(reinterpret_cast<const float4*>(xr) + i);
        v[j][0] = t.x; v[j][1] = t.y; v[j][2] = t.z; v[j][3] = t.w;
        if (bias) {  // the projection's bias (see b200ctc_softmax_fwd_bias)
          const float4 b = __ldg(reinterpret_cast<const float4*>(bias) + i);
          v[j][0] += b.x; v[j][1] += b.y; v[j][2] += b.z; v[j][3] += b.w;
        }
      } else {
        const uint4 t = __ldcs(reinterpret_cast<const uint4*>(xr) + i);
        const T* h = reinterpret_cast<const T*>(&t);
#pragma unroll
        for (int e = 0; e < VEC; ++e) v[j][e] = to_f32<T>(h[e]);
        if (bias) {
          const uint4 tb = __ldg(reinterpret_cast<const uint4*>(bias) + i);
          const T* hb = reinterpret_cast<const T*>(&tb);
#pragma unroll
          for (int e = 0; e < VEC; ++e) v[j][e] += to_f32<T>(hb[e]);
        }
      }
    } else {
#pragma unroll
      for (int e = 0; e < VEC; ++e) v[j][e] = kNegInf;
    }
  }
  float m = kNegInf;
  bool nan = false;
#pragma unroll
  for (int j = 0; j < kArgChunks; ++j)
#pragma unroll
    for (int e = 0; e < VEC; ++e) {
      m = fmaxf(m, v[j][e]);  // (ignores NaN)
      nan |= v[j][e] != v[j][e];
    }
  uint32_t idx = 0xFFFFFFFFu;
#pragma unroll
  for (int j = kArgChunks - 1; j >= 0; --j)
#pragma unroll
    for (int e = VEC - 1; e >= 0; --e) {
      const bool hit = nan ? (v[j][e] != v[j][e]) : (v[j][e] == m);
      const uint32_t i = (uint32_t)((threadIdx.x + j * 256) * VEC + e);
      idx = (hit && threadIdx.x + j * 256 < nchunk) ? i : idx;  // descending scan: first hit wins
    }
  unsigned long long best = idx == 0xFFFFFFFFu ? 0ull : pack(nan ? 0xFFFFFFFFu : order_key(m), idx);
  best = warp_max_u64(best);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = best;
  __syncthreads();
  best = (l < 8) ? red[l] : 0ull;
  return warp_max_u64(best);
}

template <typename T, int VEC>
__global__ void __launch_bounds__(256) argmax_kernel(const T* __restrict__ x,
                                                     int64_t* __restrict__ out, int V,
                                                     const T* __restrict__ bias) {
  __shared__ unsigned long long red[32];
  const int64_t row = blockIdx.x;
  unsigned long long best;
  if (VEC > 1 && V / VEC <= 256 * kArgChunks) best = row_argmax_regs<T, VEC>(x + row * V, V / VEC, red, bias);
  else best = row_argmax<T, VEC>(x + row * V, V, red, bias);
  if (threadIdx.x == 0) out[row] = (int64_t)(0xFFFFFFFFu - (uint32_t)(best & 0xFFFFFFFFull));
}

// one CTA per utterance: in-place stream compaction of the frame-wise argmax
__global__ void __launch_bounds__(256) collapse_kernel(int64_t* __restrict__ ids,
                                                       const int64_t* __restrict__ in_lens,
                                                       int64_t* __restrict__ out_lens, int T,
                                                       int blank) {
  __shared__ int warp_tot[8];
  __shared__ int base_s;
  __shared__ long long last_s;
  const int b = blockIdx.x;
  int64_t* row = ids + (int64_t)b * T;
  int len = in_lens ? (int)in_lens[b] : T;
  len = len < 0 ? 0 : (len > T ? T : len);
  if (threadIdx.x == 0) { base_s = 0; last_s = -1; }
  __syncthreads();
  for (int t0 = 0; t0 < T; t0 += 256) {
    const int t = t0 + threadIdx.x;
    const long long cur = t < len ? row[t] : -1;
    const long long prev = (t > t0) ? (t - 1 < len ? row[t - 1] : -1) : last_s;
    const bool keep = t < len && cur != prev && cur != blank;
    const unsigned ball = __ballot_sync(0xffffffffu, keep);
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    const int within = __popc(ball & ((1u << l) - 1));
    if (l == 0) warp_tot[w] = __popc(ball);
    __syncthreads();  // all reads of this tile done, totals visible
    int off = base_s;
    for (int i = 0; i < w; ++i) off += warp_tot[i];
    int total = 0;
    for (int i = 0; i < 8; ++i) total += warp_tot[i];
    const long long tile_last = (t0 + 255 < len) ? row[t0 + 255] : -1;
    __syncthreads();  // tile_last read before any write of this tile
    if (keep) row[off + within] = cur;  // off+within <= t: never clobbers unread frames
    if (threadIdx.x == 0) { base_s += total; last_s = tile_last; }
    __syncthreads();
  }
  const int n = base_s;
  for (int t = n + threadIdx.x; t < T; t += 256) row[t] = -1;
  if (threadIdx.x == 0) out_lens[b] = n;
}

template <typename T>
static int launch_argmax(const void* x, int64_t* out, int64_t rows, int V, cudaStream_t st,
                         const void* bias = nullptr) {
  const T* xi = static_cast<const T*>(x);
  const T* bi = static_cast<const T*>(bias);
  constexpr int N = sizeof(T) == 4 ? 4 : 8;
  const bool vec = (V % N == 0) && ((reinterpret_cast<uintptr_t>(x) & 15) == 0) &&
                   ((reinterpret_cast<uintptr_t>(bias) & 15) == 0);
  if (vec)
    argmax_kernel<T, N><<<(unsigned)rows, 256, 0, st>>>(xi, out, V, bi);
  else
    argmax_kernel<T, 1><<<(unsigned)rows, 256, 0, st>>>(xi, out, V, bi);
  count_launch();
  return check_launch("argmax");
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" int b200ctc_argmax(const void* x, int64_t* out, int64_t rows, int V, int dtype,
                              b200ctc_stream_t stream) {
  B2_REQUIRE(x && out, "argmax: null pointer");
  B2_REQUIRE(rows >= 0 && rows < (1ll << 31) && V > 0, "argmax: bad shape");
  if (rows == 0) return B200CTC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200CTC_F32: return launch_argmax<float>(x, out, rows, V, st);
    case B200CTC_F16: return launch_argmax<__half>(x, out, rows, V, st);
    case B200CTC_BF16: return launch_argmax<__nv_bfloat16>(x, out, rows, V, st);
  }
  set_error("argmax: unknown dtype %d", dtype);
  return B200CTC_ERR_ARG;
}

extern "C" int b200ctc_argmax_bias(const void* x, const void* bias, int64_t* out, int64_t rows, int V,
                                   int dtype, b200ctc_stream_t stream) {
  B2_REQUIRE(x && bias && out, "argmax_bias: null pointer");
  B2_REQUIRE(rows >= 0 && rows < (1ll << 31) && V > 0, "argmax_bias: bad shape");
  if (rows == 0) return B200CTC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200CTC_F32: return launch_argmax<float>(x, out, rows, V, st, bias);
    case B200CTC_F16: return launch_argmax<__half>(x, out, rows, V, st, bias);
    case B200CTC_BF16: return launch_argmax<__nv_bfloat16>(x, out, rows, V, st, bias);
  }
  set_error("argmax_bias: unknown dtype %d", dtype);
  return B200CTC_ERR_ARG;
}

extern "C" int b200ctc_greedy_decode(const void* x, const int64_t* in_lens, int B, int T, int V,
                                     int dtype, int blank, int64_t* out_ids, int64_t* out_lens,
                                     b200ctc_stream_t stream) {
  B2_REQUIRE(x && out_ids && out_lens, "greedy_decode: null pointer");
  B2_REQUIRE(B >= 0 && T >= 0 && V > 0, "greedy_decode: bad shape");
  if (B == 0) return B200CTC_OK;
  int rc = T > 0 ? b200ctc_argmax(x, out_ids, (int64_t)B * T, V, dtype, stream) : B200CTC_OK;
  if (rc != B200CTC_OK) return rc;
  collapse_kernel<<<B, 256, 0, static_cast<cudaStream_t>(stream)>>>(out_ids, in_lens, out_lens, T,
                                                                     blank);
  count_launch();
  return check_launch("greedy_decode");
}
