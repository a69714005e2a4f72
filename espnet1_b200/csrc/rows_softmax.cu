// Vocabulary soft-max family: log_softmax / softmax forward and backward.
// One CTA per (b,t) row, the whole row held in registers as 128-bit chunks,
// warp-shuffle max/sum reductions.  HBM-bound: 1 read + 1 write per element.
#include "common.cuh"

namespace b200ctc {

template <typename T>
struct Chunk;  // 16-byte chunk of T
template <>
struct Chunk<float> {
  static constexpr int N = 4;
  float v[4];
  __device__ __forceinline__ void load(const float* p) {
    float4 t = __ldcs(reinterpret_cast<const float4*>(p));
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
  }
  __device__ __forceinline__ static void store(float* p, const float* f) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(f[0], f[1], f[2], f[3]));
  }
};
template <typename H>
struct Chunk16 {
  static constexpr int N = 8;
  float v[8];
  __device__ __forceinline__ void load(const H* p) {
    uint4 t = __ldcs(reinterpret_cast<const uint4*>(p));
    const H* h = reinterpret_cast<const H*>(&t);
#pragma unroll
    for (int i = 0; i < 8; ++i) v[i] = to_f32<H>(h[i]);
  }
  __device__ __forceinline__ static void store(H* p, const float* f) {
    uint4 t;
    H* h = reinterpret_cast<H*>(&t);
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = from_f32<H>(f[i]);
    __stcs(reinterpret_cast<uint4*>(p), t);
  }
};
template <>
struct Chunk<__half> : Chunk16<__half> {};
template <>
struct Chunk<__nv_bfloat16> : Chunk16<__nv_bfloat16> {};

// block-wide reductions; `red` has 32 floats of shared memory per use
template <int NT>
__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  if (NT > 32) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) red[w] = v;
    __syncthreads();
    v = (l < NT / 32) ? red[l] : kNegInf;
    v = warp_max(v);
    __syncthreads();
  }
  return v;
}
template <int NT>
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  if (NT > 32) {
    const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    if (l == 0) red[w] = v;
    __syncthreads();
    v = (l < NT / 32) ? red[l] : 0.f;
    v = warp_sum(v);
    __syncthreads();
  }
  return v;
}

// ---------------------------------------------------------------- forward
template <typename T, int NT, int PER, int MODE>
__global__ void __launch_bounds__(NT) softmax_fwd_vec(const T* __restrict__ x, T* __restrict__ y,
                                                      int V, const T* __restrict__ bias) {
  using C = Chunk<T>;
  __shared__ float red[32];
  const int64_t row = blockIdx.x;
  const T* xr = x + row * V;
  T* yr = y + row * V;
  const int nchunk = V / C::N;
  C c[PER];
  float m = kNegInf;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = threadIdx.x + j * NT;
    if (i < nchunk) {
      c[j].load(xr + i * C::N);
      if (bias) {  // the projection's bias, added here instead of by a GEMM epilogue pass
        C bc;
        bc.load(bias + i * C::N);
#pragma unroll
        for (int e = 0; e < C::N; ++e) c[j].v[e] += bc.v[e];
      }
#pragma unroll
      for (int e = 0; e < C::N; ++e) m = fmaxf(m, c[j].v[e]);
    }
  }
  m = block_max<NT>(m, red);
  const float m2 = (m == kNegInf || m == -kNegInf) ? 0.f : m * kLog2e;
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = threadIdx.x + j * NT;
    if (i < nchunk) {
#pragma unroll
      for (int e = 0; e < C::N; ++e) {
        const float p = ex2(fmaf(c[j].v[e], kLog2e, -m2));
        if (MODE == 1) c[j].v[e] = p;
        s += p;
      }
    }
  }
  s = block_sum<NT>(s, red);
  const float mm = (m == kNegInf || m == -kNegInf) ? 0.f : m;
  const float shift = mm + logf(s);
  const float inv = 1.f / s;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = threadIdx.x + j * NT;
    if (i < nchunk) {
      float o[C::N];
#pragma unroll
      for (int e = 0; e < C::N; ++e) o[e] = MODE == 1 ? c[j].v[e] * inv : c[j].v[e] - shift;
      C::store(yr + i * C::N, o);
    }
  }
}

// generic fallback: any V / alignment, three passes (row stays in L1/L2)
template <typename T, int MODE>
__global__ void __launch_bounds__(256) softmax_fwd_any(const T* __restrict__ x, T* __restrict__ y,
                                                       int V, const T* __restrict__ bias) {
  __shared__ float red[32];
  const int64_t row = blockIdx.x;
  const T* xr = x + row * V;
  T* yr = y + row * V;
  auto at = [&](int i) { return to_f32<T>(xr[i]) + (bias ? to_f32<T>(bias[i]) : 0.f); };
  float m = kNegInf;
  for (int i = threadIdx.x; i < V; i += 256) m = fmaxf(m, at(i));
  m = block_max<256>(m, red);
  const float mm = (m == kNegInf || m == -kNegInf) ? 0.f : m;
  float s = 0.f;
  for (int i = threadIdx.x; i < V; i += 256) s += ex2((at(i) - mm) * kLog2e);
  s = block_sum<256>(s, red);
  const float shift = mm + logf(s), inv = 1.f / s;
  for (int i = threadIdx.x; i < V; i += 256) {
    const float v = at(i);
    yr[i] = from_f32<T>(MODE == 1 ? ex2((v - mm) * kLog2e) * inv : v - shift);
  }
}

// --------------------------------------------------------------- backward
// log_softmax: dx = dy - exp(y) * sum(dy) ; softmax: dx = y * (dy - sum(dy*y))
template <typename T, int NT, int PER, int MODE>
__global__ void __launch_bounds__(NT) softmax_bwd_vec(const T* __restrict__ y,
                                                      const T* __restrict__ dy,
                                                      T* __restrict__ dx, int V) {
  using C = Chunk<T>;
  __shared__ float red[32];
  const int64_t row = blockIdx.x;
  const int nchunk = V / C::N;
  C cy[PER], cd[PER];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = threadIdx.x + j * NT;
    if (i < nchunk) {
      cy[j].load(y + row * V + i * C::N);
      cd[j].load(dy + row * V + i * C::N);
#pragma unroll
      for (int e = 0; e < C::N; ++e) s += MODE == 1 ? cd[j].v[e] * cy[j].v[e] : cd[j].v[e];
    }
  }
  s = block_sum<NT>(s, red);
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const int i = threadIdx.x + j * NT;
    if (i < nchunk) {
      float o[C::N];
#pragma unroll
      for (int e = 0; e < C::N; ++e)
        o[e] = MODE == 1 ? cy[j].v[e] * (cd[j].v[e] - s)
                         : cd[j].v[e] - ex2(cy[j].v[e] * kLog2e) * s;
      C::store(dx + row * V + i * C::N, o);
    }
  }
}

template <typename T, int MODE>
__global__ void __launch_bounds__(256) softmax_bwd_any(const T* __restrict__ y,
                                                       const T* __restrict__ dy,
                                                       T* __restrict__ dx, int V) {
  __shared__ float red[32];
  const int64_t o = (int64_t)blockIdx.x * V;
  float s = 0.f;
  for (int i = threadIdx.x; i < V; i += 256) {
    const float d = to_f32<T>(dy[o + i]);
    s += MODE == 1 ? d * to_f32<T>(y[o + i]) : d;
  }
  s = block_sum<256>(s, red);
  for (int i = threadIdx.x; i < V; i += 256) {
    const float d = to_f32<T>(dy[o + i]), yy = to_f32<T>(y[o + i]);
    dx[o + i] = from_f32<T>(MODE == 1 ? yy * (d - s) : d - ex2(yy * kLog2e) * s);
  }
}

// --------------------------------------------------------------- dispatch
static inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

template <typename T, int MODE>
static int launch_fwd(const void* x, void* y, int64_t rows, int V, cudaStream_t st,
                      const void* bias = nullptr) {
  constexpr int N = Chunk<T>::N;
  constexpr int PER = (N == 4) ? 8 : 4;
  const T* xi = static_cast<const T*>(x);
  T* yo = static_cast<T*>(y);
  const T* bi = static_cast<const T*>(bias);
  const bool vec = (V % N == 0) && aligned16(x) && aligned16(y) && aligned16(bias);
  const int nchunk = V / N;
  dim3 grid((unsigned)rows);
  if (vec && nchunk <= 128 * PER) {
    softmax_fwd_vec<T, 128, PER, MODE><<<grid, 128, 0, st>>>(xi, yo, V, bi);
  } else if (vec && nchunk <= 256 * PER) {
    softmax_fwd_vec<T, 256, PER, MODE><<<grid, 256, 0, st>>>(xi, yo, V, bi);
  } else if (vec && nchunk <= 512 * PER) {
    softmax_fwd_vec<T, 512, PER, MODE><<<grid, 512, 0, st>>>(xi, yo, V, bi);
  } else if (vec && nchunk <= 1024 * PER) {
    softmax_fwd_vec<T, 1024, PER, MODE><<<grid, 1024, 0, st>>>(xi, yo, V, bi);
  } else {
    softmax_fwd_any<T, MODE><<<grid, 256, 0, st>>>(xi, yo, V, bi);
  }
  count_launch();
  return check_launch("softmax_fwd");
}

template <typename T, int MODE>
static int launch_bwd(const void* y, const void* dy, void* dx, int64_t rows, int V,
                      cudaStream_t st) {
  constexpr int N = Chunk<T>::N;
  constexpr int PER = (N == 4) ? 6 : 3;
  const T* yi = static_cast<const T*>(y);
  const T* di = static_cast<const T*>(dy);
  T* xo = static_cast<T*>(dx);
  const bool vec = (V % N == 0) && aligned16(y) && aligned16(dy) && aligned16(dx);
  const int nchunk = V / N;
  dim3 grid((unsigned)rows);
  if (vec && nchunk <= 256 * PER) {
    softmax_bwd_vec<T, 256, PER, MODE><<<grid, 256, 0, st>>>(yi, di, xo, V);
  } else if (vec && nchunk <= 1024 * PER) {
    softmax_bwd_vec<T, 1024, PER, MODE><<<grid, 1024, 0, st>>>(yi, di, xo, V);
  } else {
    softmax_bwd_any<T, MODE><<<grid, 256, 0, st>>>(yi, di, xo, V);
  }
  count_launch();
  return check_launch("softmax_bwd");
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" int b200ctc_softmax_fwd(const void* x, void* y, int64_t rows, int V, int dtype, int mode,
                                   b200ctc_stream_t stream) {
  B2_REQUIRE(x && y, "softmax_fwd: null pointer");
  B2_REQUIRE(rows >= 0 && rows < (1ll << 31) && V > 0, "softmax_fwd: bad shape rows=%lld V=%d",
             (long long)rows, V);
  B2_REQUIRE(mode == 0 || mode == 1, "softmax_fwd: mode must be 0 or 1");
  if (rows == 0) return B200CTC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype * 2 + mode) {
    case 0: return launch_fwd<float, 0>(x, y, rows, V, st);
    case 1: return launch_fwd<float, 1>(x, y, rows, V, st);
    case 2: return launch_fwd<__half, 0>(x, y, rows, V, st);
    case 3: return launch_fwd<__half, 1>(x, y, rows, V, st);
    case 4: return launch_fwd<__nv_bfloat16, 0>(x, y, rows, V, st);
    case 5: return launch_fwd<__nv_bfloat16, 1>(x, y, rows, V, st);
  }
  set_error("softmax_fwd: unknown dtype %d", dtype);
  return B200CTC_ERR_ARG;
}

extern "C" int b200ctc_softmax_fwd_bias(const void* x, const void* bias, void* y, int64_t rows, int V,
                                        int dtype, int mode, b200ctc_stream_t stream) {
  B2_REQUIRE(x && y && bias, "softmax_fwd_bias: null pointer");
  B2_REQUIRE(rows >= 0 && rows < (1ll << 31) && V > 0, "softmax_fwd_bias: bad shape rows=%lld V=%d",
             (long long)rows, V);
  B2_REQUIRE(mode == 0 || mode == 1, "softmax_fwd_bias: mode must be 0 or 1");
  if (rows == 0) return B200CTC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype * 2 + mode) {
    case 0: return launch_fwd<float, 0>(x, y, rows, V, st, bias);
    case 1: return launch_fwd<float, 1>(x, y, rows, V, st, bias);
    case 2: return launch_fwd<__half, 0>(x, y, rows, V, st, bias);
    case 3: return launch_fwd<__half, 1>(x, y, rows, V, st, bias);
    case 4: return launch_fwd<__nv_bfloat16, 0>(x, y, rows, V, st, bias);
    case 5: return launch_fwd<__nv_bfloat16, 1>(x, y, rows, V, st, bias);
  }
  set_error("softmax_fwd_bias: unknown dtype %d", dtype);
  return B200CTC_ERR_ARG;
}

extern "C" int b200ctc_softmax_bwd(const void* y, const void* dy, void* dx, int64_t rows, int V,
                                   int dtype, int mode, b200ctc_stream_t stream) {
  B2_REQUIRE(y && dy && dx, "softmax_bwd: null pointer");
  B2_REQUIRE(rows >= 0 && rows < (1ll << 31) && V > 0, "softmax_bwd: bad shape");
  B2_REQUIRE(mode == 0 || mode == 1, "softmax_bwd: mode must be 0 or 1");
  if (rows == 0) return B200CTC_OK;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype * 2 + mode) {
    case 0: return launch_bwd<float, 0>(y, dy, dx, rows, V, st);
    case 1: return launch_bwd<float, 1>(y, dy, dx, rows, V, st);
    case 2: return launch_bwd<__half, 0>(y, dy, dx, rows, V, st);
    case 3: return launch_bwd<__half, 1>(y, dy, dx, rows, V, st);
    case 4: return launch_bwd<__nv_bfloat16, 0>(y, dy, dx, rows, V, st);
    case 5: return launch_bwd<__nv_bfloat16, 1>(y, dy, dx, rows, V, st);
  }
  set_error("softmax_bwd: unknown dtype %d", dtype);
  return B200CTC_ERR_ARG;
}
