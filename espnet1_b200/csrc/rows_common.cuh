// Helpers shared by the two streaming row passes of the CTC loss (pass A:
// lse + emission gather, pass B: fused gradient).
#pragma once
#include <stdlib.h>

#include "ctc_ws.cuh"

namespace b200ctc {

template <typename T>
struct Elems { static constexpr int N = sizeof(T) == 4 ? 4 : 8; };  // elements per 16 B

template <typename T>
__device__ __forceinline__ void unpack16(const uint4& raw, float* f) {
  if (sizeof(T) == 4) {
    f[0] = __uint_as_float(raw.x); f[1] = __uint_as_float(raw.y);
    f[2] = __uint_as_float(raw.z); f[3] = __uint_as_float(raw.w);
  } else {
    const T* h = reinterpret_cast<const T*>(&raw);
#pragma unroll
    for (int i = 0; i < 8; ++i) f[i] = to_f32<T>(h[i]);
  }
}
template <typename T>
__device__ __forceinline__ void store16(T* p, const float* f) {
  if (sizeof(T) == 4) {
    __stcs(reinterpret_cast<float4*>(p), make_float4(f[0], f[1], f[2], f[3]));
  } else {
    uint4 t;
    T* h = reinterpret_cast<T*>(&t);
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = from_f32<T>(f[i]);
    __stcs(reinterpret_cast<uint4*>(p), t);
  }
}

// (b,t) of the rows a persistent CTA visits: g = first, first+stride, ...
// kept incrementally (no 64-bit division per row)
struct RowCursor {
  int b, t, T, stride_b, stride_t;
  int64_t g;
  int64_t stride;
  __device__ __forceinline__ void init(int64_t first, int64_t step, int Tfr) {
    T = Tfr; g = first; stride = step;
    b = (int)(first / Tfr); t = (int)(first - (int64_t)b * Tfr);
    stride_b = (int)(step / Tfr); stride_t = (int)(step - (int64_t)stride_b * Tfr);
  }
  __device__ __forceinline__ void next() {
    g += stride; b += stride_b; t += stride_t;
    if (t >= T) { t -= T; ++b; }
  }
};

constexpr int kMaxRowStages = 8;

struct RowLaunchCfg { int nstage, ctas_per_sm, stride; bool tma; };

// shared-memory budgeting for the row passes; env overrides are for tuning sweeps only
static inline RowLaunchCfg row_launch_cfg(const void* base, size_t row_bytes, size_t extra_smem,
                                          int nt, const char* env_prefix, int want_ctas_hint = 0,
                                          int want_stages_hint = 0) {
  RowLaunchCfg c;
  c.tma = (row_bytes % 16 == 0) && ((reinterpret_cast<uintptr_t>(base) & 15) == 0);
  c.stride = (int)align_up(row_bytes, 128);
  const size_t budget = 220 * 1024;  // of the 227 KB an SM can hand to CTAs
  // measured on B200 (tools/sweep_rows.py, C2): many single-stage CTAs beat fewer
  // double-buffered ones - the other CTAs of the SM are the pipeline
  int want_ctas = nt <= 128 ? 6 : (nt <= 256 ? 3 : 1);
  int want_stages = nt <= 128 ? 1 : 2;
  if (want_ctas_hint > 0) want_ctas = want_ctas_hint;
  if (want_stages_hint > 0) want_stages = want_stages_hint;
  char name[64];
  snprintf(name, sizeof(name), "%s_CTAS", env_prefix);
  if (const char* e = tuning_env(name)) want_ctas = atoi(e);
  snprintf(name, sizeof(name), "%s_STAGES", env_prefix);
  if (const char* e = tuning_env(name)) want_stages = atoi(e);
  if (want_ctas < 1) want_ctas = 1;
  if (want_stages < 1) want_stages = 1;
  if (want_stages > kMaxRowStages) want_stages = kMaxRowStages;
  if (!c.tma) want_stages = 1;
  // shrink until it fits (1 KB per CTA reserved by the driver)
  while (want_ctas > 1 && want_ctas * ((size_t)want_stages * c.stride + extra_smem + 1024) > budget) {
    if (want_stages > 2) --want_stages; else --want_ctas;
  }
  while (want_stages > 1 && (size_t)want_stages * c.stride + extra_smem + 1024 > budget) --want_stages;
  c.ctas_per_sm = want_ctas;
  c.nstage = want_stages;
  return c;
}

}  // namespace b200ctc
