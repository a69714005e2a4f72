// Shared device/host helpers for libb200ctc (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/b200ctc.h"

namespace b200ctc {

// ---------------------------------------------------------------- host side
void set_error(const char* fmt, ...);
void count_launch(int n = 1);
int check_launch(const char* what);  // cudaGetLastError -> status
// prefix.cu: b200ctc_prefix_step with a strided last-token array (beam_position.cu)
int prefix_step_impl(const float* lp, const float* lp_t, const float* r_prev, const float* s_prev,
                     const int64_t* last_ids, long long last_stride, const int64_t* scoring_ids,
                     const int64_t* end_frames, int B, int n_hyps, int T, int V, int snum, int out_len,
                     int start, int end, int blank, int eos, float logzero, float* r, float* log_psi,
                     float* scores, int64_t* idmap, float* rsum_ws, int lp_t_ld, const int* win_utt,
                     int r_time_major, b200ctc_stream_t stream);
bool prefix_scan_applies(bool have_lp_t, bool have_ids, int n_bh, int snum, int T);
int prefix_select_impl(const float* r, const float* log_psi, const int64_t* idmap, const int64_t* best_ids,
                       int B, int n_hyps, int W, int T, int V, int snum, float* r_new, float* s_new,
                       int r_time_major, b200ctc_stream_t stream);
int sm_count();
// profiling phases: 0 prep, 1 lse_gather (pass A), 2 lattice, 3 grad_rows (pass B),
// 4 prefix_step, 5 softmax family
constexpr int kProfPhases = 8;
void prof_mark(int phase, int end, cudaStream_t st);
// Tuning switches exist only in a -DB200CTC_TUNING build (B200CTC_NVCC_EXTRA of build.py):
// the shipped library reads no environment variable and has ONE code path per shape.
#ifdef B200CTC_TUNING
static inline const char* tuning_env(const char* name) { return getenv(name); }
#else
static inline const char* tuning_env(const char*) { return nullptr; }
#endif
static inline int tuning_int(const char* name, int dflt) {
  const char* e = tuning_env(name);
  return e ? atoi(e) : dflt;
}
// where a kernel fires griddepcontrol.launch_dependents (tuning: B200CTC_PDL_MODE):
// 0 = at its start, 1 = at its end, 2 = never (implicit at completion)
static inline int pdl_mode() {
  static const int m = tuning_int("B200CTC_PDL_MODE", 2);
  return m;
}

static inline bool pdl_enabled(int pdl_bit) {
  static const int pdl_mask = tuning_int("B200CTC_PDL", 63);
  return pdl_bit >= 0 && ((pdl_mask >> pdl_bit) & 1);
}

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                     cudaStream_t st, int pdl_bit, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  // Programmatic stream serialisation on the lattice (bit 0), the gradient row pass
  // (bit 1) and nll (bit 2).  Measured on B200 at C2 (profiles/README.md): with the
  // trigger at the START of the primary, early-resident lattice CTAs land several per SM
  // beside the running gather and the latency-bound recursion shares issue slots
  // (0.405 ms/step); with the trigger at the end, or never (implicit at completion, the
  // default), the kernel boundaries just get cheaper: 0.386 vs 0.393 ms without.
  // bits: 0 lattice, 1 gradient row pass, 2 nll, 3 emission gather, 4 rescale check,
  // 5 the kernels of a beam-search position (pre-beam, prefix scoring, pruning, state gather)
  cfg.numAttrs = pdl_enabled(pdl_bit) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

#define B2_REQUIRE(cond, ...)        \
  do {                               \
    if (!(cond)) {                   \
      b200ctc::set_error(__VA_ARGS__); \
      return B200CTC_ERR_ARG;        \
    }                                \
  } while (0)

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }
static inline int dtype_size(int dtype) { return dtype == B200CTC_F32 ? 4 : 2; }

// -------------------------------------------------------------- device side
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kNegInf = -__builtin_huge_valf();

__device__ __forceinline__ float ex2(float x) {  // MUFU.EX2, 2 ulp, ex2(-inf)=0
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2(float x) {  // MUFU.LG2
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// ---- programmatic dependent launch (PDL) ------------------------------------
// A kernel launched with launch_pdl(..., pdl=true) may start (become resident,
// run its prologue) while the previous kernel of the stream is still running;
// it must call griddep_wait() before touching anything that kernel wrote.
// Both instructions are no-ops in a normally launched kernel.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// the small kernels of a beam-search position: wait, then (tuning: -DB200CTC_BEAM_PDL_EARLY=1)
// let the next kernel's CTAs become resident at once - they park at their own wait
#ifndef B200CTC_BEAM_PDL_EARLY
#define B200CTC_BEAM_PDL_EARLY 0
#endif
__device__ __forceinline__ void griddep_wait_chain() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
#if B200CTC_BEAM_PDL_EARLY
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
#endif
}
__device__ __forceinline__ void griddep_launch() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}

// Index checks of a -DB200CTC_BOUNDS build (compute-sanitizer is not available on the GPU pool:
// the checked build runs the GPU tests instead; a violated check traps, which the next CUDA call
// reports).  Compiled out of the shipped library.
#ifdef B200CTC_BOUNDS
#define B2_DEV_CHECK(cond) do { if (!(cond)) __trap(); } while (0)
#else
#define B2_DEV_CHECK(cond) do { } while (0)
#endif

// ---- mbarrier + 1-D bulk async copy (TMA engine, UBLKCP in SASS) -----------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra DONE_%=;\n"
      "bra WAIT_%=;\n"
      "DONE_%=:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}
// global -> shared bulk copy; bytes % 16 == 0, both addresses 16-B aligned
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src, uint32_t bytes,
                                         uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
          smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
}
// ---- L2 eviction-priority policies (createpolicy; carried by the .L2::cache_hint forms) ----
// The loss step is DRAM-bound end to end (profiles/README.md, round 2): what stays in the
// 126 MB L2 between its kernels is worth 10 % of the traffic.  kind: 0 normal, 1 evict_first
// (streamed once), 2 evict_last (scratch that a later kernel re-reads).
__device__ __forceinline__ uint64_t l2_policy(int kind) {
  uint64_t p;
  if (kind == 1) asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  else if (kind == 2) asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  else asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void bulk_g2s_hint(void* dst_smem, const void* src, uint32_t bytes,
                                              uint64_t* bar, uint64_t pol) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst_smem)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(pol)
      : "memory");
}
// drop a 128-byte line from L2 WITHOUT writing it back (scratch that was read for the last time)
__device__ __forceinline__ void l2_discard_128(const void* p) {
  asm volatile("discard.global.L2 [%0], 128;" ::"l"(p) : "memory");
}
// shared -> global bulk copy (bulk async-group completion)
__device__ __forceinline__ void bulk_s2g(void* dst, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst),
               "r"(smem_u32(src_smem)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
__device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
// make generic-proxy smem writes visible to the async (TMA) proxy
__device__ __forceinline__ void fence_async_smem() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}

}  // namespace b200ctc
