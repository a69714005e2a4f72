// Sharded-loss exchange over NVLink peer memory (the one exchange step of the path).
//
// The reference all-reduces two numbers per training step - the rank's summed loss and
// its weight (utterance count) - espnet2/train/trainer.py:604-613 with
// espnet2/torch_utils/recursive_op.py:8-47.  Here that collective is the TAIL OF THE
// KERNEL THAT PRODUCES THE LOSS: the last CTA of nll_kernel pushes the rank's
// {sum, weight} into every peer's exchange buffer with plain peer stores (one 8-byte
// payload + a sequence word per peer, straight through NVSwitch), then polls its OWN
// buffer until all ranks' words for this step have landed, and adds them up in rank
// order - the same order on every rank, so the result is bit-identical everywhere.
// No NCCL launch, no host work per step, one-shot latency of a single NVLink hop.
//
// Buffer of a rank (cudaMalloc + cudaIpc, mapped by every peer):
//   header {u32 pushed step, u32 error, u32 consumed step}, then slots[2 parities][world] x
//   {f32 sum, u32 seq, f32 weight, u32 seq}: two 8-byte words that each carry the step they
//   belong to.  An aligned 8-byte store is single-copy atomic, so every word VALIDATES ITSELF and
//   the push needs no release fence (a `st.release.sys` first makes all earlier writes of the CTA
//   visible system-wide: ~6 us per step, measured at N=2 as the whole remaining overhead).
//
// DEFERRED mode (the default of sharded.PeerLossExchange): the kernel that finishes the loss
// only PUSHES; the wait for the peers' words - a barrier on the slowest rank that nothing on
// the device needs inside the step - moves to the consumer: the next step's exchange consumes
// the previous step first (its words arrived a whole step ago: no wait in steady state), and
// b200ctc_peer_wait consumes the latest step when the result is read.  A rank therefore still
// cannot push step s+1 before it has consumed step s, i.e. before every peer pushed s, and a
// peer pushes s only after consuming s-1: ranks stay within one step of each other and parity
// double-buffering remains sufficient.
#pragma once
#include "common.cuh"

namespace b200ctc {

constexpr int kMaxPeers = 64;
constexpr int kPeerSlotBytes = 16;
constexpr int kPeerHeaderBytes = 64;
constexpr size_t kPeerBufBytes = kPeerHeaderBytes + 2 * kMaxPeers * kPeerSlotBytes;
constexpr unsigned long long kPeerDefaultTimeoutNs = 30ull * 1000 * 1000 * 1000;  // then NaN + error flag

struct PeerDev {  // passed by value to kernels
  unsigned char* const* bufs;  // device array [world]: every rank's buffer (bufs[rank] = mine)
  int world, rank;
  int deferred;                   // 1: push only, consume later (see above)
  unsigned long long timeout_ns;  // 0 = kPeerDefaultTimeoutNs
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
// {value bits, seq} as ONE relaxed system-scope 8-byte access
__device__ __forceinline__ void st_relaxed_sys_pair(void* p, float v, unsigned seq) {
  asm volatile("st.relaxed.sys.global.v2.b32 [%0], {%1,%2};" ::"l"(p), "r"(__float_as_uint(v)), "r"(seq)
               : "memory");
}
__device__ __forceinline__ void ld_relaxed_sys_pair(const void* p, float& v, unsigned& seq) {
  unsigned a, b;
  asm volatile("ld.relaxed.sys.global.v2.b32 {%0,%1}, [%2];" : "=r"(a), "=r"(b) : "l"(p) : "memory");
  v = __uint_as_float(a);
  seq = b;
}

// ---- all three helpers are called by ONE full warp (all 32 lanes converged) of ONE CTA per
// kernel; kernels of a rank are stream-ordered, so the header needs no atomics.

// push my pair of `step` into slot[rank] of every rank's buffer
__device__ __forceinline__ void peer_push_warp(const PeerDev& P, unsigned step, float my_sum,
                                               float my_weight) {
  const int lane = threadIdx.x & 31;
  const unsigned par = step & 1u;
  const size_t my_slot = kPeerHeaderBytes + ((size_t)par * kMaxPeers + P.rank) * kPeerSlotBytes;
  for (int r = lane; r < P.world; r += 32) {
    unsigned char* s = P.bufs[r] + my_slot;
    st_relaxed_sys_pair(s, my_sum, step);
    st_relaxed_sys_pair(s + 8, my_weight, step);
  }
}

// wait for every rank's word of `step` in MY buffer, sum in rank order (the same order on every
// rank: bit-identical results), write out[0] = sum_r sum / sum_r weight, out[1] = sum_r weight
__device__ __forceinline__ void peer_consume_warp(const PeerDev& P, unsigned step,
                                                  float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  unsigned char* mine = P.bufs[P.rank];
  unsigned* hdr = reinterpret_cast<unsigned*>(mine);
  const unsigned par = step & 1u;
  const unsigned long long limit = P.timeout_ns ? P.timeout_ns : kPeerDefaultTimeoutNs;
  double tot_s = 0.0, tot_w = 0.0;
  bool ok = true;
  const unsigned long long t0 = globaltimer_ns();
  for (int r0 = 0; r0 < P.world; r0 += 32) {
    const int r = r0 + lane;
    float2 v = make_float2(0.f, 0.f);
    if (r < P.world) {
      const unsigned char* s = mine + kPeerHeaderBytes + ((size_t)par * kMaxPeers + r) * kPeerSlotBytes;
      unsigned q0 = 0, q1 = 0;
      for (;;) {  // both words must carry THIS step (each is atomic and validates itself)
        ld_relaxed_sys_pair(s, v.x, q0);
        ld_relaxed_sys_pair(s + 8, v.y, q1);
        if (q0 == step && q1 == step) break;
        if (globaltimer_ns() - t0 > limit) { ok = false; break; }
        __nanosleep(100);
      }
    }
    ok = __all_sync(0xffffffffu, ok);
    for (int j = 0; j < 32 && r0 + j < P.world; ++j) {  // fixed order: identical on all ranks
      tot_s += (double)__shfl_sync(0xffffffffu, v.x, j);
      tot_w += (double)__shfl_sync(0xffffffffu, v.y, j);
    }
  }
  if (lane == 0) {
    if (!ok) {
      hdr[1] = 1u;  // sticky: b200ctc_peer_error()
      out[0] = __int_as_float(0x7fc00000);
      out[1] = 0.f;
    } else {
      out[0] = tot_w > 0.0 ? (float)(tot_s / tot_w) : 0.f;
      out[1] = (float)tot_w;
    }
    hdr[2] = step;
  }
  __syncwarp();
}

// one exchange call: consume the previous step if that is still owed, push this step, and -
// unless deferred - consume it right away (out then holds THIS step's result on return)
__device__ __forceinline__ void peer_exchange_warp(const PeerDev& P, float my_sum, float my_weight,
                                                   float* __restrict__ out) {
  unsigned* hdr = reinterpret_cast<unsigned*>(P.bufs[P.rank]);
  unsigned step = 0, consumed = 0;
  if ((threadIdx.x & 31) == 0) { step = hdr[0] + 1u; consumed = hdr[2]; }
  step = __shfl_sync(0xffffffffu, step, 0);
  consumed = __shfl_sync(0xffffffffu, consumed, 0);
  if (consumed + 1u < step) peer_consume_warp(P, step - 1u, out);
  peer_push_warp(P, step, my_sum, my_weight);
  if ((threadIdx.x & 31) == 0) hdr[0] = step;
  __syncwarp();
  if (!P.deferred) peer_consume_warp(P, step, out);
}

// consume the latest pushed step if that has not happened yet (b200ctc_peer_wait)
__device__ __forceinline__ void peer_wait_warp(const PeerDev& P, float* __restrict__ out) {
  unsigned* hdr = reinterpret_cast<unsigned*>(P.bufs[P.rank]);
  unsigned step = 0, consumed = 0;
  if ((threadIdx.x & 31) == 0) { step = hdr[0]; consumed = hdr[2]; }
  step = __shfl_sync(0xffffffffu, step, 0);
  consumed = __shfl_sync(0xffffffffu, consumed, 0);
  if (consumed < step) peer_consume_warp(P, step, out);
}

}  // namespace b200ctc

// opaque host handle (include/b200ctc.h: b200ctc_peer_t)
struct b200ctc_peer {
  int world = 0, rank = 0, dev = 0;
  bool connected = false;
  int deferred = 0;                   // b200ctc_peer_configure
  unsigned long long timeout_ns = 0;  // 0 = default (30 s)
  unsigned char* local = nullptr;
  unsigned char* mapped[b200ctc::kMaxPeers] = {};
  unsigned char** d_bufs = nullptr;
};
