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
//   header {u32 step, u32 error}, then slots[2 parities][world] x {f32 sum, f32 weight,
//   u32 seq, u32 pad}.  Parity double-buffering is enough: a rank can only reach step
//   s+2 after it has seen every peer's push of step s+1, which a peer issues after it
//   has consumed step s (stream order).
#pragma once
#include "common.cuh"

namespace b200ctc {

constexpr int kMaxPeers = 64;
constexpr int kPeerSlotBytes = 16;
constexpr int kPeerHeaderBytes = 64;
constexpr size_t kPeerBufBytes = kPeerHeaderBytes + 2 * kMaxPeers * kPeerSlotBytes;
constexpr unsigned long long kPeerTimeoutNs = 5ull * 1000 * 1000 * 1000;  // then NaN + error flag

struct PeerDev {  // passed by value to kernels
  unsigned char* const* bufs;  // device array [world]: every rank's buffer (bufs[rank] = mine)
  int world, rank;
};

__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ void st_release_sys_u32(unsigned* p, unsigned v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned ld_acquire_sys_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Called by ONE full warp (all 32 lanes converged).  out[0] = sum_r sum / sum_r weight,
// out[1] = sum_r weight on every rank.
__device__ __forceinline__ void peer_exchange_warp(const PeerDev& P, float my_sum, float my_weight,
                                                   float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  unsigned char* mine = P.bufs[P.rank];
  unsigned* hdr = reinterpret_cast<unsigned*>(mine);
  unsigned step = 0;
  if (lane == 0) {
    step = hdr[0] + 1u;  // only this kernel touches it; calls are stream-ordered
    hdr[0] = step;
  }
  step = __shfl_sync(0xffffffffu, step, 0);
  const unsigned par = step & 1u;
  const size_t my_slot = kPeerHeaderBytes + ((size_t)par * kMaxPeers + P.rank) * kPeerSlotBytes;
  for (int r = lane; r < P.world; r += 32) {  // push: my pair into slot[rank] of rank r
    unsigned char* s = P.bufs[r] + my_slot;
    asm volatile("st.global.v2.f32 [%0], {%1,%2};" ::"l"(s), "f"(my_sum), "f"(my_weight) : "memory");
    st_release_sys_u32(reinterpret_cast<unsigned*>(s + 8), step);
  }
  // pull: wait for every rank's word of this step in MY buffer, then sum in rank order
  double tot_s = 0.0, tot_w = 0.0;
  bool ok = true;
  const unsigned long long t0 = globaltimer_ns();
  for (int r0 = 0; r0 < P.world; r0 += 32) {
    const int r = r0 + lane;
    float2 v = make_float2(0.f, 0.f);
    if (r < P.world) {
      const unsigned char* s = mine + kPeerHeaderBytes + ((size_t)par * kMaxPeers + r) * kPeerSlotBytes;
      while (ld_acquire_sys_u32(reinterpret_cast<const unsigned*>(s + 8)) != step) {
        if (globaltimer_ns() - t0 > kPeerTimeoutNs) { ok = false; break; }
        __nanosleep(100);
      }
      asm volatile("ld.volatile.global.v2.f32 {%0,%1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(s) : "memory");
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
  }
}

}  // namespace b200ctc

// opaque host handle (include/b200ctc.h: b200ctc_peer_t)
struct b200ctc_peer {
  int world = 0, rank = 0, dev = 0;
  bool connected = false;
  unsigned char* local = nullptr;
  unsigned char* mapped[b200ctc::kMaxPeers] = {};
  unsigned char** d_bufs = nullptr;
};
