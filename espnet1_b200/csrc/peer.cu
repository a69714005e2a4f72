// Host side of the peer-memory loss exchange (peer.cuh): buffer creation, cudaIpc
// mapping of the peers' buffers, the stand-alone exchange kernel.
#include <cstring>

#include "peer.cuh"

namespace b200ctc {

__global__ void __launch_bounds__(32) peer_exchange_kernel(PeerDev P, const float* __restrict__ pair,
                                                           float* __restrict__ out) {
  peer_exchange_warp(P, pair[0], pair[1], out);
}

__global__ void __launch_bounds__(32) peer_wait_kernel(PeerDev P, float* __restrict__ out) {
  peer_wait_warp(P, out);
}

}  // namespace b200ctc

using namespace b200ctc;

extern "C" int b200ctc_peer_configure(b200ctc_peer* p, int deferred, int timeout_ms) {
  B2_REQUIRE(p, "peer_configure: null handle");
  B2_REQUIRE(timeout_ms >= 0, "peer_configure: timeout_ms must be >= 0 (0 = default)");
  p->deferred = deferred ? 1 : 0;
  p->timeout_ns = (unsigned long long)timeout_ms * 1000000ull;
  return B200CTC_OK;
}

extern "C" int b200ctc_peer_wait(b200ctc_peer* p, float* global_out, b200ctc_stream_t stream) {
  B2_REQUIRE(p && p->connected, "peer_wait: handle is not connected");
  B2_REQUIRE(global_out, "peer_wait: null pointer");
  PeerDev P{p->d_bufs, p->world, p->rank, p->deferred, p->timeout_ns};
  peer_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(P, global_out);
  count_launch();
  return check_launch("peer wait");
}

extern "C" int b200ctc_peer_create(int world, int rank, b200ctc_peer** out,
                                   unsigned char* handle_out) {
  B2_REQUIRE(out && handle_out, "peer_create: null pointer");
  B2_REQUIRE(world >= 1 && world <= kMaxPeers && rank >= 0 && rank < world,
             "peer_create: bad world/rank %d/%d (max %d ranks)", rank, world, kMaxPeers);
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "handle size is part of the ABI");
  b200ctc_peer* p = new b200ctc_peer();
  p->world = world;
  p->rank = rank;
  cudaError_t e = cudaGetDevice(&p->dev);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&p->local), kPeerBufBytes);
  if (e == cudaSuccess) e = cudaMemset(p->local, 0, kPeerBufBytes);
  if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&p->d_bufs), sizeof(void*) * kMaxPeers);
  cudaIpcMemHandle_t h;
  memset(&h, 0, sizeof(h));
  if (e == cudaSuccess && world > 1) e = cudaIpcGetMemHandle(&h, p->local);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    set_error("peer_create: %s", cudaGetErrorString(e));
    if (p->local) cudaFree(p->local);
    if (p->d_bufs) cudaFree(p->d_bufs);
    delete p;
    cudaGetLastError();
    return B200CTC_ERR_DEVICE;
  }
  memcpy(handle_out, &h, sizeof(h));
  *out = p;
  return B200CTC_OK;
}

extern "C" int b200ctc_peer_connect(b200ctc_peer* p, const unsigned char* all_handles) {
  B2_REQUIRE(p && (all_handles || p->world == 1), "peer_connect: null pointer");
  B2_REQUIRE(!p->connected, "peer_connect: already connected");
  for (int r = 0; r < p->world; ++r) {
    if (r == p->rank) { p->mapped[r] = p->local; continue; }
    cudaIpcMemHandle_t h;
    memcpy(&h, all_handles + (size_t)r * sizeof(h), sizeof(h));
    void* ptr = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      set_error("peer_connect: cannot map rank %d's buffer: %s", r, cudaGetErrorString(e));
      for (int q = 0; q < r; ++q)
        if (q != p->rank && p->mapped[q]) { cudaIpcCloseMemHandle(p->mapped[q]); p->mapped[q] = nullptr; }
      cudaGetLastError();
      return B200CTC_ERR_DEVICE;
    }
    p->mapped[r] = static_cast<unsigned char*>(ptr);
  }
  const cudaError_t e = cudaMemcpy(p->d_bufs, p->mapped, sizeof(void*) * p->world, cudaMemcpyHostToDevice);
  if (e != cudaSuccess) {
    set_error("peer_connect: %s", cudaGetErrorString(e));
    return B200CTC_ERR_DEVICE;
  }
  p->connected = true;
  return B200CTC_OK;
}

extern "C" int b200ctc_peer_destroy(b200ctc_peer* p) {
  if (!p) return B200CTC_OK;
  cudaDeviceSynchronize();
  for (int r = 0; r < p->world; ++r)
    if (r != p->rank && p->mapped[r]) cudaIpcCloseMemHandle(p->mapped[r]);
  if (p->local) cudaFree(p->local);
  if (p->d_bufs) cudaFree(p->d_bufs);
  delete p;
  cudaGetLastError();
  return B200CTC_OK;
}

extern "C" int b200ctc_peer_error(b200ctc_peer* p) {
  if (!p || !p->local) return 0;
  unsigned hdr[2] = {0, 0};
  if (cudaMemcpy(hdr, p->local, sizeof(hdr), cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return (int)hdr[1];
}

extern "C" int b200ctc_peer_exchange(b200ctc_peer* p, const float* local_pair, float* global_out,
                                     b200ctc_stream_t stream) {
  B2_REQUIRE(p && p->connected, "peer_exchange: handle is not connected");
  B2_REQUIRE(local_pair && global_out, "peer_exchange: null pointer");
  // (stand-alone: always consumes its own step - the caller wants the result now)
  PeerDev P{p->d_bufs, p->world, p->rank, 0, p->timeout_ns};
  peer_exchange_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(P, local_pair, global_out);
  count_launch();
  return check_launch("peer exchange");
}
