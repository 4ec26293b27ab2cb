// Exchange buffers of the data-parallel gradient sum: allocation and CUDA IPC mapping (include/flyb200.h,
// FlybPeerGroup).  The kernels that use them are in ppo_kernels*.cu (peer.cuh).
#include <stdlib.h>
#include <string.h>

#include "peer.cuh"

namespace flyb {

unsigned long long peer_timeout_ns(const FlybPeerGroup& g) {
  if (g.timeout_ms > 0) return (unsigned long long)g.timeout_ms * 1000000ull;
  static const double env_s = [] {
    const char* e = getenv("FLYB_PEER_TIMEOUT_S");
    const double v = e ? atof(e) : 0.0;
    return v > 0.0 ? v : 120.0;
  }();
  return (unsigned long long)(env_s * 1e9);
}

int peer_fence_mode() {
  const char* e = getenv("FLYB_PEER_FENCE");   // read per call: one process can A/B the forms
  if (e && e[0] == 'c') return 0;
  if (e && e[0] == 'n') return 2;
  return 1;   // default: last (measured at 2 GPUs, one box: 142.9 us per optimiser step; cta 145.8, none 140.8)
}

int peer_validate(const FlybPeerGroup* g, int total) {
  FLYB_REQUIRE(g->world >= 1 && g->world <= FLYB_MAX_PEERS, "FlybPeerGroup: world %d out of range", g->world);
  FLYB_REQUIRE(g->rank >= 0 && g->rank < g->world, "FlybPeerGroup: rank %d out of range", g->rank);
  FLYB_REQUIRE(g->total_pad >= total && g->total_pad % 64 == 0, "FlybPeerGroup: total_pad %d does not fit %d entries",
               g->total_pad, total);
  for (int r = 0; r < g->world; ++r)
    FLYB_REQUIRE(g->base[r] && (reinterpret_cast<uintptr_t>(g->base[r]) & 255u) == 0,
                 "FlybPeerGroup: base[%d] NULL or not 256-byte aligned", r);
  return FLYB_OK;
}

}  // namespace flyb

using namespace flyb;

extern "C" int32_t flyb_peer_total_pad(int32_t total) { return (total + 63) / 64 * 64; }

extern "C" size_t flyb_peer_bytes(int32_t total, int32_t world) {
  if (total <= 0 || world < 1 || world > FLYB_MAX_PEERS) return 0;
  const int tp = flyb_peer_total_pad(total);
  return peer_slots_bytes(tp, world) + PEER_CTRL_BYTES + peer_ll_bytes(tp, world);
}

extern "C" int flyb_peer_alloc(size_t bytes, void** ptr_out, void* handle_out) {
  FLYB_REQUIRE(ptr_out && handle_out && bytes > 0, "flyb_peer_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* p = nullptr;
  int rc = check_cuda(cudaMalloc(&p, bytes), "cudaMalloc(peer buffer)");
  if (rc != FLYB_OK) return rc;
  if ((rc = check_cuda(cudaMemset(p, 0, bytes), "cudaMemset(peer buffer)")) != FLYB_OK) { cudaFree(p); return rc; }
  if ((rc = check_cuda(cudaDeviceSynchronize(), "cudaDeviceSynchronize")) != FLYB_OK) { cudaFree(p); return rc; }
  cudaIpcMemHandle_t h;
  if ((rc = check_cuda(cudaIpcGetMemHandle(&h, p), "cudaIpcGetMemHandle")) != FLYB_OK) { cudaFree(p); return rc; }
  memcpy(handle_out, &h, sizeof(h));
  *ptr_out = p;
  return FLYB_OK;
}

extern "C" int flyb_peer_open(const void* handle, void** ptr_out) {
  FLYB_REQUIRE(handle && ptr_out, "flyb_peer_open: NULL argument");
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  return check_cuda(cudaIpcOpenMemHandle(ptr_out, h, cudaIpcMemLazyEnablePeerAccess), "cudaIpcOpenMemHandle");
}

extern "C" int flyb_peer_close(void* ptr) {
  FLYB_REQUIRE(ptr, "flyb_peer_close: NULL");
  return check_cuda(cudaIpcCloseMemHandle(ptr), "cudaIpcCloseMemHandle");
}

extern "C" int flyb_peer_free(void* ptr) {
  FLYB_REQUIRE(ptr, "flyb_peer_free: NULL");
  return check_cuda(cudaFree(ptr), "cudaFree(peer buffer)");
}

extern "C" int flyb_peer_stamps(const FlybPeerGroup* g, uint64_t* out8) {
  FLYB_REQUIRE(g && out8, "flyb_peer_stamps: NULL argument");
  int rc = peer_validate(g, 1);
  if (rc != FLYB_OK) return rc;
  const PeerView v = make_peer_view(*g);
  if ((rc = check_cuda(cudaDeviceSynchronize(), "cudaDeviceSynchronize")) != FLYB_OK) return rc;
  return check_cuda(cudaMemcpy(out8, v.stamp, 8 * sizeof(uint64_t), cudaMemcpyDeviceToHost), "cudaMemcpy(stamps)");
}

extern "C" int flyb_peer_status(const FlybPeerGroup* g, int32_t* status_out) {
  FLYB_REQUIRE(g && status_out, "flyb_peer_status: NULL argument");
  int rc = peer_validate(g, 1);
  if (rc != FLYB_OK) return rc;
  const PeerView v = make_peer_view(*g);
  uint32_t word = 0;
  if ((rc = check_cuda(cudaMemcpy(&word, v.error, sizeof(word), cudaMemcpyDeviceToHost), "cudaMemcpy(peer status)")) !=
      FLYB_OK)
    return rc;
  *status_out = (int32_t)word;
  return FLYB_OK;
}
