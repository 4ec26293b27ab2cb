// Gradient exchange over NVLink peer memory: device-side helpers shared by the reducing kernels (push) and the norm
// kernel (wait + sum).  Protocol and buffer layout: include/flyb200.h, FlybPeerGroup.
#pragma once
#include "flyb_common.cuh"

namespace flyb {

struct PeerView {   // FlybPeerGroup resolved into typed pointers (kernel argument)
  int rank, world, total_pad;
  float* slots[FLYB_MAX_PEERS];        // slots[r]: [2][world][total_pad] in rank r's buffer
  uint32_t* flags[FLYB_MAX_PEERS];     // flags[r]: [FLYB_MAX_PEERS] in rank r's buffer
  uint32_t* epoch;                     // own: exchanges completed by this rank
  uint32_t* error;                     // own: sticky error word (0 = fine, 1 + r = gave up waiting for rank r)
  uint32_t* push_count;                // own: pushing CTAs that have finished (fence mode "last")
  int fence_mode;                      // 0: one system fence per pushing CTA; 1: one by the LAST pushing CTA; 2: none
  unsigned long long timeout_ns;       // how long peer_wait_all spins before it gives up
  unsigned long long* stamp;           // own: 8 x %globaltimer marks of the last exchange (flyb_peer_stamps)
};

inline size_t peer_slots_bytes(int total_pad, int world) { return (size_t)2 * world * total_pad * sizeof(float); }
constexpr size_t PEER_CTRL_BYTES = 256;   // flags[8], epoch, error, (pad), 8 x u64 stamps at byte 64
unsigned long long peer_timeout_ns(const FlybPeerGroup& g);   // peer.cu
int peer_fence_mode();                                        // peer.cu (FLYB_PEER_FENCE = cta | last | none)

inline PeerView make_peer_view(const FlybPeerGroup& g) {
  PeerView v{};
  v.rank = g.rank; v.world = g.world; v.total_pad = g.total_pad;
  const size_t sb = peer_slots_bytes(g.total_pad, g.world);
  for (int r = 0; r < g.world; ++r) {
    v.slots[r] = (float*)g.base[r];
    v.flags[r] = (uint32_t*)((char*)g.base[r] + sb);
  }
  v.epoch = v.flags[g.rank] + FLYB_MAX_PEERS;
  v.error = v.epoch + 1;
  v.push_count = v.epoch + 2;
  v.fence_mode = peer_fence_mode();
  v.timeout_ns = peer_timeout_ns(g);
  v.stamp = (unsigned long long*)(v.flags[g.rank] + 16);
  return v;
}

FLYB_D unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
FLYB_D uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
FLYB_D void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
FLYB_D float ld_relaxed_sys(const float* p) {
  float v;
  asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(v) : "l"(p) : "memory");
  return v;
}

// Entry i of this rank's gradient -> slot [parity][rank] of every rank (own one included): posted NVLink stores.
FLYB_D void peer_push(const PeerView& pv, uint32_t parity, int i, float val) {
  const size_t at = ((size_t)parity * pv.world + pv.rank) * pv.total_pad + i;
#pragma unroll
  for (int r = 0; r < FLYB_MAX_PEERS; ++r)
    if (r < pv.world) pv.slots[r][at] = val;
}
// Four consecutive entries (i a multiple of 4) as ONE 128-bit store per rank: a warp's push is 512 contiguous bytes.
FLYB_D void peer_push4(const PeerView& pv, uint32_t parity, int i, const float4& val) {
  const size_t at = ((size_t)parity * pv.world + pv.rank) * pv.total_pad + i;   // total_pad % 64 == 0: 16-byte aligned
#pragma unroll
  for (int r = 0; r < FLYB_MAX_PEERS; ++r)
    if (r < pv.world) *reinterpret_cast<float4*>(pv.slots[r] + at) = val;
}
// Called by EVERY thread of a CTA after its last peer_push.  The pushed data must be ordered before the flag the NEXT
// kernel raises, and not by the kernel boundary alone (round-1 review).  Three forms, FLYB_PEER_FENCE:
//   last (default)  CTA barrier + gpu-scope fence + counter; the LAST pushing CTA issues the one system-scope fence,
//                   cumulative over every CTA it observed through the counter (142.9 us per optimiser step at 2 GPUs);
//   cta             CTA barrier + one system-scope fence per pushing CTA (145.8 us);
//   none            nothing: the kernel boundary and the flag-raising kernel's own system fence — round 1's form (140.8 us).
// (A fence per pushing THREAD, the first form tried, cost 5-7 us more than `cta` at 8 ranks.)
FLYB_D void peer_push_fence(const PeerView& pv) {
  if (pv.fence_mode == 2) return;
  __syncthreads();
  if (threadIdx.x == 0) {
    if (pv.fence_mode == 0) {
      __threadfence_system();
    } else {
      __threadfence();
      if (atomicAdd(pv.push_count, 1u) == gridDim.x - 1u) {
        *pv.push_count = 0u;
        __threadfence_system();
      }
    }
  }
}

// Raise this rank's flag of epoch `want` on every rank.  Called by ONE CTA of the kernel that FOLLOWS the pushing
// kernel in the stream: the pushing threads fenced their stores at system scope and the kernel boundary orders them
// before this release, so the flag cannot overtake the data.
FLYB_D void peer_raise_flags(const PeerView& pv, uint32_t want) {
  if (threadIdx.x < pv.world) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(pv.flags[threadIdx.x] + pv.rank), "r"(want) : "memory");
  }
}

// Wait until every rank has pushed epoch `want` into this rank's buffer.  One polling lane per source rank.  A peer
// that does not show up within `timeout_ns` (FlybPeerGroup.timeout_ms / FLYB_PEER_TIMEOUT_S, default 120 s — a rank
// may be saving, evaluating or loading modules) does NOT kill the context: the lane records a sticky error word in this
// rank's control block (flyb_peer_status reads it; the host raises and the job restarts with FLYB_DP=nccl), and this and
// every later wait return at once, so the stream drains instead of hanging.  Results after that are meaningless.
FLYB_D void peer_wait_all(const PeerView& pv, uint32_t want) {
  if (threadIdx.x < pv.world) {
    const uint32_t* f = pv.flags[pv.rank] + threadIdx.x;
    const unsigned long long t0 = global_ns();
    unsigned spins = 0;
    while ((int32_t)(ld_acquire_sys(f) - want) < 0) {
      __nanosleep(64);
      if ((++spins & 1023u) == 0u) {
        if (*(volatile uint32_t*)pv.error != 0u) break;
        if (global_ns() - t0 > pv.timeout_ns) {
          if (atomicCAS(pv.error, 0u, 1u + threadIdx.x) == 0u)
            printf("flyb200: rank %d gave up waiting for rank %d (epoch %u); peer exchange disabled\n", pv.rank,
                   (int)threadIdx.x, want);
          break;
        }
      }
    }
  }
  __syncthreads();
}

// Sum of entry i over the ranks' slots, in rank order.
FLYB_D float peer_sum(const PeerView& pv, uint32_t parity, int i) {
  const float* s = pv.slots[pv.rank] + (size_t)parity * pv.world * pv.total_pad + i;
  float acc = ld_relaxed_sys(s);
  for (int r = 1; r < pv.world; ++r) acc += ld_relaxed_sys(s + (size_t)r * pv.total_pad);
  return acc;
}

int peer_validate(const FlybPeerGroup* g, int total);

}  // namespace flyb
