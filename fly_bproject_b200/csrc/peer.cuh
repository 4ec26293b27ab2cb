// Gradient exchange over NVLink peer memory: device-side helpers shared by the reducing kernels (push) and the norm
// kernel (wait + sum).  Protocol and buffer layout: include/flyb200.h, FlybPeerGroup.
#pragma once
#include "flyb_common.cuh"

namespace flyb {

struct PeerView {   // FlybPeerGroup resolved into typed pointers (kernel argument)
  int rank, world, total_pad;
  float* slots[FLYB_MAX_PEERS];        // slots[r]: [2][world][total_pad] in rank r's buffer
  uint32_t* flags[FLYB_MAX_PEERS];     // flags[r]: [FLYB_MAX_PEERS] in rank r's buffer
  unsigned long long* ll[FLYB_MAX_PEERS];   // ll[r]: [2][world][total_pad] {epoch << 32 | value bits} words in rank r's buffer
  uint32_t* epoch;                     // own: exchanges completed by this rank
  uint32_t* error;                     // own: sticky error word (0 = fine, 1 + r = gave up waiting for rank r)
  uint32_t* push_count;                // own: pushing CTAs that have finished (fence mode "last")
  int fence_mode;                      // 0: one system fence per pushing CTA; 1: one by the LAST pushing CTA; 2: none
  unsigned long long timeout_ns;       // how long peer_wait_all spins before it gives up
  unsigned long long* stamp;           // own: 8 x %globaltimer marks of the last exchange (flyb_peer_stamps)
};

inline size_t peer_slots_bytes(int total_pad, int world) { return (size_t)2 * world * total_pad * sizeof(float); }
constexpr size_t PEER_CTRL_BYTES = 256;   // flags[8], epoch, error, (pad), 8 x u64 stamps at byte 64
// second region, behind the control block: the self-validating words of the fused exchange (flyb_ppo_step_dp)
inline size_t peer_ll_bytes(int total_pad, int world) { return (size_t)2 * world * total_pad * sizeof(unsigned long long); }
unsigned long long peer_timeout_ns(const FlybPeerGroup& g);   // peer.cu
int peer_fence_mode();                                        // peer.cu (FLYB_PEER_FENCE = cta | last | none)

inline PeerView make_peer_view(const FlybPeerGroup& g) {
  PeerView v{};
  v.rank = g.rank; v.world = g.world; v.total_pad = g.total_pad;
  const size_t sb = peer_slots_bytes(g.total_pad, g.world);
  for (int r = 0; r < g.world; ++r) {
    v.slots[r] = (float*)g.base[r];
    v.flags[r] = (uint32_t*)((char*)g.base[r] + sb);
    v.ll[r] = (unsigned long long*)((char*)g.base[r] + sb + PEER_CTRL_BYTES);
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
FLYB_D void peer_wait_flag(const PeerView& pv, const uint32_t* f, uint32_t want, int src) {
  const unsigned long long t0 = global_ns();
  unsigned spins = 0;
  while ((int32_t)(ld_acquire_sys(f) - want) < 0) {
    __nanosleep(64);
    if ((++spins & 1023u) == 0u) {
      if (*(volatile uint32_t*)pv.error != 0u) break;
      if (global_ns() - t0 > pv.timeout_ns) {
        if (atomicCAS(pv.error, 0u, 1u + (uint32_t)src) == 0u)
          printf("flyb200: rank %d gave up waiting for rank %d (epoch %u); peer exchange disabled\n", pv.rank, src, want);
        break;
      }
    }
  }
}
FLYB_D void peer_wait_all(const PeerView& pv, uint32_t want) {
  if (threadIdx.x < pv.world) peer_wait_flag(pv, pv.flags[pv.rank] + threadIdx.x, want, (int)threadIdx.x);
  __syncthreads();
}

// ---- fused exchange (flyb_ppo_step_dp): no flags, no fences.  Every gradient entry travels as ONE 64-bit word
// {epoch + 1 << 32 | float bits}: 64-bit accesses are single-copy atomic, so a word that shows the awaited epoch carries
// its value (the scheme of NCCL's LL protocol).  The thread that has reduced a quad of entries stores it into its rank's
// slot on every rank and then polls the same quad of every rank's slot in its OWN buffer, adding them in rank order:
// the latency is one NVLink crossing, where push + system fence + flag + poll pays three.  Nothing here waits for
// another thread of this grid, and what a thread waits for is stored by the peers' threads of the same index before
// THEY wait: no dead-lock AS LONG AS the whole reducing grid is resident on every rank (every CTA then pushes before
// it waits, in whatever order CTAs are dispatched).  A grid larger than the GPU holds could fill the machine with CTAs
// that wait for peers' CTAs not yet scheduled: flyb_ppo_step_dp checks residency (tc_ppo_exchange_resident) and runs
// the split exchange otherwise; SMs taken by other clients of the GPU are beyond that check.  Slots are double-buffered by epoch parity and a rank's next
// reducing kernel starts only after this one has ended on that rank (stream order), so a rank can be at most one
// exchange ahead and never overwrites words a peer still has to read.  Twice the bytes of the float slots.
FLYB_D unsigned long long ll_pack(float v, uint32_t tag) {
  return ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(v);
}
FLYB_D void peer_ll_push4(const PeerView& pv, uint32_t parity, uint32_t tag, int i, const float (&g)[4]) {
  const size_t at = ((size_t)parity * pv.world + pv.rank) * pv.total_pad + i;   // i % 4 == 0: 32-byte aligned
  const unsigned long long w0 = ll_pack(g[0], tag), w1 = ll_pack(g[1], tag), w2 = ll_pack(g[2], tag), w3 = ll_pack(g[3], tag);
#pragma unroll
  for (int r = 0; r < FLYB_MAX_PEERS; ++r)
    if (r < pv.world) {
      unsigned long long* d = pv.ll[r] + at;
      asm volatile("st.relaxed.sys.global.v2.b64 [%0], {%1, %2};" ::"l"(d), "l"(w0), "l"(w1) : "memory");
      asm volatile("st.relaxed.sys.global.v2.b64 [%0], {%1, %2};" ::"l"(d + 2), "l"(w2), "l"(w3) : "memory");
    }
}
// Sum over the ranks' slots of the quad at entry i, in rank order; spins until every word shows `tag`.  A peer that does
// not show up within the timeout sets the sticky error word (see peer_wait_flag); the sum is then meaningless.
FLYB_D float4 peer_ll_sum4(const PeerView& pv, uint32_t parity, uint32_t tag, int i) {
  const unsigned long long* base = pv.ll[pv.rank] + (size_t)parity * pv.world * pv.total_pad + i;
  unsigned long long w[FLYB_MAX_PEERS][4];
  uint32_t pending = (1u << pv.world) - 1u;
  unsigned long long t0 = 0;
  unsigned spins = 0;
  while (pending) {
#pragma unroll
    for (int r = 0; r < FLYB_MAX_PEERS; ++r)
      if (pending & (1u << r)) {
        const unsigned long long* s = base + (size_t)r * pv.total_pad;
        asm volatile("ld.relaxed.sys.global.v2.b64 {%0, %1}, [%2];" : "=l"(w[r][0]), "=l"(w[r][1]) : "l"(s) : "memory");
        asm volatile("ld.relaxed.sys.global.v2.b64 {%0, %1}, [%2];" : "=l"(w[r][2]), "=l"(w[r][3]) : "l"(s + 2) : "memory");
      }
#pragma unroll
    for (int r = 0; r < FLYB_MAX_PEERS; ++r)
      if (pending & (1u << r)) {
        const bool ok = (uint32_t)(w[r][0] >> 32) == tag && (uint32_t)(w[r][1] >> 32) == tag &&
                        (uint32_t)(w[r][2] >> 32) == tag && (uint32_t)(w[r][3] >> 32) == tag;
        if (ok) pending &= ~(1u << r);
      }
    if (pending && (++spins & 255u) == 0u) {
      if (t0 == 0) t0 = global_ns();
      if (*(volatile uint32_t*)pv.error != 0u) break;
      if (global_ns() - t0 > pv.timeout_ns) {
        const int src = __ffs((int)pending) - 1;
        if (atomicCAS(pv.error, 0u, 1u + (uint32_t)src) == 0u)
          printf("flyb200: rank %d gave up waiting for rank %d (epoch %u); peer exchange disabled\n", pv.rank, src, tag);
        break;
      }
    }
  }
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
  for (int r = 0; r < FLYB_MAX_PEERS; ++r)
    if (r < pv.world) {
      const float x = __uint_as_float((uint32_t)w[r][0]), y = __uint_as_float((uint32_t)w[r][1]);
      const float z = __uint_as_float((uint32_t)w[r][2]), u = __uint_as_float((uint32_t)w[r][3]);
      if (r == 0) acc = make_float4(x, y, z, u);
      else { acc.x += x; acc.y += y; acc.z += z; acc.w += u; }
    }
  return acc;
}

// Sum of entry i over the ranks' float slots (split exchange), in rank order.
FLYB_D float peer_sum(const PeerView& pv, uint32_t parity, int i) {
  const float* s = pv.slots[pv.rank] + (size_t)parity * pv.world * pv.total_pad + i;
  float acc = ld_relaxed_sys(s);
  for (int r = 1; r < pv.world; ++r) acc += ld_relaxed_sys(s + (size_t)r * pv.total_pad);
  return acc;
}

int peer_validate(const FlybPeerGroup* g, int total);

}  // namespace flyb
