// Reset-id compaction shared by the env kernels: every group of 64 envs publishes a 64-bit mask of the envs whose
// reset flag is now set; whoever finishes LAST (a device counter tells) turns the bitmap into the ascending id list
// (== reset_buf.nonzero(), reference fly.py:447) and the count.  Nobody ever waits for anybody else.
#pragma once
#include "flyb_common.cuh"

namespace flyb {

constexpr int SCAN_ENVS = 64;   // envs per bitmap word

// cross-CTA scratch: `done` counts finished publishers (re-armed by the last one), status[b] = mask of env group b
struct ScanScratch {
  unsigned int ticket;   // unused (kept so that the scratch size / zeroing contract of ABI v2 stands)
  unsigned int done;
  unsigned long long status[1];
};

FLYB_D unsigned long long ld_status(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
FLYB_D void st_status(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// One full warp turns `nwords` masks into the id list.  Words are read in batches of 32 (lane l takes word 32 i + l:
// one coalesced 256-byte load per batch), eight batches in flight at a time — the masks come from L2, and a lane that
// walked its words one dependent load after the other would spend ~0.6 us on each.  Per batch: popcount, warp prefix sum,
// every lane writes the ids of its word.  Also stores the count and re-arms the counter.
FLYB_D void scan_masks_warp(ScanScratch* scratch, int nwords, int* __restrict__ reset_ids, int* __restrict__ reset_count) {
  const int lane = threadIdx.x & 31;
  const unsigned long long* bitmap = scratch->status;
  constexpr int NB = 8;
  int base = 0;
  for (int w0 = 0; w0 < nwords; w0 += 32 * NB) {
    unsigned long long wd[NB];
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      const int idx = w0 + 32 * i + lane;
      wd[i] = idx < nwords ? ld_status(bitmap + idx) : 0ull;
    }
#pragma unroll
    for (int i = 0; i < NB; ++i) {
      const int cnt = __popcll(wd[i]);
      int incl = cnt;
#pragma unroll
      for (int o2 = 1; o2 < 32; o2 <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o2);
        if (lane >= o2) incl += t;
      }
      const int total = __shfl_sync(0xffffffffu, incl, 31);
      int pos = base + incl - cnt;
      const int env0 = (w0 + 32 * i + lane) * SCAN_ENVS;
      unsigned long long word = wd[i];
      while (word != 0ull) {
        const int b = __ffsll((long long)word) - 1;
        reset_ids[pos++] = env0 + b;
        word &= word - 1ull;
      }
      base += total;
    }
  }
  if (lane == 0) {
    *reset_count = base;
    scratch->done = 0u;   // re-armed for the next launch
  }
}

}  // namespace flyb
