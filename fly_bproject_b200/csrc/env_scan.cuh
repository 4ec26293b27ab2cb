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

// One full warp turns `nwords` masks into the id list: each lane owns a contiguous run of words (count, warp scan,
// write).  Also stores the count and re-arms the counter.
FLYB_D void scan_masks_warp(ScanScratch* scratch, int nwords, int* __restrict__ reset_ids, int* __restrict__ reset_count) {
  const int lane = threadIdx.x & 31;
  const unsigned long long* bitmap = scratch->status;
  const int W = (nwords + 31) / 32;
  const int w_begin = min(lane * W, nwords), w_end = min(w_begin + W, nwords);
  int cnt = 0;
  for (int w = w_begin; w < w_end; ++w) cnt += __popcll(ld_status(bitmap + w));
  int incl = cnt;
#pragma unroll
  for (int o2 = 1; o2 < 32; o2 <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o2);
    if (lane >= o2) incl += t;
  }
  const int total = __shfl_sync(0xffffffffu, incl, 31);
  int pos = incl - cnt;
  for (int w = w_begin; w < w_end; ++w) {
    unsigned long long word = ld_status(bitmap + w);
    while (word != 0ull) {
      const int b = __ffsll((long long)word) - 1;
      reset_ids[pos++] = w * SCAN_ENVS + b;
      word &= word - 1ull;
    }
  }
  if (lane == 0) {
    *reset_count = total;
    scratch->done = 0u;   // re-armed for the next launch
  }
}

}  // namespace flyb
