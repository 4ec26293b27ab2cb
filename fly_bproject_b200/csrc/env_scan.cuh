// Reset-id compaction shared by the env kernels: every group of 64 envs publishes a 64-bit mask of the envs whose
// reset flag is now set; whoever finishes LAST (a device counter tells) turns the bitmap into the ascending id list
// (== reset_buf.nonzero(), reference fly.py:447) and the count.  Nobody ever waits for anybody else.
#pragma once
#include "flyb_common.cuh"

namespace flyb {

constexpr int SCAN_ENVS = 64;   // envs per bitmap word

// cross-CTA scratch: `done` counts finished publishers (re-armed by the last one), status[b] = mask of env group b;
// behind the masks: one tagged count per 128-env tile (the fused env step's distributed compaction, below)
struct ScanScratch {
  unsigned int ticket;   // launches of the fused env step that have completed their compaction (tag of the tile counts)
  unsigned int done;
  unsigned long long status[1];
};
__host__ __device__ inline size_t scan_words(int num_envs) { return (size_t)(num_envs + SCAN_ENVS - 1) / SCAN_ENVS + 1; }
__host__ __device__ inline size_t scan_tiles(int num_envs) { return (size_t)(num_envs + 127) / 128 + 1; }
__host__ __device__ inline size_t scan_scratch_bytes(int num_envs) {
  return 16 + scan_words(num_envs) * sizeof(unsigned long long) + ((scan_tiles(num_envs) + 1) / 2) * 8;
}
inline unsigned* scan_tile_counts(ScanScratch* s, int num_envs) {
  return reinterpret_cast<unsigned*>(s->status + scan_words(num_envs));
}

FLYB_D unsigned long long ld_status(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
FLYB_D void st_status(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

FLYB_D unsigned ld_relaxed_u32(const unsigned* p) {
  unsigned v;
  asm volatile("ld.relaxed.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
FLYB_D void st_relaxed_u32(unsigned* p, unsigned v) {
  asm volatile("st.relaxed.gpu.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// Distributed compaction for a grid whose tiles are all in flight together (the fused env step: one 128-env tile per
// SM).  One warp per tile: publish the tile's count as a word tagged with this launch's number, add up the counts of the
// tiles in front (spinning on the tag: every tile publishes before it waits, so nothing can dead-lock as long as tiles
// are taken in ascending order), and write this tile's own ids — ascending, at the right place.  The id list of N envs
// is thus written by N / 128 warps at once, not by the last warp to finish (which took ~20 us for 8000 ids and made its
// CTA the kernel's straggler).  m[k] bit l = env tile * 128 + 32 k + l.  The tile that holds the last envs also stores
// the count.  `tag` = ScanScratch::ticket + 1 as read at kernel entry, identical in every CTA of a launch.
FLYB_D void compact_tile_warp(unsigned* __restrict__ tile_cnt, unsigned tag, long long tile, long long ntiles,
                              const unsigned (&m)[4], int* __restrict__ reset_ids, int* __restrict__ reset_count) {
  const int lane = threadIdx.x & 31;
  const int c0 = __popc(m[0]), c1 = __popc(m[1]), c2 = __popc(m[2]), c3 = __popc(m[3]);
  const int cnt = c0 + c1 + c2 + c3;
  const unsigned t24 = tag & 0xffffffu;
  if (lane == 0) st_relaxed_u32(tile_cnt + tile, (t24 << 8) | (unsigned)cnt);
  int off = 0;
  for (long long u = lane; u < tile; u += 32) {
    unsigned v = ld_relaxed_u32(tile_cnt + u);
    while ((v >> 8) != t24) {
      __nanosleep(40);
      v = ld_relaxed_u32(tile_cnt + u);
    }
    off += (int)(v & 0xffu);
  }
#pragma unroll
  for (int o2 = 16; o2 > 0; o2 >>= 1) off += __shfl_xor_sync(0xffffffffu, off, o2);
  const unsigned below = (1u << lane) - 1u;
  const int env0 = (int)tile * 128 + lane;
  if ((m[0] >> lane) & 1u) reset_ids[off + __popc(m[0] & below)] = env0;
  if ((m[1] >> lane) & 1u) reset_ids[off + c0 + __popc(m[1] & below)] = env0 + 32;
  if ((m[2] >> lane) & 1u) reset_ids[off + c0 + c1 + __popc(m[2] & below)] = env0 + 64;
  if ((m[3] >> lane) & 1u) reset_ids[off + c0 + c1 + c2 + __popc(m[3] & below)] = env0 + 96;
  if (tile == ntiles - 1 && lane == 0) *reset_count = off + cnt;
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
