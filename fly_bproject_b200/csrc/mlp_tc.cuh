// Tensor-core forward of the actor-critic MLP (reference ppo.py:10-38, 88-102) on 128-row tiles:
// tcgen05.mma kind::tf32 (3xTF32, fp32 accumulate in tensor memory), weights streamed L2 -> shared memory by
// 1-D bulk TMA copies through a 32-stage mbarrier ring, activations kept in TENSOR MEMORY between layers
// (the A operand of every GEMM is read from TMEM, "TS" mode), so that an observation row never leaves the SM
// between the first layer and the action / value it produces.
//
// One CTA = 10 warps:  warps 0-7  epilogue (TMEM -> registers -> bias + ELU -> hi/lo split -> TMEM),
//                      warp  8    TMA producer (one lane): observation tiles + the packed weight stream,
//                      warp  9    MMA issuer (one lane).
// Epilogue warp w owns TMEM lanes 32*(w%4)..+31 (row r of the tile = lane r) and column half w/4.
//
// Layer schedule of one tile (K0 = 80 padded obs, trunk 256 -> 128, heads 64 -> {A<=32, 1}):
//   the 256-wide first layer is produced in 8 chunks of 32 columns; chunk c is  D1[c&1] = A1 * W1[c]^T  (N = 32),
//   its epilogue writes ELU(D1 + b1) as the K-chunk A2[c&1] of the second layer, and  D2 += A2[c&1] * W2[:, c]^T
//   (N = 128) consumes it.  D1 / A2 are double buffered: the epilogue of chunk c overlaps the MMAs of chunks
//   c-1 (layer 2) and c+1 (layer 1).  Then  A3 = ELU(D2 + b2),  D3 = A3 * W3^T (mean and/or value head, N = 64
//   each),  A4 = ELU(D3pi + b3),  D4 = A4 * W4^T (N = 32); the 64 -> 1 value output is a register dot product.
//
// TMEM columns (512 allocated):   A1 hi|lo [0,160)   D1[2] [160,224)   A2[2] hi|lo [224,352)   D2 [352,480)
//   after the trunk:              A3 hi|lo [0,256)   D3 pi|v [256,384) A4 hi|lo [384,512)      D4 [0,32)
//
// Packed weight stream (flyb_pack_weights): 4 KB stages in consumption order, each k-step (8 k) of an N-row
// operand stored as [hi slab | lo slab], slab = N x 32 bytes in the no-swizzle K-major core-matrix layout
// (tc_common.cuh).  Trunk = 8 x (W1 chunk: 5 stages, W2 K-chunk: 8 stages), then W3pi (16), W3v (16), W4 (4).
#pragma once
#include "tc_common.cuh"

namespace flyb {
namespace tcm {

constexpr int TM = 128;
constexpr int K0 = 80;
constexpr int H1 = 256, H2 = 128, H3 = 64, APAD = 32;
constexpr int CH = 32, NCH = H1 / CH;
constexpr int STAGE = 4096;
constexpr int NST = 32;                        // ring stages (128 KB)
constexpr int EPI_WARPS = 8, EPI_THREADS = EPI_WARPS * 32;
constexpr int PRODUCER_WARP = 8, MMA_WARP = 9;
constexpr int THREADS = 320;

constexpr int W1C_STAGES = (K0 / 8) * 2 * (CH * 32) / STAGE;   // 5
constexpr int W2C_STAGES = (CH / 8) * 2 * (H2 * 32) / STAGE;   // 8
constexpr int TRUNK_STAGES = NCH * (W1C_STAGES + W2C_STAGES);  // 104
constexpr int W3_STAGES = (H2 / 8) * 2 * (H3 * 32) / STAGE;    // 16
constexpr int W4_STAGES = (H3 / 8) * 2 * (APAD * 32) / STAGE;  // 4
constexpr int OFF_W3PI = TRUNK_STAGES, OFF_W3V = OFF_W3PI + W3_STAGES, OFF_W4 = OFF_W3V + W3_STAGES;
constexpr int FWD_STAGES = OFF_W4 + W4_STAGES;                 // 140 stages = 560 KB

// TMEM column map
constexpr uint32_t C_A1H = 0, C_A1L = K0, C_D1 = 160, C_A2 = 224, C_D2 = 352;
constexpr uint32_t C_A3H = 0, C_A3L = H2, C_D3 = 256, C_A4H = 384, C_A4L = 448, C_D4 = 0;

// bias block in shared memory (floats)
constexpr int B_B1 = 0, B_B2 = 256, B_B3PI = 384, B_B3V = 448, B_B4 = 512, B_W4V = 544, B_B4V = 608, B_TOTAL = 640;

constexpr size_t OBS_STAGE_BYTES = (size_t)TM * K0 * 4;   // 40960: room for any O <= 80

struct Smem {
  unsigned char* ring;
  float* obs;
  float* out;
  float* bias;
  float* vpart;
  uint64_t *full, *empty, *bar_a, *bar_d, *obs_full, *obs_empty;
  uint32_t* tmem_slot;
};

__host__ __device__ inline size_t smem_bytes(size_t out_bytes) {
  return 1024 + (size_t)NST * STAGE + OBS_STAGE_BYTES + ((out_bytes + 127) / 128 * 128) + B_TOTAL * 4 + TM * 4 +
         (2 * NST + 6) * 8 + 16;
}

FLYB_D Smem carve(unsigned char* raw, size_t out_bytes) {
  Smem s;
  uintptr_t p = (reinterpret_cast<uintptr_t>(raw) + 1023) & ~(uintptr_t)1023;
  s.ring = reinterpret_cast<unsigned char*>(p);
  p += (size_t)NST * STAGE;
  s.obs = reinterpret_cast<float*>(p);
  p += OBS_STAGE_BYTES;
  s.out = reinterpret_cast<float*>(p);
  p += (out_bytes + 127) / 128 * 128;
  s.bias = reinterpret_cast<float*>(p);
  p += B_TOTAL * 4;
  s.vpart = reinterpret_cast<float*>(p);
  p += TM * 4;
  s.full = reinterpret_cast<uint64_t*>(p);
  s.empty = s.full + NST;
  s.bar_a = s.empty + NST;
  s.bar_d = s.bar_a + 2;
  s.obs_full = s.bar_d + 2;
  s.obs_empty = s.obs_full + 1;
  s.tmem_slot = reinterpret_cast<uint32_t*>(s.obs_empty + 1);
  return s;
}

FLYB_D void epi_bar_sync() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// ---------------------------------------------------------------------------------------------- MMA issuer
struct MmaState {
  uint32_t tb, ring;
  uint64_t *full, *empty;
  uint32_t g;          // stages consumed so far (running over tiles)
  uint32_t acquired;   // stages whose full barrier has been waited for
};

// One GEMM segment: D[128 x N] (+)= A[128 x 8*ksteps] * B^T, A hi/lo in TMEM, B = the next 2*ksteps slabs of the ring.
template <int N>
FLYB_D void mma_segment(MmaState& m, uint32_t a_hi, uint32_t a_lo, uint32_t d, int ksteps, bool accumulate) {
  constexpr uint32_t SLAB = N * 32;
  constexpr uint32_t SPS = STAGE / SLAB;
  const uint32_t idesc = tc::umma_idesc_tf32(N);
  for (int ks = 0; ks < ksteps; ++ks) {
    const uint32_t slab_hi = 2 * ks, slab_lo = 2 * ks + 1;
    const uint32_t st_hi = m.g + slab_hi / SPS, st_lo = m.g + slab_lo / SPS;
    if (m.acquired <= st_lo) {
      while (m.acquired <= st_lo) {
        tc::mbar_wait(&m.full[m.acquired % NST], (m.acquired / NST) & 1u);
        ++m.acquired;
      }
      tc::tc_fence_after();
    }
    const uint32_t addr_hi = m.ring + (st_hi % NST) * STAGE + (slab_hi % SPS) * SLAB;
    const uint32_t addr_lo = m.ring + (st_lo % NST) * STAGE + (slab_lo % SPS) * SLAB;
    const uint64_t b_hi = tc::umma_smem_desc(addr_hi, N * 16, 128);
    const uint64_t b_lo = tc::umma_smem_desc(addr_lo, N * 16, 128);
    tc::umma_tf32_ts(d, a_lo + 8 * ks, b_hi, idesc, accumulate || ks > 0);
    tc::umma_tf32_ts(d, a_hi + 8 * ks, b_lo, idesc, true);
    tc::umma_tf32_ts(d, a_hi + 8 * ks, b_hi, idesc, true);
    if (SPS == 1) {
      tc::umma_commit(&m.empty[st_hi % NST]);
      tc::umma_commit(&m.empty[st_lo % NST]);
    } else if (slab_lo % SPS == SPS - 1) {
      tc::umma_commit(&m.empty[st_lo % NST]);
    }
  }
  m.g += (uint32_t)(2 * ksteps) / SPS;
}

// ---------------------------------------------------------------------------------------------- epilogue helpers
// y = ELU(acc + bias) for 16 consecutive accumulator columns, split to hi/lo and stored as the next A operand
FLYB_D void epi_elu16(uint32_t tb, uint32_t lane_base, uint32_t d_col, uint32_t ah_col, uint32_t al_col,
                      const float* bias, float* keep) {
  uint32_t v[16], hi[16], lo[16];
  tc::tmem_ld16(tc::tmem_addr(tb, lane_base, d_col), v);
  tc::tmem_wait_ld();
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const float y = elu(__uint_as_float(v[j]) + bias[j]);
    if (keep) keep[j] = y;
    tc::split_tf32(y, hi[j], lo[j]);
  }
  tc::tmem_st16(tc::tmem_addr(tb, lane_base, ah_col), hi);
  tc::tmem_st16(tc::tmem_addr(tb, lane_base, al_col), lo);
}

struct FwdArgs {
  const float* obs;          // (rows, O)
  const unsigned char* wpack;
  const float* params;       // flat fp32 parameters (biases and the 64 -> 1 value row are read from here)
  long long rows;
  int O, A;
  size_t out_bytes;
};

// Per-tile consumer of the results; called by the 256 epilogue threads.
//   begin_tile(...)                      after the observation tile is staged in shared memory (obs_s, row-major)
//   pi(row_in_tile, half, mu[32], ...)   both column halves call it with the full mu row (ELU applied)
//   value(row_in_tile, v)                half 0 only
// See policy_kernels_tc.cu for the implementations.

template <bool WANT_PI, bool WANT_V, class Final>
FLYB_D void forward_persistent(const FlybNetLayout& net, const FwdArgs& a, Final& fin, unsigned char* smem_raw) {
  const Smem s = carve(smem_raw, a.out_bytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int O = a.O;
  const long long ntiles = (a.rows + TM - 1) / TM;

  // ---- one-time setup
  if (threadIdx.x == 0) {
    for (int i = 0; i < NST; ++i) {
      tc::mbar_init(&s.full[i], 1);
      tc::mbar_init(&s.empty[i], 1);
    }
    tc::mbar_init(&s.bar_a[0], EPI_WARPS);
    tc::mbar_init(&s.bar_a[1], EPI_WARPS);
    tc::mbar_init(&s.bar_d[0], 1);
    tc::mbar_init(&s.bar_d[1], 1);
    tc::mbar_init(s.obs_full, 1);
    tc::mbar_init(s.obs_empty, EPI_WARPS);
    tc::fence_mbar_init();
  }
  if (warp == MMA_WARP) tc::tmem_alloc(s.tmem_slot, 512);
  {  // biases + value output row -> shared memory
    const float* p = a.params;
    for (int i = threadIdx.x; i < B_TOTAL; i += THREADS) {
      float v = 0.0f;
      if (i < B_B2) v = p[net.off[1] + i];
      else if (i < B_B3PI) v = p[net.off[3] + i - B_B2];
      else if (i < B_B3V) v = p[net.off[5] + i - B_B3PI];
      else if (i < B_B4) v = p[net.off[9] + i - B_B3V];
      else if (i < B_W4V) v = (i - B_B4 < a.A) ? p[net.off[7] + i - B_B4] : 0.0f;
      else if (i < B_B4V) v = p[net.off[10] + i - B_W4V];
      else if (i == B_B4V) v = p[net.off[11]];
      s.bias[i] = v;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tb = *s.tmem_slot;

  if (warp == PRODUCER_WARP) {
    // =========================================================================== TMA producer
    if (lane == 0) {
      uint32_t gs = 0, obs_phase = 0;
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        const long long row0 = tile * TM;
        const int nrows = (int)min((long long)TM, a.rows - row0);
        const uint32_t bytes = (uint32_t)nrows * (uint32_t)O * 4u;
        const float* src = a.obs + (size_t)row0 * O;
        tc::mbar_wait(s.obs_empty, obs_phase ^ 1u);   // the epilogue warps are done with the previous tile's rows
        obs_phase ^= 1u;
        if ((bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
          tc::mbar_arrive_expect_tx(s.obs_full, bytes);
          tc::tma_load_1d(s.obs, src, bytes, s.obs_full);
        }
        auto stream = [&](int first, int count) {
          for (int i = 0; i < count; ++i, ++gs) {
            const uint32_t slot = gs % NST;
            tc::mbar_wait(&s.empty[slot], ((gs / NST) & 1u) ^ 1u);
            tc::mbar_arrive_expect_tx(&s.full[slot], STAGE);
            tc::tma_load_1d(s.ring + (size_t)slot * STAGE, a.wpack + (size_t)(first + i) * STAGE, STAGE, &s.full[slot]);
          }
        };
        stream(0, TRUNK_STAGES);
        if (WANT_PI) stream(OFF_W3PI, W3_STAGES);
        if (WANT_V) stream(OFF_W3V, W3_STAGES);
        if (WANT_PI) stream(OFF_W4, W4_STAGES);
      }
    }
  } else if (warp == MMA_WARP) {
    // =========================================================================== MMA issuer
    if (lane == 0) {
      MmaState m{tb, tc::smem_u32(s.ring), s.full, s.empty, 0u, 0u};
      uint32_t pa[2] = {0u, 0u};
      auto wait_a = [&](int b) {
        tc::mbar_wait(&s.bar_a[b], pa[b]);
        pa[b] ^= 1u;
        tc::tc_fence_after();
      };
      for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
        wait_a(0);   // A1 staged
        // issue order (== packed stream order): W1[0], W1[1], then (W2[c], W1[c+2]) for c = 0..7, so that the
        // epilogue of chunk c overlaps the layer-2 MMAs of chunk c-1 and the layer-1 MMAs of chunk c+1
        mma_segment<CH>(m, tb + C_A1H, tb + C_A1L, tb + C_D1, K0 / 8, false);
        tc::umma_commit(&s.bar_d[0]);
        mma_segment<CH>(m, tb + C_A1H, tb + C_A1L, tb + C_D1 + 32, K0 / 8, false);
        tc::umma_commit(&s.bar_d[1]);
        for (int c = 0; c < NCH; ++c) {
          const int b = c & 1;
          wait_a(b);   // A2[b] = ELU(D1[b] + b1) written, D1[b] drained
          mma_segment<H2>(m, tb + C_A2 + 64 * b, tb + C_A2 + 64 * b + 32, tb + C_D2, CH / 8, c > 0);
          if (c + 2 < NCH) {
            mma_segment<CH>(m, tb + C_A1H, tb + C_A1L, tb + C_D1 + 32 * b, K0 / 8, false);
            tc::umma_commit(&s.bar_d[b]);
          }
        }
        tc::umma_commit(&s.bar_d[0]);   // D2 complete
        wait_a(0);                      // A3
        if (WANT_PI) mma_segment<H3>(m, tb + C_A3H, tb + C_A3L, tb + C_D3, H2 / 8, false);
        if (WANT_V) mma_segment<H3>(m, tb + C_A3H, tb + C_A3L, tb + C_D3 + 64, H2 / 8, false);
        tc::umma_commit(&s.bar_d[0]);   // D3
        if (WANT_PI) {
          wait_a(0);                    // A4
          mma_segment<APAD>(m, tb + C_A4H, tb + C_A4L, tb + C_D4, H3 / 8, false);
          tc::umma_commit(&s.bar_d[0]); // D4
        }
      }
    }
  } else {
    // =========================================================================== epilogue warps
    const int q = warp & 3, h = warp >> 2;
    const uint32_t lane_base = (uint32_t)q * 32u;
    const int r = q * 32 + lane;   // row inside the tile
    uint32_t pd[2] = {0u, 0u};
    uint32_t obs_phase = 0;
    auto wait_d = [&](int b) {
      tc::mbar_wait(&s.bar_d[b], pd[b]);
      pd[b] ^= 1u;
      tc::tc_fence_after();
    };
    auto arrive_a = [&](int b) {
      tc::tmem_wait_st();
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&s.bar_a[b]);
    };
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long long row0 = tile * TM;
      const int nrows = (int)min((long long)TM, a.rows - row0);
      const uint32_t bytes = (uint32_t)nrows * (uint32_t)O * 4u;
      const float* src = a.obs + (size_t)row0 * O;
      const bool by_tma = (bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0;
      if (by_tma) {
        tc::mbar_wait(s.obs_full, obs_phase);
        obs_phase ^= 1u;
      } else {   // ragged / unaligned tile: cooperative loads
        for (int i = threadIdx.x; i < nrows * O; i += EPI_THREADS) s.obs[i] = __ldg(src + i);
        tc::fence_proxy_async();
        epi_bar_sync();
      }
      fin.begin_tile(s, row0, nrows, by_tma);

      // ---- A1 = split(obs row), zero padded to K0 columns; rows beyond nrows are zero
      {
        const float* xr = s.obs + r * O;
#pragma unroll
        for (int gI = 0; gI < K0 / 16; ++gI) {
          const int c0 = h * (K0 / 2) + gI * 8;
          uint32_t hi[8], lo[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int c = c0 + j;
            const float x = (r < nrows && c < O) ? xr[c] : 0.0f;
            tc::split_tf32(x, hi[j], lo[j]);
          }
          tc::tmem_st8(tc::tmem_addr(tb, lane_base, C_A1H + c0), hi);
          tc::tmem_st8(tc::tmem_addr(tb, lane_base, C_A1L + c0), lo);
        }
      }
      fin.obs_consumed(s, by_tma);   // releases the staging buffer (arrives on obs_empty when the tile came by TMA)
      arrive_a(0);

      // ---- trunk: 8 chunks of layer 1 -> layer-2 K-chunks
      for (int c = 0; c < NCH; ++c) {
        const int b = c & 1;
        wait_d(b);
        epi_elu16(tb, lane_base, C_D1 + 32 * b + 16 * h, C_A2 + 64 * b + 16 * h, C_A2 + 64 * b + 32 + 16 * h,
                  s.bias + B_B1 + c * CH + 16 * h, nullptr);
        arrive_a(b);
      }
      // ---- A3 = ELU(D2 + b2): 64 columns per thread
      wait_d(0);
#pragma unroll
      for (int gI = 0; gI < 4; ++gI) {
        const int c0 = h * 64 + gI * 16;
        epi_elu16(tb, lane_base, C_D2 + c0, C_A3H + c0, C_A3L + c0, s.bias + B_B2 + c0, nullptr);
      }
      arrive_a(0);
      // ---- heads, first layer
      wait_d(0);
      if (WANT_PI) {   // A4 = ELU(D3pi + b3pi): 32 columns per thread
#pragma unroll
        for (int gI = 0; gI < 2; ++gI) {
          const int c0 = h * 32 + gI * 16;
          epi_elu16(tb, lane_base, C_D3 + c0, C_A4H + c0, C_A4L + c0, s.bias + B_B3PI + c0, nullptr);
        }
        arrive_a(0);
      }
      if (WANT_V) {    // value = ELU(D3v + b3v) . w + b : each half sums 32 columns, half 0 finishes
        float part = 0.0f;
#pragma unroll
        for (int gI = 0; gI < 2; ++gI) {
          const int c0 = h * 32 + gI * 16;
          uint32_t v[16];
          tc::tmem_ld16(tc::tmem_addr(tb, lane_base, C_D3 + 64 + c0), v);
          tc::tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            part = fmaf(elu(__uint_as_float(v[j]) + s.bias[B_B3V + c0 + j]), s.bias[B_W4V + c0 + j], part);
        }
        if (h == 1) s.vpart[r] = part;
        epi_bar_sync();
        if (h == 0) fin.value(row0, r, nrows, part + s.vpart[r] + s.bias[B_B4V]);
        if (!WANT_PI) tc::tc_fence_before();   // D3 reads done before the next tile's MMAs overwrite the region
      }
      if (WANT_PI) {
        wait_d(0);
        float mu[APAD];
#pragma unroll
        for (int gI = 0; gI < 2; ++gI) {
          uint32_t v[16];
          tc::tmem_ld16(tc::tmem_addr(tb, lane_base, C_D4 + gI * 16), v);
          tc::tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j) mu[gI * 16 + j] = elu(__uint_as_float(v[j]) + s.bias[B_B4 + gI * 16 + j]);
        }
        tc::tc_fence_before();
        fin.pi(s, row0, r, h, nrows, mu);
      }
      if (WANT_V && !WANT_PI) epi_bar_sync();   // vpart is rewritten by the next tile
    }
    fin.finish(s);
  }

  // ---- teardown
  tc::tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tc::tmem_dealloc(tb, 512);
}

}  // namespace tcm
}  // namespace flyb
