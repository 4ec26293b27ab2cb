// Tensor-core forward of the actor-critic MLP (reference ppo.py:10-38, 88-102) on 128-row tiles:
// tcgen05.mma kind::tf32 (3xTF32, fp32 accumulate in tensor memory), weights streamed L2 -> shared memory by
// 1-D bulk TMA copies through an 8-stage (64 KB) mbarrier ring.  The observation tile is the only activation that lives
// in shared memory (hi/lo split, UMMA K-major layout: the A operand of layer 1); every later activation stays
// in TENSOR MEMORY and is fed back as the A operand of the next GEMM ("TS" mode), so a row never leaves the SM
// between the observation and the action / value it produces.
//
// One CTA = 10 warps:  warps 0-7  epilogue (TMEM -> registers -> bias + ELU -> hi/lo split -> TMEM),
//                      warp  8    TMA producer: observation tiles + the packed weight stream,
//                      warp  9    MMA issuer (warp-uniform code, one elected lane issues).
// Epilogue warp w owns TMEM lanes 32*(w%4)..+31 (row r of the tile = lane r) and column half w/4.
//
// Layer schedule of one tile (K0 = 80 padded obs, trunk 256 -> 128, heads 64 -> {A<=32, 1}):
//   the 256-wide first layer is produced in 4 chunks of 64 columns; chunk c is  D1[c&1] = X * W1[c]^T  (N = 64),
//   its epilogue writes ELU(D1 + b1) as the K-chunk A2[c&1] of the second layer, and  D2 += A2[c&1] * W2[:, c]^T
//   (N = 128) consumes it.  D1 / A2 are double buffered and the MMA warp issues  W1[0], W1[1], (W2[c], W1[c+2])...
//   so the epilogue of chunk c overlaps the layer-2 MMAs of chunk c-1 and the layer-1 MMAs of chunk c+1.
//   Then  A3 = ELU(D2 + b2),  D3 = A3 * W3^T (mean and/or value head, N = 64 each),  A4 = ELU(D3pi + b3),
//   D4 = A4 * W4^T (N = 32); the 64 -> 1 value output is a register dot product.
//
// TMEM columns (512 allocated):   D1[2] [0,128)      A2[2] hi|lo [128,384)     D2 [384,512)
//   after the trunk:              A3 hi|lo [0,256)   D3 pi|v [256,384)         A4 hi|lo [384,512)    D4 [0,32)
//
// Packed weight stream (flyb_pack_weights): 8 KB stages in consumption order, each k-step (8 k) of an N-row
// operand stored as [hi slab | lo slab], slab = N x 32 bytes in the no-swizzle K-major core-matrix layout
// (tc_common.cuh).  Trunk = W1[0], W1[1], W2[0], W1[2], W2[1], W1[3], W2[2], W2[3]  (W1 chunk: 5 stages,
// W2 K-chunk: 8 stages), then W3pi (8), W3v (8), W4 (2).
//
// TWO SHAPES, one source: the file is compiled once per shape, each into its own namespace (flyb::tcn / flyb::tcw):
//   narrow (default)      K0 =  80, APAD = 32   num_obs <= 80,  num_act <= 32   (the 18-joint actuation: O = 73, A = 18)
//   wide (FLYB_TC_WIDE)   K0 = 160, APAD = 48   num_obs <= 160, num_act <= 48   (the 42-joint set: O = 145, A = 42)
// The wide X planes alone are 161 KB, so the wide shape has no raw observation stage: the epilogue warps pull the tile
// with 4-byte cp.async straight into the hi plane (chunk pitch padded by 16 bytes so that consecutive columns do not
// land on one bank) and split it in place; the finaliser's staging aliases the X planes, which are dead once the last
// layer-1 MMA has completed (mu-independent staging work moves from `early` to that point); the weight ring has 6 stages.
#pragma once
#include "tc_common.cuh"

#ifndef FLYB_TC_WIDE
#define FLYB_TC_WIDE 0
#endif
#if FLYB_TC_WIDE
#define FLYB_TC_VNS tcw
#else
#define FLYB_TC_VNS tcn
#endif

namespace flyb {
namespace FLYB_TC_VNS {
namespace tcm {

constexpr bool WIDE = FLYB_TC_WIDE != 0;
constexpr int TM = 128;
constexpr int K0 = WIDE ? 160 : 80;
constexpr int H1 = 256, H2 = 128, H3 = 64, APAD = WIDE ? 48 : 32;
constexpr int CH = 64, NCH = H1 / CH;
constexpr int STAGE = 8192;
constexpr int NST = WIDE ? 6 : 8;              // ring stages (64 KB / 48 KB)
FLYB_HD uint32_t ring_slot(uint32_t stage) { return stage % (uint32_t)NST; }
FLYB_HD uint32_t ring_phase(uint32_t stage) { return (stage / (uint32_t)NST) & 1u; }
constexpr int EPI_WARPS = 8, EPI_THREADS = EPI_WARPS * 32;
constexpr int PRODUCER_WARP = 8, MMA_WARP = 9;
constexpr int THREADS = 320;

// an operand with N rows: one k-step (8 k) = [hi slab | lo slab] of N x 32 bytes each; a stage holds KPS whole k-steps
// (N = 48 leaves a quarter of its stages unused)
constexpr int kps_for(int n) { return STAGE / (2 * n * 32); }
constexpr int stages_for(int n, int ksteps) { return ksteps / kps_for(n); }
constexpr int W1C_STAGES = stages_for(CH, K0 / 8);             // 5 / 10
constexpr int W2C_STAGES = stages_for(H2, CH / 8);             // 8
constexpr int TRUNK_STAGES = NCH * (W1C_STAGES + W2C_STAGES);  // 52 / 72
constexpr int W3_STAGES = stages_for(H3, H2 / 8);              // 8
constexpr int W4_STAGES = stages_for(APAD, H3 / 8);            // 2 / 4
constexpr int OFF_W3PI = TRUNK_STAGES, OFF_W3V = OFF_W3PI + W3_STAGES, OFF_W4 = OFF_W3V + W3_STAGES;
constexpr int FWD_STAGES = OFF_W4 + W4_STAGES;                 // 70 stages = 560 KB / 92 stages

// TMEM column map
constexpr uint32_t C_D1 = 0, C_A2 = 128, C_D2 = 384;
constexpr uint32_t C_A3H = 0, C_A3L = H2, C_D3 = 256, C_A4H = 384, C_A4L = 448, C_D4 = 0;

// bias block in shared memory (floats)
constexpr int B_B1 = 0, B_B2 = 256, B_B3PI = 384, B_B3V = 448, B_B4 = 512, B_W4V = B_B4 + APAD, B_B4V = B_W4V + 64,
              B_TOTAL = B_B4V + 32;

__host__ __device__ inline size_t obs_stage_bytes(int O) { return WIDE ? 0 : ((size_t)TM * O * 4 + 127) / 128 * 128; }
// K-major no-swizzle A operand: [k/4][row][4 floats]; the wide shape pads the chunk pitch (see the header comment)
constexpr uint32_t X_LBO = TM * 16 + (WIDE ? 16 : 0), X_SBO = 128;
constexpr size_t X_PLANE_BYTES = (size_t)(K0 / 4) * X_LBO;   // 40960 / 82560: one of the hi / lo planes
FLYB_D uint4* x_chunk(float* plane, int kc, int r) {
  return reinterpret_cast<uint4*>(reinterpret_cast<unsigned char*>(plane) + (size_t)kc * X_LBO) + r;
}

struct Smem {
  unsigned char* ring;
  float* xhi;            // layer-1 A operand, hi plane (UMMA layout)
  float* xlo;
  float* obs;            // raw observation tile (row-major, O floats per row)
  float* out;
  float* bias;
  float* vpart;
  uint64_t *full, *empty, *bar_a, *bar_d, *obs_full, *obs_empty;
  uint32_t* tmem_slot;
};

__host__ __device__ inline size_t smem_bytes(int O, size_t out_bytes) {
  // wide: the finaliser's staging aliases the X planes (out_bytes <= 2 * X_PLANE_BYTES, checked by the launchers)
  return 1024 + (size_t)NST * STAGE + 2 * X_PLANE_BYTES + obs_stage_bytes(O) +
         (WIDE ? 0 : ((out_bytes + 127) / 128 * 128)) + B_TOTAL * 4 + TM * 4 + (2 * NST + 6) * 8 + 16;
}

// `raw` must be the kernel's `extern __shared__` array itself: every pointer below is derived from it by plain
// pointer arithmetic, so that the compiler keeps the shared state space (LDS / STS instead of generic LD / ST).
FLYB_D Smem carve(unsigned char* raw, int O, size_t out_bytes) {
  Smem s;
  const uint32_t pad = (1024u - (tc::smem_u32(raw) & 1023u)) & 1023u;
  unsigned char* p = raw + pad;
  s.ring = p;
  p += (size_t)NST * STAGE;
  s.xhi = reinterpret_cast<float*>(p);
  p += X_PLANE_BYTES;
  s.xlo = reinterpret_cast<float*>(p);
  p += X_PLANE_BYTES;
  s.obs = WIDE ? nullptr : reinterpret_cast<float*>(p);
  p += obs_stage_bytes(O);
  s.out = WIDE ? s.xhi : reinterpret_cast<float*>(p);
  p += WIDE ? 0 : (out_bytes + 127) / 128 * 128;
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
// The MMA warp runs warp-uniform code (all 32 lanes wait on the barriers and compute the descriptors, which keeps
// them in uniform registers); `leader` (elect.sync) predicates the tcgen05 instructions themselves.
FLYB_D void umma_ts_if(uint32_t leader, uint32_t d, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(acc), "r"(leader)
      : "memory");
}
FLYB_D void umma_ss_if(uint32_t leader, uint32_t d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\tsetp.ne.b32 p, %4, 0;\n\tsetp.ne.b32 q, %5, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(acc), "r"(leader)
      : "memory");
}
FLYB_D void umma_commit_if(uint32_t leader, uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %1, 0;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(tc::smem_u32(bar)),
      "r"(leader)
      : "memory");
}

struct MmaState {
  uint32_t leader;
  uint32_t ring;       // shared-memory byte address of the weight ring
  uint64_t *full, *empty;
  uint32_t stage;      // stages consumed so far (running over tiles)
};

// One GEMM segment: D[128 x N] (+)= A[128 x 8*KSTEPS] * B^T with B = the next 2*KSTEPS slabs of the ring.
// A_SMEM = false: a_hi / a_lo are TMEM addresses (8 columns per k-step);
// A_SMEM = true:  a_hi / a_lo are shared-memory byte addresses of the X planes (2 * X_LBO bytes per k-step).
template <int N, int KSTEPS, bool A_SMEM>
FLYB_D void mma_segment(MmaState& m, uint32_t a_hi, uint32_t a_lo, uint32_t d, bool accumulate) {
  constexpr uint32_t SLAB = N * 32;
  const uint32_t idesc = tc::umma_idesc_tf32(N);
  auto issue3 = [&](int ks, uint32_t hi_addr, uint32_t lo_addr) {
    const uint64_t b_hi = tc::umma_smem_desc(hi_addr, N * 16, 128);
    const uint64_t b_lo = tc::umma_smem_desc(lo_addr, N * 16, 128);
    const uint32_t acc0 = (accumulate || ks > 0) ? 1u : 0u;
    if constexpr (A_SMEM) {
      const uint64_t xa_hi = tc::umma_smem_desc(a_hi + (uint32_t)ks * 2u * X_LBO, X_LBO, X_SBO);
      const uint64_t xa_lo = tc::umma_smem_desc(a_lo + (uint32_t)ks * 2u * X_LBO, X_LBO, X_SBO);
      umma_ss_if(m.leader, d, xa_lo, b_hi, idesc, acc0);
      umma_ss_if(m.leader, d, xa_hi, b_lo, idesc, 1u);
      umma_ss_if(m.leader, d, xa_hi, b_hi, idesc, 1u);
    } else {
      umma_ts_if(m.leader, d, a_lo + 8 * ks, b_hi, idesc, acc0);
      umma_ts_if(m.leader, d, a_hi + 8 * ks, b_lo, idesc, 1u);
      umma_ts_if(m.leader, d, a_hi + 8 * ks, b_hi, idesc, 1u);
    }
  };
  static_assert(2 * SLAB <= STAGE, "a k-step (hi + lo slab) must fit one stage");
  constexpr int KPS = kps_for(N);   // whole k-steps per stage
  static_assert(KSTEPS % KPS == 0, "segment must end on a stage boundary");
  // rolled over stages (the whole kernel runs once per tile: straight-line code would only miss in the
  // instruction cache), unrolled over the KPS k-steps of a stage
#pragma unroll 1
  for (int ks0 = 0; ks0 < KSTEPS; ks0 += KPS) {
    const uint32_t slot = ring_slot(m.stage);
    tc::mbar_wait(&m.full[slot], ring_phase(m.stage));
    tc::tc_fence_after();
    const uint32_t base = m.ring + slot * STAGE;
#pragma unroll
    for (int j = 0; j < KPS; ++j) {
      const uint32_t hi_addr = base + (uint32_t)j * 2u * SLAB;
      issue3(ks0 + j, hi_addr, hi_addr + SLAB);
    }
    umma_commit_if(m.leader, &m.empty[slot]);
    ++m.stage;
  }
}

// ---------------------------------------------------------------------------------------------- epilogue helpers
// ELU of the tensor-core path: exp through ex2.approx (|error| <= ~2.4e-7 absolute), x > 0 ? x : exp(x) - 1
FLYB_D float elu_fast(float y) {
  float e;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(e) : "f"(y * 1.4426950408889634f));
  return y > 0.0f ? y : e - 1.0f;
}
// hi = rna_tf32(y); lo = y - hi (exact; the tensor core reads its upper 19 bits)
// The rounding is done on the integer pipe: adding half an ulp of the 10-bit mantissa to the magnitude bits and
// clearing the 13 low bits is round-to-nearest, ties away from zero — the same bits cvt.rna.tf32.f32 gives for finite
// inputs — but costs two full-rate ALU instructions instead of one quarter-rate conversion.
FLYB_D void split_fast(float y, uint32_t& hi, uint32_t& lo) {
  hi = (__float_as_uint(y) + 0x1000u) & 0xFFFFE000u;
  lo = __float_as_uint(y - __uint_as_float(hi));
}

// y = ELU(acc + bias) for 32 consecutive accumulator columns (two independent 16-column groups in flight),
// split to hi/lo and stored as the next A operand
FLYB_D void epi_elu32(uint32_t d_addr, uint32_t ah_addr, uint32_t al_addr, const float* bias) {
  uint32_t v0[16], v1[16];
  tc::tmem_ld16(d_addr, v0);
  tc::tmem_ld16(d_addr + 16, v1);
  tc::tmem_wait_ld();
  const float4* b4 = reinterpret_cast<const float4*>(bias);
#pragma unroll
  for (int g = 0; g < 2; ++g) {
    uint32_t hi[16], lo[16];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 b = b4[4 * g + q];
      const uint32_t* v = g ? v1 : v0;
      split_fast(elu_fast(__uint_as_float(v[4 * q + 0]) + b.x), hi[4 * q + 0], lo[4 * q + 0]);
      split_fast(elu_fast(__uint_as_float(v[4 * q + 1]) + b.y), hi[4 * q + 1], lo[4 * q + 1]);
      split_fast(elu_fast(__uint_as_float(v[4 * q + 2]) + b.z), hi[4 * q + 2], lo[4 * q + 2]);
      split_fast(elu_fast(__uint_as_float(v[4 * q + 3]) + b.w), hi[4 * q + 3], lo[4 * q + 3]);
    }
    tc::tmem_st16(ah_addr + 16 * g, hi);
    tc::tmem_st16(al_addr + 16 * g, lo);
  }
}
// `ncols` (multiple of 32) columns starting at d_col -> hi at ah_col, lo at al_col
FLYB_D void epi_elu(uint32_t tb, uint32_t lane_base, uint32_t d_col, uint32_t ah_col, uint32_t al_col,
                    const float* bias, int ncols) {
  const uint32_t t0 = tc::tmem_addr(tb, lane_base, 0);
#pragma unroll 1
  for (int c = 0; c < ncols; c += 32) epi_elu32(t0 + d_col + c, t0 + ah_col + c, t0 + al_col + c, bias + c);
}

struct FwdArgs {
  const float* obs;          // (rows, O)
  const unsigned char* wpack;
  const float* params;       // flat fp32 parameters (biases and the 64 -> 1 value row are read from here)
  long long rows;
  int O, A;
  size_t out_bytes;
  unsigned long long* trace;   // debug: clock64 timeline of CTA 0 (nullable; flyb_debug_set_trace)
  // Final::kBuildsObs kernels only: a per-row block of `pre_row_bytes` bytes (the env's whole contact-force block) is
  // bulk-copied for the CTA's first tile into the region that starts at the X planes (X | obs | out: all free until
  // the observation rows are built), signalled on obs_full.  NULL / 0: nothing.
  const float* pre_src = nullptr;
  unsigned int pre_row_bytes = 0;
};
// bytes of the contiguous shared-memory region [X hi | X lo | obs stage | out staging]
__host__ __device__ inline size_t build_region_bytes(int O, size_t out_bytes) {
  return 2 * X_PLANE_BYTES + obs_stage_bytes(O) + ((out_bytes + 127) / 128 * 128);
}

// trace slots: [0,64) MMA warp, [64,128) epilogue warp 0 lane 0
FLYB_D void trace_mark(unsigned long long* trace, int& idx, int limit) {
  if (trace != nullptr && blockIdx.x == 0 && idx < limit) trace[idx++] = clock64();
}

// Barrier set-up + the first loads of a CTA, issued by the producer warp BEFORE the kernel's other set-up work
// (tables, biases, TMEM allocation) so that their latency is hidden behind it: all mbarriers are initialised, then the
// first tile's observation rows and the first NST stages of the weight stream are put in flight.  The producer loop
// later skips exactly these (it still does its barrier bookkeeping).  Call once, by all 32 lanes of warp
// PRODUCER_WARP, at kernel start; every other thread must pass a __syncthreads() before touching a barrier.
FLYB_D void producer_kickoff(const Smem& s, const FwdArgs& a, bool load_obs = true) {
  if (tc::elect_one()) {
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
    const long long ntiles = (a.rows + TM - 1) / TM;
    if ((long long)blockIdx.x < ntiles) {
      const long long row0 = (long long)blockIdx.x * TM;
      const int nrows = (int)min((long long)TM, a.rows - row0);
      const uint32_t bytes = (uint32_t)nrows * (uint32_t)a.O * 4u;
      const float* src = a.obs + (size_t)row0 * a.O;
      if (load_obs && (bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
        tc::mbar_arrive_expect_tx(s.obs_full, bytes);
        tc::tma_load_1d(s.obs, src, bytes, s.obs_full);
      }
      if (!load_obs && a.pre_src != nullptr) {   // the first tile's contact-force blocks (see FwdArgs::pre_src)
        const uint32_t pbytes = (uint32_t)nrows * a.pre_row_bytes;
        if ((pbytes & 15u) == 0) {
          tc::mbar_arrive_expect_tx(s.obs_full, pbytes);
          tc::tma_load_1d(s.xhi, reinterpret_cast<const char*>(a.pre_src) + (size_t)row0 * a.pre_row_bytes, pbytes,
                          s.obs_full);
        }
      }
      for (int i = 0; i < NST; ++i) {
        tc::mbar_arrive_expect_tx(&s.full[i], STAGE);
        tc::tma_load_1d(s.ring + (size_t)i * STAGE, a.wpack + (size_t)i * STAGE, STAGE, &s.full[i]);
      }
    }
  }
  __syncwarp();
}

// biases + the 64 -> 1 value row -> the shared-memory bias block, by `nthreads` cooperating threads
FLYB_D void load_bias_block(const Smem& s, const FlybNetLayout& net, const float* p, int A, int first, int nthreads) {
  constexpr int BATCH = 20;   // loads in flight per thread: the block is filled by one warp, latency is what counts
  for (int i0 = first; i0 < B_TOTAL; i0 += BATCH * nthreads) {
    float v[BATCH];
#pragma unroll
    for (int u = 0; u < BATCH; ++u) {
      const int i = i0 + u * nthreads;
      int src = -1;
      if (i < B_B2) src = net.off[1] + i;
      else if (i < B_B3PI) src = net.off[3] + i - B_B2;
      else if (i < B_B3V) src = net.off[5] + i - B_B3PI;
      else if (i < B_B4) src = net.off[9] + i - B_B3V;
      else if (i < B_W4V) src = (i - B_B4 < A) ? net.off[7] + i - B_B4 : -1;
      else if (i < B_B4V) src = net.off[10] + i - B_W4V;
      else if (i == B_B4V) src = net.off[11];
      v[u] = src >= 0 ? __ldg(p + src) : 0.0f;
    }
#pragma unroll
    for (int u = 0; u < BATCH; ++u) {
      const int i = i0 + u * nthreads;
      if (i < B_TOTAL) s.bias[i] = v[u];
    }
  }
}
// Late set-up hand-off: the producer warp fills the bias block / the finaliser's tables while the epilogue warps are
// already splitting the first observation tile; they meet on named barrier 2 before the first bias is read.
constexpr int SETUP_BAR_THREADS = EPI_THREADS + 32;
FLYB_D void setup_arrive() {
  __threadfence_block();
  asm volatile("bar.arrive 2, %0;" ::"n"(SETUP_BAR_THREADS) : "memory");
}
FLYB_D void setup_wait() { asm volatile("bar.sync 2, %0;" ::"n"(SETUP_BAR_THREADS) : "memory"); }

// Per-tile consumer of the results (`Final`), called by the 256 epilogue threads:
//   setup(lane)                            once, by the 32 lanes of the producer warp: per-kernel tables (any thread
//                                          may read them after its setup_wait())
//   begin_tile(s, row0, r, nrows)          after the raw observation tile is in shared memory (s.obs, row-major)
//   obs_consumed(s)                        after the A operand of layer 1 was built; must arrive on s.obs_empty
//   early(s, row0, nrows)                  right after the layer-1 operand was handed to the MMA warp
//   pre_pi(s, row0, r, h, nrows)           before the last accumulator is waited for (mu-independent work)
//   pi(s, row0, r, h, nrows, mu[32])       both column halves call it with the full mu row (ELU applied)
//   value(row0, r, nrows, v)               half 0 only
//   finish(s)                              once, after the last tile
//   Final::kBuildsObs (static constexpr bool)   true: the observation tile is not loaded from `a.obs` but BUILT by the
//                                          epilogue warps — build_obs(s, row0, nrows) must leave the tile's rows in
//                                          s.obs (row-major, O floats per row), fenced for the async proxy and with all
//                                          256 threads synchronised (the fused env step: obs + reward + reset + act)
//   aux_warp(s, ntiles)                    body of any warp beyond the ten above (a kernel launched with more threads)
// Implementations: policy_kernels_tc.cu.
// KICKED: producer_kickoff() already ran (at the top of the kernel); otherwise it runs here.
template <bool WANT_PI, bool WANT_V, class Final, bool KICKED = false>
FLYB_D void forward_persistent(const FlybNetLayout& net, const FwdArgs& a, Final& fin) {
  extern __shared__ unsigned char smem_raw[];
  const Smem s = carve(smem_raw, a.O, a.out_bytes);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int O = a.O;
  const long long ntiles = (a.rows + TM - 1) / TM;

  // ---- one-time setup
  constexpr bool BUILD = Final::kBuildsObs;
  constexpr bool NO_TMA_OBS = BUILD || WIDE;   // the producer warp does not load observation tiles
  static_assert(!(BUILD && WIDE), "the fused env step exists for the narrow shape only");
  if (!KICKED && warp == PRODUCER_WARP) producer_kickoff(s, a, !NO_TMA_OBS);
  if (warp == MMA_WARP) tc::tmem_alloc(s.tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tb = *s.tmem_slot;

  if (warp == PRODUCER_WARP) {
    // =========================================================================== TMA producer (warp-uniform)
    // the first loads are in flight (producer_kickoff): fill the bias block and the finaliser's tables in their shadow
    load_bias_block(s, net, a.params, a.A, lane, 32);
    fin.setup(lane);
    setup_arrive();
    const uint32_t leader = tc::elect_one() ? 1u : 0u;
    const uint32_t ring = tc::smem_u32(s.ring);
    uint32_t gs = 0, obs_phase = 0;
    int kicked = NST;   // stages (and one observation tile) producer_kickoff() already put in flight
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long long row0 = tile * TM;
      const int nrows = (int)min((long long)TM, a.rows - row0);
      const uint32_t bytes = (uint32_t)nrows * (uint32_t)O * 4u;
      const float* src = a.obs + (size_t)row0 * O;
      // The first tile's rows were requested by producer_kickoff(): no wait for it (the epilogue warps may already
      // have released that buffer by now — waiting for "the phase before" would then wait for the NEXT release).
      if (kicked == 0 && !NO_TMA_OBS) {
        tc::mbar_wait_relaxed(s.obs_empty, obs_phase);   // the epilogue warps are done with the previous tile's rows
        obs_phase ^= 1u;
        if ((bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0 && leader) {
          tc::mbar_arrive_expect_tx(s.obs_full, bytes);
          tc::tma_load_1d(s.obs, src, bytes, s.obs_full);
        }
      }
      auto stream = [&](int first, int count) {
        const unsigned char* g = a.wpack + (size_t)first * STAGE;
        for (int i = 0; i < count; ++i, ++gs, g += STAGE) {
          if (kicked > 0) { --kicked; continue; }
          const uint32_t slot = ring_slot(gs);
          tc::mbar_wait_relaxed(&s.empty[slot], ring_phase(gs) ^ 1u);
          const uint32_t bar = tc::smem_u32(&s.full[slot]);
          asm volatile(
              "{\n\t.reg .pred q;\n\t.reg .b64 st;\n\tsetp.ne.b32 q, %4, 0;\n\t"
              "@q mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %3;\n\t"
              "@q cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%1], [%2], %3, [%0];\n\t}" ::"r"(bar),
              "r"(ring + slot * STAGE), "l"(g), "r"((uint32_t)STAGE), "r"(leader)
              : "memory");
        }
      };
      stream(0, TRUNK_STAGES);
      if (WANT_PI) stream(OFF_W3PI, W3_STAGES);
      if (WANT_V) stream(OFF_W3V, W3_STAGES);
      if (WANT_PI) stream(OFF_W4, W4_STAGES);
    }
  } else if (warp == MMA_WARP) {
    // =========================================================================== MMA issuer (warp-uniform)
    MmaState m{tc::elect_one() ? 1u : 0u, tc::smem_u32(s.ring), s.full, s.empty, 0u};
    const uint32_t xh = tc::smem_u32(s.xhi), xl = tc::smem_u32(s.xlo);
    uint32_t pa = 0u;   // bit b = phase of bar_a[b]
    int ti = 0;
    unsigned long long* mtrace = (lane == 0) ? a.trace : nullptr;
    auto wait_a = [&](int b) {
      tc::mbar_wait_relaxed(&s.bar_a[b], (pa >> b) & 1u);
      pa ^= 1u << b;
      tc::tc_fence_after();
    };
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      trace_mark(mtrace, ti, 64);
      wait_a(0);   // X staged
      trace_mark(mtrace, ti, 64);
      // issue order (== packed stream order): W1[0], W1[1], then (W2[c], W1[c+2]) for c = 0..3
#pragma unroll 1
      for (int c = -2; c < NCH; ++c) {
        if (c >= 0) {
          const int b = c & 1;
          wait_a(b);                    // A2[b] = ELU(D1[b] + b1) written, D1[b] drained
          trace_mark(mtrace, ti, 64);
          mma_segment<H2, CH / 8, false>(m, tb + C_A2 + 128 * b, tb + C_A2 + 128 * b + CH, tb + C_D2, c > 0);
        }
        if (c + 2 < NCH) {
          const int b2 = (c + 2) & 1;
          mma_segment<CH, K0 / 8, true>(m, xh, xl, tb + C_D1 + CH * b2, false);
          umma_commit_if(m.leader, &s.bar_d[b2]);
        }
      }
      umma_commit_if(m.leader, &s.bar_d[0]);   // D2 complete
      trace_mark(mtrace, ti, 64);
      wait_a(0);                               // A3
      trace_mark(mtrace, ti, 64);
      if (WANT_PI) mma_segment<H3, H2 / 8, false>(m, tb + C_A3H, tb + C_A3L, tb + C_D3, false);
      if (WANT_V) mma_segment<H3, H2 / 8, false>(m, tb + C_A3H, tb + C_A3L, tb + C_D3 + H3, false);
      umma_commit_if(m.leader, &s.bar_d[0]);   // D3
      if (WANT_PI) {
        trace_mark(mtrace, ti, 64);
        wait_a(0);                             // A4
        trace_mark(mtrace, ti, 64);
        mma_segment<APAD, H3 / 8, false>(m, tb + C_A4H, tb + C_A4L, tb + C_D4, false);
        umma_commit_if(m.leader, &s.bar_d[0]); // D4
        trace_mark(mtrace, ti, 64);
      }
    }
  } else if (warp < EPI_WARPS) {
    // =========================================================================== epilogue warps
    const int q = warp & 3, h = warp >> 2;
    const uint32_t lane_base = (uint32_t)q * 32u;
    const int r = q * 32 + lane;   // row inside the tile
    uint32_t pd = 0u, obs_phase = 0u;   // bit b of pd = phase of bar_d[b]
    bool setup_seen = false;
    int ti = 64;
    unsigned long long* etrace = (threadIdx.x == 0) ? a.trace : nullptr;
    auto wait_d = [&](int b) {
      tc::mbar_wait(&s.bar_d[b], (pd >> b) & 1u);
      pd ^= 1u << b;
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
      if constexpr (WIDE) {
        // the previous tile's finaliser is done with the staging that aliases the X planes (its bulk store included)
        if (threadIdx.x == 0) tc::tma_store_wait_read<0>();
        epi_bar_sync();
        // the tile's nrows x O floats are contiguous: lane-consecutive 4-byte cp.async into the hi plane's K-major slots
        {
          const int n = nrows * O;
          const int r_step = EPI_THREADS / O, c_step = EPI_THREADS % O;
          int rr = threadIdx.x / O, c = threadIdx.x - rr * O;
          unsigned char* plane = reinterpret_cast<unsigned char*>(s.xhi);
          for (int e = threadIdx.x; e < n; e += EPI_THREADS) {
            const uint32_t dst = tc::smem_u32(plane + (size_t)(c >> 2) * X_LBO + (size_t)rr * 16 + (c & 3) * 4);
            asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(dst), "l"(src + e) : "memory");
            rr += r_step;
            c += c_step;
            if (c >= O) { c -= O; ++rr; }
          }
          cp_async_commit();
          cp_async_wait<0>();
          epi_bar_sync();
        }
      } else if constexpr (BUILD) {
        fin.build_obs(s, row0, nrows);
      } else if (by_tma) {
        tc::mbar_wait(s.obs_full, obs_phase);
        obs_phase ^= 1u;
      } else {   // ragged / unaligned tile: cooperative loads
        for (int i = threadIdx.x; i < nrows * O; i += EPI_THREADS) s.obs[i] = __ldg(src + i);
        tc::fence_proxy_async();
        epi_bar_sync();
      }
      trace_mark(etrace, ti, 128);
      fin.begin_tile(s, row0, r, nrows);

      // ---- X = split(obs row) in the UMMA K-major layout, zero padded to K0 columns and TM rows
      //      (wide: the raw values already sit in their hi-plane slots and are split in place)
      {
        const float* xr = WIDE ? nullptr : s.obs + r * O;
        const bool live = r < nrows;
#pragma unroll 2
        for (int g = 0; g < K0 / 8; ++g) {
          const int kc = h * (K0 / 8) + g;   // 16-byte K chunk
          uint32_t hi[4], lo[4];
          float raw[4];
          if constexpr (WIDE) {
            const uint4 v = *x_chunk(s.xhi, kc, r);
            raw[0] = __uint_as_float(v.x); raw[1] = __uint_as_float(v.y);
            raw[2] = __uint_as_float(v.z); raw[3] = __uint_as_float(v.w);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = kc * 4 + j;
            float x = 0.0f;
            if (live && c < O) x = WIDE ? raw[j] : xr[c];
            split_fast(x, hi[j], lo[j]);
          }
          *x_chunk(s.xhi, kc, r) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *x_chunk(s.xlo, kc, r) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
        }
        tc::fence_proxy_async();   // generic-proxy writes -> visible to the tensor core's shared-memory reads
      }
      trace_mark(etrace, ti, 128);
      fin.obs_consumed(s);
      arrive_a(0);
      if constexpr (!WIDE) fin.early(s, row0, nrows);   // mu-independent work, overlapped with the first MMAs of the tile
      if (!setup_seen) {           // bias block + tables are complete from here on
        setup_wait();
        setup_seen = true;
      }

      // ---- trunk: 4 chunks of layer 1 -> layer-2 K-chunks (32 columns per thread)
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const int b = c & 1;
        wait_d(b);
        trace_mark(etrace, ti, 128);
        epi_elu(tb, lane_base, C_D1 + CH * b + 32 * h, C_A2 + 128 * b + 32 * h, C_A2 + 128 * b + CH + 32 * h,
                s.bias + B_B1 + c * CH + 32 * h, 32);
        arrive_a(b);
        trace_mark(etrace, ti, 128);
      }
      // wide: every layer-1 MMA has completed (chunk 3's accumulator was seen), the X planes are dead -> staging
      if constexpr (WIDE) fin.early(s, row0, nrows);
      // ---- A3 = ELU(D2 + b2): 64 columns per thread
      wait_d(0);
      trace_mark(etrace, ti, 128);
      epi_elu(tb, lane_base, C_D2 + 64 * h, C_A3H + 64 * h, C_A3L + 64 * h, s.bias + B_B2 + 64 * h, 64);
      arrive_a(0);
      trace_mark(etrace, ti, 128);
      // ---- heads, first layer
      wait_d(0);
      trace_mark(etrace, ti, 128);
      if (WANT_PI) {   // A4 = ELU(D3pi + b3pi): 32 columns per thread
        epi_elu(tb, lane_base, C_D3 + 32 * h, C_A4H + 32 * h, C_A4L + 32 * h, s.bias + B_B3PI + 32 * h, 32);
        arrive_a(0);
      }
      if (WANT_V) {    // value = ELU(D3v + b3v) . w + b : each half sums 32 columns, half 0 finishes
        float part = 0.0f;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const int c0 = h * 32 + g * 16;
          uint32_t v[16];
          tc::tmem_ld16(tc::tmem_addr(tb, lane_base, C_D3 + H3 + c0), v);
          tc::tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j)
            part = fmaf(elu_fast(__uint_as_float(v[j]) + s.bias[B_B3V + c0 + j]), s.bias[B_W4V + c0 + j], part);
        }
        if (h == 1) s.vpart[r] = part;
        tc::tc_fence_before();
        epi_bar_sync();
        if (h == 0) fin.value(row0, r, nrows, part + s.vpart[r] + s.bias[B_B4V]);
      }
      if (WANT_PI) {
        trace_mark(etrace, ti, 128);
        fin.pre_pi(s, row0, r, h, nrows);   // work that does not depend on mu, done while the last MMAs run
        wait_d(0);
        trace_mark(etrace, ti, 128);
        float mu[APAD];
#pragma unroll
        for (int g = 0; g < APAD / 16; ++g) {
          uint32_t v[16];
          tc::tmem_ld16(tc::tmem_addr(tb, lane_base, C_D4 + g * 16), v);
          tc::tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j) mu[g * 16 + j] = elu_fast(__uint_as_float(v[j]) + s.bias[B_B4 + g * 16 + j]);
        }
        tc::tc_fence_before();
        fin.pi(s, row0, r, h, nrows, mu);
        trace_mark(etrace, ti, 128);
      }
      if (WANT_V && !WANT_PI) epi_bar_sync();   // vpart is rewritten by the next tile
    }
    if (!setup_seen) setup_wait();
    fin.finish(s);
  } else {
    fin.aux_warp(s, ntiles);
  }

  // ---- teardown
  tc::tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tc::tmem_dealloc(tb, 512);
}

}  // namespace tcm
}  // namespace FLYB_TC_VNS
}  // namespace flyb
