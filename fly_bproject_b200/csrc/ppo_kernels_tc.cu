// Kernel (d), tensor-core path: one optimiser step's gradient of PPO.update (reference ppo.py:183-197) on
// tcgen05 / TMEM — forward of both heads on ONE trunk evaluation, clipped-surrogate + SmoothL1 loss, backward
// chain and weight gradients as 3xTF32 GEMMs with fp32 accumulation.  Same results contract as the CUDA-core path
// in ppo_kernels.cu.
//
//   ppo_fwd_bwd_tc_kernel  persistent over 128-row tiles, same role split as the forward (mlp_tc.cuh): 8 epilogue
//       warps, a TMA producer warp, an MMA issuer warp.  The forward is the forward of mlp_tc.cuh; its epilogues
//       additionally write every layer input (X, h1, h2, [m1|v1]) to the workspace as fp32 "row images".  The loss
//       terms are evaluated per row in registers; dL/dmu and dL/dv start the backward chain, which runs as TS-mode
//       MMAs (A = dZ in tensor memory) against TRANSPOSED weight slabs of the packed stream:
//           dM1 = dZmu * W4          dH2 = [dZm1 | dZv1] * [W3pi ; W3v]          dH1 = dZh2 * W2  (two N = 128 halves)
//       and each epilogue multiplies by ELU'(y) = y > 0 ? 1 : y + 1.  Every dZ is written as a row image too.
//   ppo_dw_tc_kernel       dW_l = dZ_l^T * X_l: the sample dimension is the MMA's K.  Row images are streamed by bulk
//       TMA, transposed + hi/lo split by 16 warps into K-major UMMA operands in shared memory (SS MMAs), accumulators
//       stay in TMEM over all tiles of a CTA; bias gradients are exact fp32 column sums taken by the transposers.
//       CTAs alternate between two roles (layers {2,4} and {1,3}) so that a role's accumulators fit 512 columns.
//   ppo_reduce_tc_kernel   grad = sum over the per-CTA partials in a fixed order (deterministic), loss.
//
// Row image of an operand with F features, one 128-row tile: [16 k-steps][F/4][8 rows][4 floats]: a thread that owns
// row r writes/reads 16-byte chunks, a warp's access is 4 x 128 contiguous bytes, and a k-step group is one
// contiguous block for the bulk copy.
//
// TMEM columns of the fwd/bwd kernel.  Trunk as in mlp_tc.cuh.  Afterwards:
//   A3 (h2) hi|lo [0,256)   D3 pi|v [256,384)   A4 (m1) hi|lo [384,512)   D4 (mu) [256,288)
//   backward:  A5 (dZmu) hi [0,32) lo [32,64)      D5 (dM1) [320,384)
//              A6 ([dZm1|dZv1]) hi [0,128) lo [128,256)      D6 (dH2) [384,512)
//              A7 (dZh2) hi [0,128) lo [128,256)             D7 (dH1) [256,512)
#include <algorithm>
#include <cstdlib>

#include "adam.cuh"
#include "mlp_tc.cuh"
#include "peer.cuh"

namespace flyb {

int num_sms();
int validate_net(const FlybNetLayout* net, const float* params);
unsigned long long* tc_get_trace();

#ifndef FLYB_STEP_FUSE_DEFINED
#define FLYB_STEP_FUSE_DEFINED
struct StepFuse {   // see ppo_kernels.cu
  long long* step;
  const float* sq_part;
  int sq_parts;
  int tail;
  float *params, *exp_avg, *exp_avg_sq, *grad_norm_out;
  double lr, beta1, beta2, eps, max_norm;
  int done;
  float grad_scale;   // 1 / world when the tail also does the data-parallel exchange (peers != NULL), else 1
  int exchange;
};
#endif

namespace FLYB_TC_VNS {   // tcn (narrow shape) or tcw (wide shape, FLYB_TC_WIDE): see mlp_tc.cuh
bool tc_supported(const FlybNetLayout* net);
const int* tc_wpack_index(const FlybNetLayout* net, const void* wpack);

namespace tcb {
using namespace tcm;

constexpr int W4T_STAGES = stages_for(H3, APAD / 8);                  // 2 / 3
constexpr int W3T_STAGES = stages_for(H2, 2 * H3 / 8);                // 16
constexpr int W2T_STAGES = 2 * stages_for(128, H2 / 8);               // 32
constexpr int ALL_STAGES = FWD_STAGES + W4T_STAGES + W3T_STAGES + W2T_STAGES;   // 120 / 143

// image feature counts
constexpr int F_X = K0, F_H1 = H1, F_H2 = H2, F_MV = 2 * H3, F_DZMUV = 64, F_DZMV = 2 * H3, F_DZH2 = H2, F_DZH1 = H1;

struct Images {   // per-operand base pointers; tile t of an operand with F features starts at base + t * 128 * F
  float *x, *h1, *h2, *mv, *dzmuv, *dzmv, *dzh2, *dzh1;
};

struct GradArgs {
  const float* action;
  const float* old_logp;
  const float* target;
  const float* advantage;
  const float* action_var;
  float clip_lo, clip_hi, inv_rows;
  float* loss_part;   // [gridDim.x][2]
};

FLYB_D float* img_ptr(float* tile_base, int F, int r, int f) {
  return tile_base + (r >> 3) * (F * 8) + (f >> 2) * 32 + (r & 7) * 4;
}
FLYB_D void img_store16(float* p, const float (&y)[16]) {
#pragma unroll
  for (int g = 0; g < 4; ++g)
    *reinterpret_cast<float4*>(p + g * 32) = make_float4(y[4 * g], y[4 * g + 1], y[4 * g + 2], y[4 * g + 3]);
}
FLYB_D void img_load16(const float* p, float (&y)[16]) {
#pragma unroll
  for (int g = 0; g < 4; ++g) {
    const float4 v = *reinterpret_cast<const float4*>(p + g * 32);
    y[4 * g] = v.x; y[4 * g + 1] = v.y; y[4 * g + 2] = v.z; y[4 * g + 3] = v.w;
  }
}
// 16 fp32 values -> hi / lo planes of an A operand in tensor memory
FLYB_D void tmem_put16(uint32_t hi_addr, uint32_t lo_addr, const float (&y)[16]) {
  uint32_t hi[16], lo[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) split_fast(y[j], hi[j], lo[j]);
  tc::tmem_st16(hi_addr, hi);
  tc::tmem_st16(lo_addr, lo);
}
// y = ELU(D + bias) for `ncols` (multiple of 16) accumulator columns: A operand (hi/lo) + row image
FLYB_D void epi_elu_save(uint32_t d_addr, uint32_t ah_addr, uint32_t al_addr, const float* bias, float* img, int ncols) {
#pragma unroll 1
  for (int c = 0; c < ncols; c += 16) {
    uint32_t v[16];
    tc::tmem_ld16(d_addr + c, v);
    tc::tmem_wait_ld();
    float y[16];
    const float4* b4 = reinterpret_cast<const float4*>(bias + c);   // bias block offsets are multiples of 16 floats
#pragma unroll
    for (int q4 = 0; q4 < 4; ++q4) {
      const float4 b = b4[q4];
      y[4 * q4 + 0] = elu_fast(__uint_as_float(v[4 * q4 + 0]) + b.x);
      y[4 * q4 + 1] = elu_fast(__uint_as_float(v[4 * q4 + 1]) + b.y);
      y[4 * q4 + 2] = elu_fast(__uint_as_float(v[4 * q4 + 2]) + b.z);
      y[4 * q4 + 3] = elu_fast(__uint_as_float(v[4 * q4 + 3]) + b.w);
    }
    tmem_put16(ah_addr + c, al_addr + c, y);
    img_store16(img + (c >> 2) * 32, y);
  }
}

__global__ void __launch_bounds__(THREADS, 1)
ppo_fwd_bwd_tc_kernel(const __grid_constant__ FlybNetLayout net, const __grid_constant__ FwdArgs a,
                      const __grid_constant__ GradArgs ga, const __grid_constant__ Images im) {
  extern __shared__ unsigned char smem_raw[];
  const Smem s = carve(smem_raw, a.O, 0);
  __shared__ float s_std[APAD], s_inv_std[APAD], s_inv_var[APAD];
  // the tile's actions, row-major with the real stride A.  Wide shape: no room for 24 KB next to the 161 KB X planes —
  // the tile lives IN the hi plane, which is dead once the last layer-1 MMA has completed (staged after the trunk)
  __shared__ __align__(16) float s_act_static[WIDE ? 4 : TM * APAD];
  float* const s_act = WIDE ? s.xhi : s_act_static;
  __shared__ float s_half_log_det;
  __shared__ float s_vp[2 * TM];
  __shared__ float s_red[EPI_WARPS][2];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int O = a.O, A = a.A;
  const long long ntiles = (a.rows + TM - 1) / TM;

  // ---- one-time setup (barriers + the first loads go out first, see producer_kickoff)
  if (warp == PRODUCER_WARP) producer_kickoff(s, a, !WIDE);
  if (warp == MMA_WARP) tc::tmem_alloc(s.tmem_slot, 512);
  {
    const float* p = a.params;
    for (int i = threadIdx.x; i < B_TOTAL; i += THREADS) {
      float v = 0.0f;
      if (i < B_B2) v = p[net.off[1] + i];
      else if (i < B_B3PI) v = p[net.off[3] + i - B_B2];
      else if (i < B_B3V) v = p[net.off[5] + i - B_B3PI];
      else if (i < B_B4) v = p[net.off[9] + i - B_B3V];
      else if (i < B_W4V) v = (i - B_B4 < A) ? p[net.off[7] + i - B_B4] : 0.0f;
      else if (i < B_B4V) v = p[net.off[10] + i - B_W4V];
      else if (i == B_B4V) v = p[net.off[11]];
      s.bias[i] = v;
    }
    if (threadIdx.x < APAD) {
      const float var = threadIdx.x < A ? __ldg(ga.action_var + threadIdx.x) : 1.0f;
      const float sd = sqrtf(var);
      s_std[threadIdx.x] = sd;
      s_inv_std[threadIdx.x] = 1.0f / sd;
      s_inv_var[threadIdx.x] = 1.0f / (sd * sd);
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  if (threadIdx.x == 0) {
    float hl = 0.0f;
    for (int k = 0; k < A; ++k) hl += logf(s_std[k]);
    s_half_log_det = hl;
  }
  const uint32_t tb = *s.tmem_slot;
  // (wide: dZmu's lo plane cannot sit at [48,96) — [64,128) already holds dZv1's hi columns of A6 — it takes [128,176))
  constexpr uint32_t C_D4B = 256, C_A5H = 0, C_A5L = WIDE ? 128 : 32, C_D5 = 320, C_A6H = 0, C_A6L = 128, C_D6 = 384,
                     C_A7H = 0, C_A7L = 128, C_D7 = 256;

  if (warp == PRODUCER_WARP) {
    // =========================================================================== TMA producer (warp-uniform)
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
      if (kicked == 0 && !WIDE) {
        tc::mbar_wait_relaxed(s.obs_empty, obs_phase);   // the epilogue warps are done with the previous tile's rows
        obs_phase ^= 1u;
        if ((bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0 && leader) {
          tc::mbar_arrive_expect_tx(s.obs_full, bytes);
          tc::tma_load_1d(s.obs, src, bytes, s.obs_full);
        }
      }
      const unsigned char* g = a.wpack;
      for (int i = 0; i < ALL_STAGES; ++i, ++gs, g += STAGE) {
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
    }
  } else if (warp == MMA_WARP) {
    // =========================================================================== MMA issuer (warp-uniform)
    MmaState m{tc::elect_one() ? 1u : 0u, tc::smem_u32(s.ring), s.full, s.empty, 0u};
    const uint32_t xh = tc::smem_u32(s.xhi), xl = tc::smem_u32(s.xlo);
    uint32_t pa = 0u;
    auto wait_a = [&](int b) {
      tc::mbar_wait_relaxed(&s.bar_a[b], (pa >> b) & 1u);
      pa ^= 1u << b;
      tc::tc_fence_after();
    };
    int ti = 0;
    unsigned long long* mtrace = (lane == 0) ? a.trace : nullptr;   // slots [0,64): "operand ready" / "MMAs issued" pairs
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      wait_a(0);   // X staged (and the previous tile's accumulators drained)
      trace_mark(mtrace, ti, 64);
#pragma unroll 1
      for (int c = -2; c < NCH; ++c) {
        if (c >= 0) {
          const int b = c & 1;
          wait_a(b);
          mma_segment<H2, CH / 8, false>(m, tb + C_A2 + 128 * b, tb + C_A2 + 128 * b + CH, tb + C_D2, c > 0);
        }
        if (c + 2 < NCH) {
          const int b2 = (c + 2) & 1;
          mma_segment<CH, K0 / 8, true>(m, xh, xl, tb + C_D1 + CH * b2, false);
          umma_commit_if(m.leader, &s.bar_d[b2]);
        }
      }
      umma_commit_if(m.leader, &s.bar_d[0]);   // D2
      trace_mark(mtrace, ti, 64);
      wait_a(0);                               // A3
      trace_mark(mtrace, ti, 64);
      // D3 = [pi | v]: committed per half, so that the m1 epilogue runs while the value half's MMAs do.  The pi half
      // goes to bar_d[0] (its next phase, D4, is only issued after every thread has waited for this one), the value half
      // to bar_d[1] (next phase there: D7b, long after the value head has waited for it).
      mma_segment<H3, H2 / 8, false>(m, tb + C_A3H, tb + C_A3L, tb + C_D3, false);
      umma_commit_if(m.leader, &s.bar_d[0]);   // D3 pi
      mma_segment<H3, H2 / 8, false>(m, tb + C_A3H, tb + C_A3L, tb + C_D3 + H3, false);
      umma_commit_if(m.leader, &s.bar_d[1]);   // D3 v
      trace_mark(mtrace, ti, 64);
      wait_a(0);                               // A4
      trace_mark(mtrace, ti, 64);
      mma_segment<APAD, H3 / 8, false>(m, tb + C_A4H, tb + C_A4L, tb + C_D4B, false);
      umma_commit_if(m.leader, &s.bar_d[0]);   // D4
      trace_mark(mtrace, ti, 64);
      wait_a(0);                               // A5 = dZmu (and dZv1 in its A6 slots)
      trace_mark(mtrace, ti, 64);
      mma_segment<H3, APAD / 8, false>(m, tb + C_A5H, tb + C_A5L, tb + C_D5, false);
      umma_commit_if(m.leader, &s.bar_d[0]);   // D5 = dM1
      trace_mark(mtrace, ti, 64);
      wait_a(0);                               // A6 = [dZm1 | dZv1]
      trace_mark(mtrace, ti, 64);
      mma_segment<H2, 2 * H3 / 8, false>(m, tb + C_A6H, tb + C_A6L, tb + C_D6, false);
      umma_commit_if(m.leader, &s.bar_d[0]);   // D6 = dH2
      trace_mark(mtrace, ti, 64);
      wait_a(0);                               // A7 = dZh2
      trace_mark(mtrace, ti, 64);
      mma_segment<128, H2 / 8, false>(m, tb + C_A7H, tb + C_A7L, tb + C_D7, false);
      umma_commit_if(m.leader, &s.bar_d[0]);   // D7, h1 features [0,128)
      mma_segment<128, H2 / 8, false>(m, tb + C_A7H, tb + C_A7L, tb + C_D7 + 128, false);
      umma_commit_if(m.leader, &s.bar_d[1]);   // D7, h1 features [128,256)
      trace_mark(mtrace, ti, 64);
    }
  } else {
    // =========================================================================== epilogue warps
    const int q = warp & 3, h = warp >> 2;
    const uint32_t lane_base = (uint32_t)q * 32u;
    const int r = q * 32 + lane;
    const uint32_t t0 = tc::tmem_addr(tb, lane_base, 0);
    uint32_t pd = 0u, obs_phase = 0u;
    float surr_sum = 0.0f, sl1_sum = 0.0f;
    const float a_log_2pi = (float)((double)A * 1.8378770664093453);
    int ti = 64;
    unsigned long long* etrace = (threadIdx.x == 0) ? a.trace : nullptr;   // slots [64,128): accumulator seen / operand handed over
    auto wait_d = [&](int b) {
      tc::mbar_wait(&s.bar_d[b], (pd >> b) & 1u);
      pd ^= 1u << b;
      tc::tc_fence_after();
      trace_mark(etrace, ti, 128);
    };
    auto arrive_a = [&](int b) {
      tc::tmem_wait_st();
      tc::tc_fence_before();
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(&s.bar_a[b]);
      trace_mark(etrace, ti, 128);
    };
    for (long long tile = blockIdx.x; tile < ntiles; tile += gridDim.x) {
      const long long row0 = tile * TM;
      const int nrows = (int)min((long long)TM, a.rows - row0);
      const bool live = r < nrows;
      const long long row = row0 + r;
      float* x_t = im.x + (size_t)tile * TM * F_X;
      float* h1_t = im.h1 + (size_t)tile * TM * F_H1;
      float* h2_t = im.h2 + (size_t)tile * TM * F_H2;
      float* mv_t = im.mv + (size_t)tile * TM * F_MV;
      float* dzmuv_t = im.dzmuv + (size_t)tile * TM * F_DZMUV;
      float* dzmv_t = im.dzmv + (size_t)tile * TM * F_DZMV;
      float* dzh2_t = im.dzh2 + (size_t)tile * TM * F_DZH2;
      float* dzh1_t = im.dzh1 + (size_t)tile * TM * F_DZH1;

      const uint32_t bytes = (uint32_t)nrows * (uint32_t)O * 4u;
      const float* src = a.obs + (size_t)row0 * O;
      const bool by_tma = (bytes & 15u) == 0 && (reinterpret_cast<uintptr_t>(src) & 15u) == 0;
      if constexpr (WIDE) {
        // (the previous tile's policy head is done with the action tile that lives in the hi plane)
        epi_bar_sync();
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
      } else if (by_tma) {
        tc::mbar_wait(s.obs_full, obs_phase);
        obs_phase ^= 1u;
      } else {
        for (int i = threadIdx.x; i < nrows * O; i += EPI_THREADS) s.obs[i] = __ldg(src + i);
        tc::fence_proxy_async();
        epi_bar_sync();
      }
      // ---- X = split(obs row) in the UMMA K-major layout (layer-1 A operand) + row image (dW1's B operand)
      {
        const float* xr = WIDE ? nullptr : s.obs + r * O;
#pragma unroll 2
        for (int g = 0; g < K0 / 8; ++g) {
          const int kc = h * (K0 / 8) + g;
          uint32_t hi[4], lo[4];
          float xv[4], raw[4];
          if constexpr (WIDE) {
            const uint4 v = *x_chunk(s.xhi, kc, r);
            raw[0] = __uint_as_float(v.x); raw[1] = __uint_as_float(v.y);
            raw[2] = __uint_as_float(v.z); raw[3] = __uint_as_float(v.w);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const int c = kc * 4 + j;
            xv[j] = 0.0f;
            if (live && c < O) xv[j] = WIDE ? raw[j] : xr[c];
            split_fast(xv[j], hi[j], lo[j]);
          }
          *x_chunk(s.xhi, kc, r) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
          *x_chunk(s.xlo, kc, r) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
          *reinterpret_cast<float4*>(img_ptr(x_t, F_X, r, kc * 4)) = make_float4(xv[0], xv[1], xv[2], xv[3]);
        }
        tc::fence_proxy_async();
      }
      __syncwarp();
      if (lane == 0) tc::mbar_arrive(s.obs_empty);
      arrive_a(0);
      // the tile's actions -> shared memory, asynchronously (consumed by the policy head, behind a barrier)
      auto stage_actions = [&]() {
        const float* asrc = ga.action + (size_t)row0 * A;
        const int nfl = nrows * A;
        if ((nfl & 3) == 0 && (reinterpret_cast<uintptr_t>(asrc) & 15u) == 0) {
          for (int i = threadIdx.x * 4; i < nfl; i += EPI_THREADS * 4) cp_async16(s_act + i, asrc + i);
        } else {
          for (int i = threadIdx.x; i < nfl; i += EPI_THREADS) s_act[i] = __ldg(asrc + i);
        }
        cp_async_commit();
      };
      if constexpr (!WIDE) stage_actions();

      // per-row scalars of the heads, fetched now so that their latency is long gone when the heads need them
      const float tgt = live ? __ldg(ga.target + row) : 0.0f;
      const float old_lp = live ? __ldg(ga.old_logp + row) : 0.0f;
      const float adv = live ? __ldg(ga.advantage + row) : 0.0f;

      // ---- trunk
#pragma unroll 1
      for (int c = 0; c < NCH; ++c) {
        const int b = c & 1;
        wait_d(b);
        epi_elu_save(t0 + C_D1 + CH * b + 32 * h, t0 + C_A2 + 128 * b + 32 * h, t0 + C_A2 + 128 * b + CH + 32 * h,
                     s.bias + B_B1 + c * CH + 32 * h, img_ptr(h1_t, F_H1, r, c * CH + 32 * h), 32);
        arrive_a(b);
      }
      if constexpr (WIDE) stage_actions();   // chunk 3's accumulator was seen: every layer-1 MMA has completed
      wait_d(0);   // D2
      epi_elu_save(t0 + C_D2 + 64 * h, t0 + C_A3H + 64 * h, t0 + C_A3L + 64 * h, s.bias + B_B2 + 64 * h,
                   img_ptr(h2_t, F_H2, r, 64 * h), 64);
      arrive_a(0);
      wait_d(0);   // D3 pi
      epi_elu_save(t0 + C_D3 + 32 * h, t0 + C_A4H + 32 * h, t0 + C_A4L + 32 * h, s.bias + B_B3PI + 32 * h,
                   img_ptr(mv_t, F_MV, r, 32 * h), 32);
      arrive_a(0);   // -> D4

      // ---- value head: v1 = ELU(D3v + b3v) (32 columns per thread), value, dL/dv, dZv1
      float dval;
      wait_d(1);   // D3 v
      {
        float v1[32];
        float part = 0.0f;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const int c0 = h * 32 + g * 16;
          uint32_t v[16];
          tc::tmem_ld16(t0 + C_D3 + H3 + c0, v);
          tc::tmem_wait_ld();
          float y[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            y[j] = elu_fast(__uint_as_float(v[j]) + s.bias[B_B3V + c0 + j]);
            v1[g * 16 + j] = y[j];
            part = fmaf(y[j], s.bias[B_W4V + c0 + j], part);
          }
          img_store16(img_ptr(mv_t, F_MV, r, H3 + c0), y);
        }
        s_vp[h * TM + r] = part;
        cp_async_wait<0>();   // this thread's share of the action tile has landed; the barrier publishes all of it
        tc::tc_fence_before();
        epi_bar_sync();
        const float value = s_vp[r] + s_vp[TM + r] + s.bias[B_B4V];
        const float z = value - tgt;
        const float az = fabsf(z);
        dval = live ? ga.inv_rows * (az < 1.0f ? z : (z > 0.0f ? 1.0f : -1.0f)) : 0.0f;   // smooth_l1, beta = 1
        if (h == 0 && live) sl1_sum += az < 1.0f ? 0.5f * z * z : az - 0.5f;
#pragma unroll
        for (int g = 0; g < 2; ++g) {
          const int c0 = h * 32 + g * 16;
          float dz[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) dz[j] = dval * s.bias[B_W4V + c0 + j] * elu_grad_from_out(v1[g * 16 + j]);
          tmem_put16(t0 + C_A6H + H3 + c0, t0 + C_A6L + H3 + c0, dz);
          img_store16(img_ptr(dzmv_t, F_DZMV, r, H3 + c0), dz);
        }
      }
      // ---- policy head: loss terms of this row (ppo.py:184-194), dL/dmu -> dZmu
      {
        float act[APAD];
#pragma unroll
        for (int k = 0; k < APAD; ++k) act[k] = (live && k < A) ? s_act[r * A + k] : 0.0f;
        wait_d(0);   // D4
        float mu[APAD];
#pragma unroll
        for (int g = 0; g < APAD / 16; ++g) {
          uint32_t v[16];
          tc::tmem_ld16(t0 + C_D4B + g * 16, v);
          tc::tmem_wait_ld();
#pragma unroll
          for (int j = 0; j < 16; ++j) mu[g * 16 + j] = elu_fast(__uint_as_float(v[j]) + s.bias[B_B4 + g * 16 + j]);
        }
        float msum = 0.0f;
#pragma unroll
        for (int k = 0; k < APAD; ++k)
          if (k < A) {
            const float diff = (act[k] - mu[k]) * s_inv_std[k];
            msum += diff * diff;
          }
        const float logp = -0.5f * (a_log_2pi + msum) - s_half_log_det;
        const float ratio = expf(logp - old_lp);
        const float s1 = ratio * adv;
        const float s2 = fminf(fmaxf(ratio, ga.clip_lo), ga.clip_hi) * adv;
        const bool in_range = ratio >= ga.clip_lo && ratio <= ga.clip_hi;
        float gsel;   // d min(s1, s2) / d ratio with torch.minimum's tie rule (half to each side)
        if (s1 < s2) gsel = adv;
        else if (s1 > s2) gsel = in_range ? adv : 0.0f;
        else gsel = 0.5f * adv + (in_range ? 0.5f * adv : 0.0f);
        const float dlogp = live ? -ga.inv_rows * gsel * ratio : 0.0f;
        if (h == 0 && live) surr_sum += fminf(s1, s2);
        float dz[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          // columns 16h .. 16h+15 (static register indexing: select after computing both candidates)
          const float mu_k = h ? mu[16 + j] : mu[j];
          const float a_k = h ? act[16 + j] : act[j];
          const int k = 16 * h + j;
          const float dmu = dlogp * ((a_k - mu_k) * s_inv_var[k]);
          dz[j] = (k < A) ? dmu * elu_grad_from_out(mu_k) : 0.0f;   // trailing ELU on mu (ppo.py:30)
        }
        tmem_put16(t0 + C_A5H + 16 * h, t0 + C_A5L + 16 * h, dz);   // A3 columns, dead since D3
        // image [dZmu (APAD) | dZval | 0 ...] in 16-feature blocks; half h writes blocks h and h + 2.  Blocks below
        // APAD / 16 are dZmu (and go to the A5 operand as well), block APAD / 16 opens with dZval, the rest is zero.
        img_store16(img_ptr(dzmuv_t, F_DZMUV, r, 16 * h), dz);
        float tail[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) tail[j] = 0.0f;
        if constexpr (APAD / 16 == 2) {
          if (h == 0) tail[0] = dval;   // feature 32
          img_store16(img_ptr(dzmuv_t, F_DZMUV, r, 32 + 16 * h), tail);
        } else {
          if (h == 0) {                 // block 2: dZmu columns 32 .. 47
            float dz2[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
              const int k = 32 + j;
              const float mu_k = mu[APAD - 16 + j];
              const float dmu = dlogp * ((act[APAD - 16 + j] - mu_k) * s_inv_var[k]);
              dz2[j] = (k < A) ? dmu * elu_grad_from_out(mu_k) : 0.0f;
            }
            tmem_put16(t0 + C_A5H + 32, t0 + C_A5L + 32, dz2);
            img_store16(img_ptr(dzmuv_t, F_DZMUV, r, 32), dz2);
          } else {                      // block 3: [dZval | 0 ...], feature 48
            tail[0] = dval;
            img_store16(img_ptr(dzmuv_t, F_DZMUV, r, 48), tail);
          }
        }
        arrive_a(0);   // -> D5
      }
      // ---- dZm1 = (dZmu * W4) .* ELU'(m1): 32 columns per thread
      wait_d(0);   // D5
#pragma unroll 1
      for (int g = 0; g < 2; ++g) {
        const int c0 = 32 * h + 16 * g;
        uint32_t d[16], mh[16], ml[16];
        tc::tmem_ld16(t0 + C_D5 + c0, d);
        tc::tmem_ld16(t0 + C_A4H + c0, mh);
        tc::tmem_ld16(t0 + C_A4L + c0, ml);
        tc::tmem_wait_ld();
        float dz[16];
#pragma unroll
        for (int j = 0; j < 16; ++j)
          dz[j] = __uint_as_float(d[j]) * elu_grad_from_out(__uint_as_float(mh[j]) + __uint_as_float(ml[j]));
        tmem_put16(t0 + C_A6H + c0, t0 + C_A6L + c0, dz);
        img_store16(img_ptr(dzmv_t, F_DZMV, r, c0), dz);
      }
      arrive_a(0);   // -> D6
      // ---- dZh2 = (dZm1 * W3pi + dZv1 * W3v) .* ELU'(h2): 64 columns per thread; h2 is re-read from its row image
      //      (written by this same thread) BEFORE the accumulator is waited for, so the loads fly during the MMAs
      {
        float y2[4][16];
#pragma unroll
        for (int g = 0; g < 4; ++g) img_load16(img_ptr(h2_t, F_H2, r, 64 * h + 16 * g), y2[g]);
        wait_d(0);   // D6
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          const int c0 = 64 * h + 16 * g;
          uint32_t d[16];
          tc::tmem_ld16(t0 + C_D6 + c0, d);
          tc::tmem_wait_ld();
          float dz[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) dz[j] = __uint_as_float(d[j]) * elu_grad_from_out(y2[g][j]);
          tmem_put16(t0 + C_A7H + c0, t0 + C_A7L + c0, dz);
          img_store16(img_ptr(dzh2_t, F_DZH2, r, c0), dz);
        }
      }
      arrive_a(0);   // -> D7
      // ---- dZh1 = (dZh2 * W2) .* ELU'(h1): chunks c of 64 features, 32 per thread; image only.  The h1 rows of
      //      chunk c + 1 are fetched while chunk c is processed.
      {
        float yb[2][16];
#pragma unroll
        for (int g = 0; g < 2; ++g) img_load16(img_ptr(h1_t, F_H1, r, 32 * h + 16 * g), yb[g]);
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
          if (c == 0) wait_d(0);
          if (c == 2) wait_d(1);
          float yn[2][16];
          if (c + 1 < NCH) {
#pragma unroll
            for (int g = 0; g < 2; ++g) img_load16(img_ptr(h1_t, F_H1, r, (c + 1) * CH + 32 * h + 16 * g), yn[g]);
          }
#pragma unroll
          for (int g = 0; g < 2; ++g) {
            const int c0 = c * CH + 32 * h + 16 * g;
            uint32_t d[16];
            tc::tmem_ld16(t0 + C_D7 + c0, d);
            tc::tmem_wait_ld();
            float dz[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) dz[j] = __uint_as_float(d[j]) * elu_grad_from_out(yb[g][j]);
            img_store16(img_ptr(dzh1_t, F_DZH1, r, c0), dz);
          }
          if (c + 1 < NCH) {
#pragma unroll
            for (int g = 0; g < 2; ++g)
#pragma unroll
              for (int j = 0; j < 16; ++j) yb[g][j] = yn[g][j];
          }
        }
      }
      tc::tc_fence_before();
    }
    // ---- loss partials of this CTA
    surr_sum = warp_sum(surr_sum);
    sl1_sum = warp_sum(sl1_sum);
    if (lane == 0) { s_red[warp][0] = surr_sum; s_red[warp][1] = sl1_sum; }
    epi_bar_sync();
    if (threadIdx.x == 0) {
      float x = 0.0f, y = 0.0f;
      for (int i = 0; i < EPI_WARPS; ++i) { x += s_red[i][0]; y += s_red[i][1]; }
      ga.loss_part[blockIdx.x * 2 + 0] = x;
      ga.loss_part[blockIdx.x * 2 + 1] = y;
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (warp == MMA_WARP) tc::tmem_dealloc(tb, 512);
}

// ================================================================================================ weight gradients
// One "pair" = one layer's dW = dZ^T X:  A = dZ image (MA features -> D lanes), B = X image (NB features -> columns).
static_assert(APAD / 16 == 2 || APAD / 16 == 3, "the policy head handles two or three 16-column blocks of dZmu");
constexpr int RED_GROUPS_C = 4;                // interleaved partial groups of the gradient reduction (ppo_reduce_tc_kernel)
constexpr int DW_KG = 2;                       // k-steps (of 8 rows) per pipeline stage
constexpr int DW_RAW_SLOTS = 4, DW_OP_SLOTS = 2;
constexpr int DW_TRANS_WARPS = 16, DW_THREADS = (2 + DW_TRANS_WARPS) * 32;
FLYB_D void dw_bar_sync() { asm volatile("bar.sync 1, %0;" ::"n"(DW_TRANS_WARPS * 32) : "memory"); }
constexpr int DW_MAX_FEATS = WIDE ? 416 : 384; // MA + NB of the widest pair (dZh2 128 + h1 256; wide: dZh1 256 + X 160)
constexpr uint32_t DW_RAW_SLOT = DW_KG * DW_MAX_FEATS * 32;               // 24576 B
// operand stage: per k-step and operand [hi plane | lo plane], plane = 2 K-chunks x (F * 16 + 16) bytes
__host__ __device__ constexpr uint32_t dw_plane_bytes(int F) { return 2u * ((uint32_t)F * 16u + 16u); }
constexpr uint32_t DW_OP_SLOT =   // the pair with the largest planes, + over-read slack
    DW_KG * 2 * (WIDE ? dw_plane_bytes(256) + dw_plane_bytes(160) : dw_plane_bytes(128) + dw_plane_bytes(256)) + 4096;

struct DwPair {
  const float* a_img;   // dZ row images
  const float* b_img;   // X row images
  int MA, NB;           // features
  int mhalves;          // MA / 128 (1 or 2); MA = 64 runs as one M = 128 MMA whose lanes >= 64 are ignored
  uint32_t d_col[2];    // accumulator columns per M half
  uint32_t bias_col[2]; // 16-column bias accumulators
};
struct DwRole {
  DwPair pair[2];
};
// flyb_ppo_step: the rest of the optimiser step rides in the tail of this kernel (cooperative launch, every CTA
// resident): grid barrier, fixed-order sum of the per-CTA partials (the same association as ppo_reduce_tc_kernel, so
// the gradient is bit-identical), per-CTA squared norms, grid barrier, clip coefficient, Adam + in-place refresh of the
// tensor-core weight stream.  Two launches per optimiser step instead of four.
struct DwTail {
  int enabled;
  int off[12], size[12];
  int total, nparts, nloss;
  float inv_rows;
  const float* loss_part;
  float* loss;
  float* grad;
  float* sq_part;           // [gridDim.x]
  long long* step;
  float *params, *exp_avg, *exp_avg_sq, *grad_norm_out;
  AdamArgs adam;
  float* wpack;
  const int4* wpack_index;
  int has_peers;            // data-parallel: the gradient exchange over NVLink peer memory happens in the tail too
  PeerView pv;
};
__device__ unsigned g_dw_grid_sync[2];   // grid barrier state of the cooperative launch (count, generation)

struct DwArgs {
  DwTail tail;
  DwRole role[2];
  long long ntiles;
  int ctas_per_role;
  float* dw_part;       // [ctas_per_role][net.total]
  int net_total;
  int off[12];
  int A;
  int dbg;              // timing experiments only (FLYB_DW_DBG): 1 skip transposition, 2 skip MMAs, 4 skip flush
  unsigned long long* trace;   // debug (flyb_debug_set_trace): CTA 0's clock64 marks, [0,64) producer: slot free seen,
                               // [64,128) transposer warp 0: stage in hand / handed on, [128,190) MMA warp: operands seen
};

template <bool TAIL>
__global__ void __launch_bounds__(DW_THREADS, 1)
ppo_dw_tc_kernel(const __grid_constant__ DwArgs a) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t pad = (1024u - (tc::smem_u32(smem_raw) & 1023u)) & 1023u;
  unsigned char* base = smem_raw + pad;
  unsigned char* raw = base;                                       // DW_RAW_SLOTS x DW_RAW_SLOT
  unsigned char* ops = raw + DW_RAW_SLOTS * DW_RAW_SLOT;           // DW_OP_SLOTS x DW_OP_SLOT
  float* ones = reinterpret_cast<float*>(ops + DW_OP_SLOTS * DW_OP_SLOT);   // K-major [2 chunks][16 feats][4 rows]
  float* bias_s = ones + 128;                                              // [2 pairs][256 features]
  uint64_t* bars = reinterpret_cast<uint64_t*>(bias_s + 512);
  uint64_t* raw_full = bars;
  uint64_t* raw_empty = raw_full + DW_RAW_SLOTS;
  uint64_t* op_full = raw_empty + DW_RAW_SLOTS;
  uint64_t* op_empty = op_full + DW_OP_SLOTS;
  uint64_t* acc_done = op_empty + DW_OP_SLOTS;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_done + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int role_id = blockIdx.x & 1, ci = blockIdx.x >> 1;
  const DwRole& role = a.role[role_id];
  // Work is dealt out by STAGE, not by tile: stage g = (tile, pair, k-step group) in a linear order, CTA ci of a role
  // takes g = ci, ci + cpr, ci + 2 cpr, ...  (4096 stages over 74 CTAs: 55 or 56 each, where whole tiles gave 48 or 64).
  // Any subset of stages may be summed in any CTA — a pair's accumulator is a sum over rows — and since cpr <= ntiles
  // every CTA meets at least 16 consecutive residues of g mod 16, i.e. both pairs.
  constexpr int stages_per_tile = 2 * (16 / DW_KG);                // both pairs
  static_assert(stages_per_tile == 16, "stage index -> (tile, pair, k-step group) by shifts");
  const long long total_stages = a.ntiles * stages_per_tile;
  const long long nstages = (total_stages - ci + a.ctas_per_role - 1) / a.ctas_per_role;

  if (threadIdx.x == 0) {
    for (int i = 0; i < DW_RAW_SLOTS; ++i) {
      tc::mbar_init(&raw_full[i], 1);
      tc::mbar_init(&raw_empty[i], DW_TRANS_WARPS);
    }
    for (int i = 0; i < DW_OP_SLOTS; ++i) {
      tc::mbar_init(&op_full[i], DW_TRANS_WARPS);
      tc::mbar_init(&op_empty[i], 1);
    }
    tc::mbar_init(acc_done, 1);
    tc::fence_mbar_init();
  }
  for (int i = threadIdx.x; i < 128; i += DW_THREADS) ones[i] = ((i & 63) < 4) ? 1.0f : 0.0f;   // feature 0, all 8 rows
  tc::fence_proxy_async();
  if (warp == 1) tc::tmem_alloc(tmem_slot, 512);
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tb = *tmem_slot;

  if (warp == 0) {
    // =========================================================================== TMA producer
    if (tc::elect_one()) {
      uint32_t slot = 0, phase = 0;
      // Newest tiles first: the chain kernel has just written 153 MB of row images through a 126 MB L2.  Reading them
      // back oldest-first would miss on everything (each miss evicts the next-oldest lines, the ones wanted next);
      // youngest-first finds most of them still resident.  CTA ci owns tiles ntiles-1-ci, ntiles-1-ci-cpr, ...
      long long g = ci;
      for (long long st = 0; st < nstages; ++st, g += a.ctas_per_role) {
        const long long tile = a.ntiles - 1 - (g >> 4);
        const int within = (int)(g & 15);
        const int pi = within >= (16 / DW_KG) ? 1 : 0, kg = within - pi * (16 / DW_KG);
        const DwPair& p = role.pair[pi];
        tc::mbar_wait_relaxed(&raw_empty[slot], phase ^ 1u);
        if (a.trace != nullptr && blockIdx.x == 0 && st < 64) a.trace[st] = clock64();
        const uint32_t a_bytes = (uint32_t)DW_KG * p.MA * 32u, b_bytes = (uint32_t)DW_KG * p.NB * 32u;
        tc::mbar_arrive_expect_tx(&raw_full[slot], a_bytes + b_bytes);
        unsigned char* dst = raw + slot * DW_RAW_SLOT;
        const float* ga = p.a_img + (size_t)tile * TM * p.MA + (size_t)kg * DW_KG * p.MA * 8;
        const float* gb = p.b_img + (size_t)tile * TM * p.NB + (size_t)kg * DW_KG * p.NB * 8;
        tc::tma_load_1d(dst, ga, a_bytes, &raw_full[slot]);
        tc::tma_load_1d(dst + a_bytes, gb, b_bytes, &raw_full[slot]);
        if (++slot == DW_RAW_SLOTS) { slot = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // =========================================================================== MMA issuer
    const uint32_t leader = tc::elect_one() ? 1u : 0u;
    uint32_t slot = 0, phase = 0;
    bool seen0 = false, seen1 = false;   // has this CTA already accumulated into the pair?
    long long g = ci;
    for (long long st = 0; st < nstages; ++st, g += a.ctas_per_role) {
      const int pi = (int)(g & 15) >= (16 / DW_KG) ? 1 : 0;
      const DwPair& p = role.pair[pi];
      tc::mbar_wait(&op_full[slot], phase);
      tc::tc_fence_after();
      if (a.trace != nullptr && blockIdx.x == 0 && lane == 0 && st < 62) a.trace[128 + st] = clock64();
      const uint32_t pa_b = dw_plane_bytes(p.MA), pb_b = dw_plane_bytes(p.NB);
      const uint32_t a_lbo = (uint32_t)p.MA * 16u + 16u, b_lbo = (uint32_t)p.NB * 16u + 16u;
      const uint32_t sbase = tc::smem_u32(ops) + slot * DW_OP_SLOT;
      const uint32_t idesc = tc::umma_idesc_tf32(p.NB);
      const bool first_stage_of_pair = pi ? !seen1 : !seen0;
      if (pi) seen1 = true; else seen0 = true;
#pragma unroll
      for (int j = 0; j < DW_KG; ++j) {
        const uint32_t a_hi = sbase + (uint32_t)j * 2u * (pa_b + pb_b), a_lo = a_hi + pa_b;
        const uint32_t b_hi = a_lo + pa_b, b_lo = b_hi + pb_b;
        const uint64_t bd_hi = tc::umma_smem_desc(b_hi, b_lbo, 128u), bd_lo = tc::umma_smem_desc(b_lo, b_lbo, 128u);
        const uint32_t acc0 = (first_stage_of_pair && j == 0) ? 0u : 1u;
        for (int mh = 0; mh < ((a.dbg & 2) ? 0 : p.mhalves); ++mh) {
          const uint64_t ad_hi = tc::umma_smem_desc(a_hi + (uint32_t)mh * 2048u, a_lbo, 128u);
          const uint64_t ad_lo = tc::umma_smem_desc(a_lo + (uint32_t)mh * 2048u, a_lbo, 128u);
          umma_ss_if(leader, tb + p.d_col[mh], ad_lo, bd_hi, idesc, acc0);
          umma_ss_if(leader, tb + p.d_col[mh], ad_hi, bd_lo, idesc, 1u);
          umma_ss_if(leader, tb + p.d_col[mh], ad_hi, bd_hi, idesc, 1u);
        }
      }
      umma_commit_if(leader, &op_empty[slot]);
      if (++slot == DW_OP_SLOTS) { slot = 0; phase ^= 1u; }
    }
    if (!(seen0 && seen1)) __trap();   // the host picks cpr so that every CTA meets both pairs (carve_ws)
    umma_commit_if(leader, acc_done);
  } else {
    // =========================================================================== transposers, then flush
    const int tw = warp - 2;
    const int tid = tw * 32 + lane;   // 0..255
    const bool rot = (lane & 16) != 0;
    const uint32_t eoff = rot ? 32u : 0u;
    static_assert(DW_TRANS_WARPS * 32 == 512, "a thread meets ONE (feature group, row-in-k-step) slot per pair");
    float4 bsum[2];   // [pair]: column sums of the dZ chunks this thread transposes
#pragma unroll
    for (int u = 0; u < 2; ++u) bsum[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (int i = tid; i < 2 * 256; i += DW_TRANS_WARPS * 32) bias_s[i] = 0.0f;
    dw_bar_sync();
    // Per-thread chunk plan, fixed for the whole kernel: one float4 chunk = (k-step j, feature group g, row r8) of the
    // raw stage [A: DW_KG x (MA/4) x 8 x 4 | B: DW_KG x (NB/4) x 8 x 4].  An operand with F features has 4 F chunks per
    // stage; a thread owns chunk `tid` and, where 4 F > 512 (F = 256: all threads, F = 160: the first 128), `tid + 512`.
    static_assert(DW_KG == 2, "the k-step of a chunk is found by one compare");
    int t_src[2][2], t_n[2][2];
    uint32_t t_dst[2][2][2], t_plane[2][2];
#pragma unroll
    for (int pi = 0; pi < 2; ++pi) {
      const DwPair& p = role.pair[pi];
      const uint32_t pa_b = dw_plane_bytes(p.MA), pb_b = dw_plane_bytes(p.NB);
      const uint32_t kstep_bytes = 2u * (pa_b + pb_b);   // one k-step block [A hi | A lo | B hi | B lo] in the operand stage
#pragma unroll
      for (int op = 0; op < 2; ++op) {
        const int F = op ? p.NB : p.MA;
        const int per_k = 2 * F, total = DW_KG * per_k;
        t_src[pi][op] = (op ? DW_KG * 2 * p.MA : 0) + tid;
        t_n[pi][op] = (a.dbg & 1) ? 0 : (tid < total ? (tid + DW_TRANS_WARPS * 32 < total ? 2 : 1) : 0);
        t_plane[pi][op] = op ? pb_b : pa_b;
#pragma unroll
        for (int it = 0; it < 2; ++it) {
          const int q = tid + it * DW_TRANS_WARPS * 32;
          const int j = q >= per_k ? 1 : 0;
          const int rem = q - j * per_k;
          const int g = rem >> 3, r8 = rem & 7;
          // K-major operand: plane[(chunk = r8 >> 2)][feature][r8 & 3], chunk stride F * 16 + 16 bytes
          t_dst[pi][op][it] = (op ? 2u * pa_b : 0u) + (uint32_t)j * kstep_bytes +
                              (uint32_t)(r8 >> 2) * ((uint32_t)F * 16u + 16u) + (uint32_t)g * 64u + (uint32_t)(r8 & 3) * 4u;
        }
      }
    }
    // Lanes 16..31 store their four features in the order 2,3,0,1: with the 16-byte feature pitch of the core matrix
    // this spreads a warp's 4-byte stores over all 32 banks.
    const uint32_t eo0 = eoff, eo1 = eoff + 16u, eo2 = eoff ^ 32u, eo3 = (eoff ^ 32u) + 16u;
    uint32_t rslot = 0, rphase = 0, oslot = 0, ophase = 0;
    long long g = ci;
    for (long long st = 0; st < nstages; ++st, g += a.ctas_per_role) {
      const int pi = (int)(g & 15) >= (16 / DW_KG) ? 1 : 0;
      tc::mbar_wait(&raw_full[rslot], rphase);
      if (a.trace != nullptr && blockIdx.x == 0 && tid == 0 && st < 21) a.trace[64 + 3 * st] = clock64();
      tc::mbar_wait(&op_empty[oslot], ophase ^ 1u);
      tc::tc_fence_after();
      if (a.trace != nullptr && blockIdx.x == 0 && tid == 0 && st < 21) a.trace[65 + 3 * st] = clock64();
      const float4* src4 = reinterpret_cast<const float4*>(raw + rslot * DW_RAW_SLOT);
      unsigned char* odst = ops + oslot * DW_OP_SLOT;
#pragma unroll
      for (int op = 0; op < 2; ++op) {
        const int src = pi ? t_src[1][op] : t_src[0][op];
        const int n = pi ? t_n[1][op] : t_n[0][op];
        const uint32_t plane = pi ? t_plane[1][op] : t_plane[0][op];
#pragma unroll
        for (int it = 0; it < 2; ++it) {
          if (it < n) {
            const float4 v = src4[src + it * (DW_TRANS_WARPS * 32)];
            if (op == 0) {   // bias gradient = column sums of dZ: exact fp32 adds, reduced over the 8 row lanes at the end
              float4& acc = pi == 0 ? bsum[0] : bsum[1];
              acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
            }
            unsigned char* kb = odst + (pi ? t_dst[1][op][it] : t_dst[0][op][it]);
            uint32_t hi, lo;
            split_fast(rot ? v.z : v.x, hi, lo);
            *reinterpret_cast<uint32_t*>(kb + eo0) = hi;
            *reinterpret_cast<uint32_t*>(kb + plane + eo0) = lo;
            split_fast(rot ? v.w : v.y, hi, lo);
            *reinterpret_cast<uint32_t*>(kb + eo1) = hi;
            *reinterpret_cast<uint32_t*>(kb + plane + eo1) = lo;
            split_fast(rot ? v.x : v.z, hi, lo);
            *reinterpret_cast<uint32_t*>(kb + eo2) = hi;
            *reinterpret_cast<uint32_t*>(kb + plane + eo2) = lo;
            split_fast(rot ? v.y : v.w, hi, lo);
            *reinterpret_cast<uint32_t*>(kb + eo3) = hi;
            *reinterpret_cast<uint32_t*>(kb + plane + eo3) = lo;
          }
        }
      }
      tc::fence_proxy_async();
      __syncwarp();
      if (lane == 0) {
        tc::mbar_arrive(&raw_empty[rslot]);
        tc::mbar_arrive(&op_full[oslot]);
      }
      if (a.trace != nullptr && blockIdx.x == 0 && tid == 0 && st < 21) a.trace[66 + 3 * st] = clock64();
      if (++rslot == DW_RAW_SLOTS) { rslot = 0; rphase ^= 1u; }
      if (++oslot == DW_OP_SLOTS) { oslot = 0; ophase ^= 1u; }
    }
    // ---- bias gradients: sum the 8 row lanes of every feature group, then one shared-memory add per contributor
    //      (at most two per slot, so the order cannot change the result)
#pragma unroll
    for (int u = 0; u < 2; ++u) {
#pragma unroll
      for (int o2 = 1; o2 < 8; o2 <<= 1) {
        bsum[u].x += __shfl_xor_sync(0xffffffffu, bsum[u].x, o2);
        bsum[u].y += __shfl_xor_sync(0xffffffffu, bsum[u].y, o2);
        bsum[u].z += __shfl_xor_sync(0xffffffffu, bsum[u].z, o2);
        bsum[u].w += __shfl_xor_sync(0xffffffffu, bsum[u].w, o2);
      }
    }
    if ((lane & 7) == 0) {
#pragma unroll
      for (int pi = 0; pi < 2; ++pi) {
        const int per_k = role.pair[pi].MA * 2;
        if (tid < DW_KG * per_k) {   // threads beyond the stage's chunk count never held a chunk of this pair
          float* b = bias_s + pi * 256 + 4 * ((tid & (per_k - 1)) >> 3);
          atomicAdd(b + 0, bsum[pi].x); atomicAdd(b + 1, bsum[pi].y);
          atomicAdd(b + 2, bsum[pi].z); atomicAdd(b + 3, bsum[pi].w);
        }
      }
    }
    dw_bar_sync();
    // ---- flush the accumulators of this CTA to its partial: lane = output row (feature of dZ)
    tc::mbar_wait(acc_done, 0u);
    tc::tc_fence_after();
    {
      const int q = warp & 3;                 // TMEM lane quarter this warp may read
      const int hsel = tw >> 2;               // four warps per quarter: 16-column blocks are dealt round-robin
      const int m = q * 32 + lane;
      const uint32_t t0 = tc::tmem_addr(tb, (uint32_t)q * 32u, 0);
      float* out = a.dw_part + (size_t)ci * a.net_total;
      auto dump = [&](uint32_t col0, int ncols, float* dst_row, int col_lo, int col_hi) {
        // writes columns [col_lo, col_hi) of the accumulator row to dst_row[col - col_lo]; this warp takes every
        // other 16-column block
        for (int c = 16 * hsel; c < ncols; c += 16 * (DW_TRANS_WARPS / 4)) {
          uint32_t v[16];
          tc::tmem_ld16(t0 + col0 + c, v);
          tc::tmem_wait_ld();
          if (dst_row != nullptr && c >= col_lo && c < col_hi) {   // windows are multiples of 16 columns, rows 16-byte aligned
            float4* d4 = reinterpret_cast<float4*>(dst_row + (c - col_lo));
#pragma unroll
            for (int j = 0; j < 4; ++j)
              d4[j] = make_float4(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1]), __uint_as_float(v[4 * j + 2]),
                                  __uint_as_float(v[4 * j + 3]));
          }
        }
      };
      auto bias_val = [&](int pi, int feature) { return bias_s[pi * 256 + feature]; };
      if (a.dbg & 4) {
      } else if (role_id == 0) {
        {   // layer 2: dW2[m][n] (128 x 256), db2[m]
          const DwPair& p = role.pair[0];
          dump(p.d_col[0], 256, out + a.off[2] + (size_t)m * 256, 0, 256);
          const float b = bias_val(0, m);
          if (hsel == 0) out[a.off[3] + m] = b;
        }
        {   // layer 4: lanes [0,A) = to_mean.2 rows over m1 columns [0,64); lane APAD = to_value.2 over v1 columns [64,128)
          const DwPair& p = role.pair[1];
          float* dst = nullptr;
          int lo = 0, hi = 0;
          if (m < a.A) { dst = out + a.off[6] + (size_t)m * 64; lo = 0; hi = 64; }
          else if (m == APAD) { dst = out + a.off[10]; lo = 64; hi = 128; }
          dump(p.d_col[0], 128, dst, lo, hi);
          const float b = bias_val(1, m & 63);
          if (hsel == 0) {
            if (m < a.A) out[a.off[7] + m] = b;
            else if (m == APAD) out[a.off[11]] = b;
          }
        }
      } else {
        {   // layer 1: dW1[128 mh + m][n] (256 x K0), db1
          const DwPair& p = role.pair[0];
          for (int mh = 0; mh < 2; ++mh) {
            // K0 columns in blocks of 16, dealt round-robin to the four warps of the quarter
            dump(p.d_col[mh], K0, out + a.off[0] + (size_t)(128 * mh + m) * K0, 0, K0);
            const float b = bias_val(0, 128 * mh + m);
            if (hsel == 0) out[a.off[1] + 128 * mh + m] = b;
          }
        }
        {   // layer 3: lanes [0,64) = to_mean.0 rows, lanes [64,128) = to_value.0 rows, 128 columns (h2)
          const DwPair& p = role.pair[1];
          float* dst = (m < 64) ? out + a.off[4] + (size_t)m * 128 : out + a.off[8] + (size_t)(m - 64) * 128;
          dump(p.d_col[0], 128, dst, 0, 128);
          const float b = bias_val(1, m);
          if (hsel == 0) {
            if (m < 64) out[a.off[5] + m] = b;
            else out[a.off[9] + m - 64] = b;
          }
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 1) tc::tmem_dealloc(tb, 512);

  if constexpr (TAIL) {
    // ======================================================================= reduction + clip + Adam (flyb_ppo_step)
    const DwTail& t = a.tail;
    __shared__ float s_red[DW_THREADS / 32];
    __shared__ AdamCoef s_c;
    constexpr int TB = 512;                     // entries per block of work
    const int tid = threadIdx.x;
    uint32_t epoch = 0;
    if (t.has_peers) {
      epoch = *(volatile const uint32_t*)t.pv.epoch;   // exchanges completed before this one (bumped at the very end)
      if (blockIdx.x == 0 && tid == 0) t.pv.stamp[0] = global_ns();
    }
    __threadfence();                            // this CTA's partial is written
    grid_barrier(g_dw_grid_sync, gridDim.x);
    float sq = 0.0f;
    for (int base = blockIdx.x * TB; base < t.total; base += gridDim.x * TB) {
      const int i = base + tid;
      if (tid < TB && i < t.total) {
        bool valid = false;
#pragma unroll
        for (int sg = 0; sg < 12; ++sg) valid = valid || (i >= t.off[sg] && i < t.off[sg] + t.size[sg]);
        float acc[RED_GROUPS_C] = {0.f, 0.f, 0.f, 0.f};   // group g sums partials g, g + 4, ... in order
        if (valid) {
          const float* src = a.dw_part + i;
          int p = 0;
          for (; p + 16 <= t.nparts; p += 16) {   // 16 loads in flight; each group still adds its partials in order
            float v[16];
#pragma unroll
            for (int u = 0; u < 16; ++u) v[u] = __ldcg(src + (size_t)(p + u) * t.total);
#pragma unroll
            for (int u = 0; u < 16; ++u) acc[u & 3] += v[u];
          }
          for (; p + 4 <= t.nparts; p += 4) {
            const float v0 = __ldcg(src + (size_t)p * t.total), v1 = __ldcg(src + (size_t)(p + 1) * t.total);
            const float v2 = __ldcg(src + (size_t)(p + 2) * t.total), v3 = __ldcg(src + (size_t)(p + 3) * t.total);
            acc[0] += v0; acc[1] += v1; acc[2] += v2; acc[3] += v3;
          }
          if (p < t.nparts) acc[0] += __ldcg(src + (size_t)p * t.total);
          if (p + 1 < t.nparts) acc[1] += __ldcg(src + (size_t)(p + 1) * t.total);
          if (p + 2 < t.nparts) acc[2] += __ldcg(src + (size_t)(p + 2) * t.total);
        }
        const float g = (acc[0] + acc[1]) + (acc[2] + acc[3]);
        t.grad[i] = g;
        if (t.has_peers) peer_push(t.pv, epoch & 1u, i, g);   // posted NVLink stores
        else sq = fmaf(g, g, sq);
      }
    }
    if (t.has_peers) {
      peer_push_fence(t.pv);
      // every push of this rank is fenced -> raise this rank's flag everywhere, wait for everybody's, add the slots in
      // rank order (bit-identical on every rank)
      grid_barrier(g_dw_grid_sync, gridDim.x);
      if (blockIdx.x == 0) {
        if (tid == 0) t.pv.stamp[3] = global_ns();
        peer_raise_flags(t.pv, epoch + 1u);
      }
      peer_wait_all(t.pv, epoch + 1u);
      if (blockIdx.x == 0 && tid == 0) t.pv.stamp[4] = global_ns();
      for (int base = blockIdx.x * TB; base < t.total; base += gridDim.x * TB) {
        const int i = base + tid;
        if (tid < TB && i < t.total) {
          const float g = peer_sum(t.pv, epoch & 1u, i);
          t.grad[i] = g;
          const float gs = g * t.adam.scale;
          sq = fmaf(gs, gs, sq);
        }
      }
    }
    sq = warp_sum(sq);
    if (lane == 0) s_red[warp] = sq;
    __syncthreads();
    if (tid == 0) {
      float tot = 0.0f;
      for (int w = 0; w < DW_THREADS / 32; ++w) tot += s_red[w];
      t.sq_part[blockIdx.x] = tot;
      if (blockIdx.x == 0) *t.step += 1;        // Adam's step counter
    }
    if (blockIdx.x == 0 && warp == 1) {         // loss = mean(-surr) + smooth_l1 mean (ppo.py:194, 197)
      float x = 0.0f, y = 0.0f;
      for (int k = lane; k < t.nloss; k += 32) {
        x += t.loss_part[k * 2 + 0];
        y += t.loss_part[k * 2 + 1];
      }
      x = warp_sum(x);
      y = warp_sum(y);
      if (lane == 0) *t.loss = -x * t.inv_rows + y * t.inv_rows;
    }
    __threadfence();
    grid_barrier(g_dw_grid_sync, gridDim.x);
    if (warp == 0) {
      float tot = 0.0f;   // fixed order: lane l adds partials l, l + 32, ...; then the butterfly
      for (int k = lane; k < (int)gridDim.x; k += 32) tot += __ldcg(t.sq_part + k);
      tot = warp_sum(tot);
      if (lane == 0) {
        s_c = adam_coefficients(t.adam, tot, *(volatile long long*)t.step);
        if (blockIdx.x == 0 && t.grad_norm_out) *t.grad_norm_out = s_c.total_norm;
        if (blockIdx.x == 0 && t.has_peers) {
          *t.pv.epoch = epoch + 1u;   // this exchange is over (every CTA read the old value before the first barrier)
          t.pv.stamp[5] = global_ns();
        }
      }
    }
    __syncthreads();
    const AdamCoef c = s_c;
    for (int base = blockIdx.x * TB; base < t.total; base += gridDim.x * TB) {
      const int i = base + tid;
      if (tid < TB && i < t.total)
        adam_entry(i, t.adam, c, t.params, t.grad, t.exp_avg, t.exp_avg_sq, t.wpack, t.wpack_index);
    }
  }
}

// grad = sum of the per-CTA partials (fixed order); entries outside the 12 parameter tensors (alignment padding) = 0
struct ReduceArgs {
  int off[12], size[12];
  int total, nparts, nloss;
  float inv_rows;
  float* sq_part;   // nullable: [gridDim.x] sums of squares of this CTA's entries (flyb_ppo_step: saves the norm kernel)
  long long* step;  // with sq_part: Adam's step counter, += 1 (the norm kernel's other job)
  float scale;      // PEERS == 2: the squares are taken of scale * (sum over ranks), scale = 1 / world
};
// ADAM: clip + Adam ride in the same launch.  Every CTA publishes its sum of squares as a 64-bit word tagged with the
// launch's number, warp 0 of every CTA gathers all of them (spinning on the tag; the grid is small enough to be resident
// as a whole — checked by the host — and every CTA publishes before it waits), adds them in ppo_adam_kernel's order (same
// clip coefficient, bit for bit) and the threads that hold the gradient quads apply Adam to their own entries.  One
// launch and one launch gap fewer per optimiser step than reduce -> ppo_adam_kernel — and 4 us slower (opt-in).
struct ReduceAdam {
  unsigned long long* sq_tag;   // [gridDim.x]  {launch tag << 32 | float bits}
  float *params, *exp_avg, *exp_avg_sq, *grad_norm_out, *wpack;
  const int4* wpack_index;
  uint32_t* peer_epoch;         // nullable: data-parallel exchange epoch, closed here
  AdamArgs adam;
};
__device__ unsigned g_red_seq;   // launches of the ADAM form completed on this device (strictly increasing): the tag
// 256 threads: 64 QUADS of gradient entries (four consecutive floats, one 128-bit load per partial) x 4 interleaved
// partial groups.  Per entry the association is what it was with one entry per thread — group g adds partials g, g + 4,
// ... in order, then (s0 + s1) + (s2 + s3) — so the gradient bits do not depend on the vector width.
// PEERS: 0 one GPU; 1 push the result to every rank, the norm kernel of flyb_ppo_apply completes the sum (fence + flags);
// 2 the FUSED exchange (flyb_ppo_step_dp): the thread that has reduced a quad pushes it as self-validating 64-bit words,
// awaits the peers' words of the same quad and adds them in rank order right here (peer.cuh: peer_ll_push4 /
// peer_ll_sum4), so that `grad` leaves this kernel as the sum over ranks, with its squared-norm partials, and clip + Adam
// follow with no norm kernel.
constexpr int RED_QUADS = 64, RED_ELEMS = 4 * RED_QUADS, RED_GROUPS = 4;
template <int PEERS, bool ADAM>
__global__ void __launch_bounds__(RED_QUADS * RED_GROUPS)
ppo_reduce_tc_kernel(const __grid_constant__ ReduceArgs a, const float* __restrict__ dw_part,
                     const float* __restrict__ loss_part, float* __restrict__ grad, float* __restrict__ loss,
                     const __grid_constant__ PeerView pv, const __grid_constant__ ReduceAdam ad) {
  __shared__ float4 s_acc[RED_GROUPS][RED_QUADS];
  __shared__ float s_sq[2];
  __shared__ AdamCoef s_coef;
  const int e = threadIdx.x & (RED_QUADS - 1), pg = threadIdx.x / RED_QUADS;
  const int i = blockIdx.x * RED_ELEMS + 4 * e;      // first entry of this thread's quad (total is a multiple of 4)
  // ADAM: both are advanced by CTA 0 only after every CTA has published, i.e. after every thread has read them here
  unsigned tag = 0u;
  long long step_new = 0;
  if (ADAM) {
    tag = *(volatile const unsigned*)&g_red_seq + 1u;
    step_new = *(volatile const long long*)a.step + 1;
  }
  uint32_t epoch = 0;
  if (PEERS) {
    epoch = *(volatile const uint32_t*)pv.epoch;   // bumped by ppo_adam_kernel, i.e. after this exchange
    if (blockIdx.x == 0 && threadIdx.x == 0) pv.stamp[0] = global_ns();
  }
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  bool valid[4] = {false, false, false, false};
  if (i < a.total) {
#pragma unroll
    for (int j = 0; j < 4; ++j)
#pragma unroll
      for (int sg = 0; sg < 12; ++sg) valid[j] = valid[j] || (i + j >= a.off[sg] && i + j < a.off[sg] + a.size[sg]);
    if (valid[0] || valid[1] || valid[2] || valid[3]) {   // (alignment padding is never summed: it stays 0)
      const float4* src = reinterpret_cast<const float4*>(dw_part + i);
      const size_t stride4 = (size_t)a.total / 4;
#pragma unroll 8
      for (int p = pg; p < a.nparts; p += RED_GROUPS) {
        const float4 v = src[(size_t)p * stride4];
        acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
      }
    }
  }
  s_acc[pg][e] = acc;
  __syncthreads();
  float g[4] = {0.f, 0.f, 0.f, 0.f};
  if (pg == 0) {   // warps 0 and 1
    if (i < a.total) {
      const float4 s0 = s_acc[0][e], s1 = s_acc[1][e], s2 = s_acc[2][e], s3 = s_acc[3][e];
      g[0] = valid[0] ? (s0.x + s1.x) + (s2.x + s3.x) : 0.0f;
      g[1] = valid[1] ? (s0.y + s1.y) + (s2.y + s3.y) : 0.0f;
      g[2] = valid[2] ? (s0.z + s1.z) + (s2.z + s3.z) : 0.0f;
      g[3] = valid[3] ? (s0.w + s1.w) + (s2.w + s3.w) : 0.0f;
      if (PEERS != 2) *reinterpret_cast<float4*>(grad + i) = make_float4(g[0], g[1], g[2], g[3]);
      // posted NVLink stores into this rank's slot on every rank; the flags are raised by the next kernel
      if (PEERS == 1) peer_push4(pv, epoch & 1u, i, make_float4(g[0], g[1], g[2], g[3]));
      if (PEERS == 2) {
        peer_ll_push4(pv, epoch & 1u, epoch + 1u, i, g);
        if (blockIdx.x == 0 && threadIdx.x == 0) pv.stamp[3] = global_ns();
        const float4 t = peer_ll_sum4(pv, epoch & 1u, epoch + 1u, i);
        if (blockIdx.x == 0 && threadIdx.x == 0) pv.stamp[4] = global_ns();
        *reinterpret_cast<float4*>(grad + i) = t;
        g[0] = t.x * a.scale; g[1] = t.y * a.scale; g[2] = t.z * a.scale; g[3] = t.w * a.scale;
      }
    }
  }
  if (pg == 0 && a.sq_part) {
    const float sq = warp_sum(((g[0] * g[0] + g[1] * g[1]) + g[2] * g[2]) + g[3] * g[3]);
    if ((threadIdx.x & 31) == 0) s_sq[threadIdx.x >> 5] = sq;
  }
  if (PEERS == 1) peer_push_fence(pv);
  if (a.sq_part) {
    __syncthreads();
    if (threadIdx.x == 0) {
      const float sq = s_sq[0] + s_sq[1];
      if (ADAM) {
        const unsigned long long w = ((unsigned long long)tag << 32) | (unsigned long long)__float_as_uint(sq);
        asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(ad.sq_tag + blockIdx.x), "l"(w) : "memory");
      } else {
        a.sq_part[blockIdx.x] = sq;
        if (blockIdx.x == 0) *a.step += 1;
      }
      if (PEERS == 2 && blockIdx.x == 0) pv.stamp[5] = global_ns();
    }
  }
  if (blockIdx.x == 0 && threadIdx.x < 32) {   // loss = mean(-surr) + smooth_l1 mean (ppo.py:194, 197)
    float x = 0.0f, y = 0.0f;
    for (int t = threadIdx.x; t < a.nloss; t += 32) {
      x += loss_part[t * 2 + 0];
      y += loss_part[t * 2 + 1];
    }
    x = warp_sum(x);
    y = warp_sum(y);
    if (threadIdx.x == 0) *loss = -x * a.inv_rows + y * a.inv_rows;
  }
  if (ADAM) {
    if (threadIdx.x < 32) {
      float t = 0.0f;   // ppo_adam_kernel's order: lane l adds partials l, l + 32, ...; then the butterfly
      for (int k = threadIdx.x; k < (int)gridDim.x; k += 32) {
        unsigned long long w;
        for (;;) {
          asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(w) : "l"(ad.sq_tag + k) : "memory");
          if ((unsigned)(w >> 32) == tag) break;
          __nanosleep(32);
        }
        t += __uint_as_float((unsigned)w);
      }
      t = warp_sum(t);
      if (threadIdx.x == 0) {
        s_coef = adam_coefficients(ad.adam, t, step_new);
        if (blockIdx.x == 0) {   // every CTA has published, so every thread of the grid has read step / tag / epoch
          if (ad.grad_norm_out) *ad.grad_norm_out = s_coef.total_norm;
          *a.step = step_new;
          if (ad.peer_epoch) *ad.peer_epoch += 1;   // this exchange is over
          atomicAdd(&g_red_seq, 1u);
        }
      }
    }
    __syncthreads();
    if (pg == 0 && i < a.total) {
      const AdamCoef c = s_coef;
#pragma unroll
      for (int j = 0; j < 4; ++j)   // grad[i + j] was stored by this very thread
        adam_entry(i + j, ad.adam, c, ad.params, grad, ad.exp_avg, ad.exp_avg_sq, ad.wpack, ad.wpack_index);
    }
  }
}

struct Ws {
  float* norm_part;   // first slot, shared with flyb_ppo_apply
  Images im;
  float* loss_part;
  float* dw_part;
  float* sq_part;     // [red_grid] per-CTA sums of squared gradient entries (flyb_ppo_step)
  unsigned long long* sq_tag;   // [red_grid] the same as tagged words (reduction with clip + Adam inside)
  int grid1, ctas_per_role, red_grid;
  size_t bytes;
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

static Ws carve_ws(const FlybNetLayout& net, long long rows, void* basep) {
  Ws w{};
  char* p = (char*)basep;
  size_t off = 0;
  auto take = [&](size_t nfloats) {
    float* r = (float*)(p + off);
    off += align_up(nfloats * sizeof(float), 256);
    return r;
  };
  const long long ntiles = (rows + TM - 1) / TM;
  const size_t R = (size_t)ntiles * TM;
  const int sms = num_sms();
  w.grid1 = (int)std::min<long long>(ntiles, sms);
  w.ctas_per_role = (int)std::max<long long>(1, std::min<long long>(ntiles, sms / 2));
  // stages are dealt with stride cpr (ppo_dw_tc_kernel): a stride that is a multiple of 16 would pin a CTA to one
  // (pair, k-step group) and leave its other pair's accumulators untouched
  if (w.ctas_per_role > 1 && w.ctas_per_role % 16 == 0) w.ctas_per_role -= 1;
  w.norm_part = take(256);   // NORM_BLOCKS floats (ppo_kernels.cu)
  w.im.x = take(R * F_X); w.im.h1 = take(R * F_H1); w.im.h2 = take(R * F_H2); w.im.mv = take(R * F_MV);
  w.im.dzmuv = take(R * F_DZMUV + 4096);   // + slack: the M = 128 MMA of the 64-feature operand never reads it, TMA does not either
  w.im.dzmv = take(R * F_DZMV); w.im.dzh2 = take(R * F_DZH2); w.im.dzh1 = take(R * F_DZH1);
  w.loss_part = take((size_t)sms * 2 + 64);
  w.dw_part = take((size_t)w.ctas_per_role * net.total);
  w.red_grid = (net.total + RED_ELEMS - 1) / RED_ELEMS;
  w.sq_part = take((size_t)w.red_grid);
  w.sq_tag = reinterpret_cast<unsigned long long*>(take((size_t)2 * w.red_grid));
  w.bytes = off;
  return w;
}

}  // namespace tcb

// the chain / weight-gradient kernels assume the first layer's row pitch IS the shape's K0 (80 or 160)
bool tc_ppo_supported(const FlybNetLayout* net) { return net->k0_pad == tcm::K0 && net->num_act <= tcm::APAD; }

size_t tc_ppo_workspace_bytes(const FlybNetLayout* net, long long rows) {
  if (!tc_ppo_supported(net)) return 0;
  return tcb::carve_ws(*net, rows, nullptr).bytes;
}

int tc_ppo_grad(const FlybNetLayout* net, const FlybPpoCfg* cfg, const float* params, const void* wpack,
                const float* obs, const float* action, const float* old_logp, const float* target,
                const float* advantage, const float* action_var, long long rows, float* grad, float* loss,
                void* workspace, const FlybPeerGroup* peers, cudaStream_t st, StepFuse* fuse) {
  using namespace tcb;
  const Ws w = carve_ws(*net, rows, workspace);
  const long long ntiles = (rows + TM - 1) / TM;
  static const bool trace_dw = getenv("FLYB_TRACE_DW") != nullptr;   // which of the two kernels fills the debug trace
  FwdArgs fa{obs, (const unsigned char*)wpack, params, rows, net->num_obs, net->num_act, 0,
             trace_dw ? nullptr : tc_get_trace()};
  GradArgs ga{};
  ga.action = action; ga.old_logp = old_logp; ga.target = target; ga.advantage = advantage; ga.action_var = action_var;
  ga.clip_lo = (float)(1.0 - cfg->clip);
  ga.clip_hi = (float)(1.0 + cfg->clip);
  ga.inv_rows = 1.0f / (float)rows;
  ga.loss_part = w.loss_part;
  const size_t smem1 = smem_bytes(net->num_obs, 0);
  int rc;
  if ((rc = ensure_smem(reinterpret_cast<const void*>(ppo_fwd_bwd_tc_kernel), smem1)) != FLYB_OK) return rc;
  ppo_fwd_bwd_tc_kernel<<<w.grid1, THREADS, smem1, st>>>(*net, fa, ga, w.im);
  FLYB_LAUNCH_CHECK("ppo_fwd_bwd_tc_kernel");

  DwArgs d{};
  // role 0: layer 2 (dZh2 x h1) and layer 4 ([dZmu|dZval] x [m1|v1]);  role 1: layer 1 (dZh1 x X), layer 3 ([dZm1|dZv1] x h2)
  d.role[0].pair[0] = DwPair{w.im.dzh2, w.im.h1, F_DZH2, F_H1, 1, {0u, 0u}, {256u, 256u}};
  d.role[0].pair[1] = DwPair{w.im.dzmuv, w.im.mv, F_DZMUV, F_MV, 1, {288u, 288u}, {416u, 416u}};
  if (WIDE) {   // layer-1 accumulators are 160 columns per M half
    d.role[1].pair[0] = DwPair{w.im.dzh1, w.im.x, F_DZH1, F_X, 2, {0u, 160u}, {448u, 464u}};
    d.role[1].pair[1] = DwPair{w.im.dzmv, w.im.h2, F_DZMV, F_H2, 1, {320u, 320u}, {480u, 480u}};
  } else {
    d.role[1].pair[0] = DwPair{w.im.dzh1, w.im.x, F_DZH1, F_X, 2, {0u, 96u}, {192u, 208u}};
    d.role[1].pair[1] = DwPair{w.im.dzmv, w.im.h2, F_DZMV, F_H2, 1, {256u, 256u}, {384u, 384u}};
  }
  d.ntiles = ntiles;
  d.ctas_per_role = w.ctas_per_role;
  d.dw_part = w.dw_part;
  d.net_total = net->total;
  for (int i = 0; i < 12; ++i) d.off[i] = net->off[i];
  d.A = net->num_act;
  d.trace = trace_dw ? tc_get_trace() : nullptr;
  { const char* e = getenv("FLYB_DW_DBG"); d.dbg = e ? atoi(e) : 0; }
  const size_t smem2 = 1024 + (size_t)DW_RAW_SLOTS * DW_RAW_SLOT + (size_t)DW_OP_SLOTS * DW_OP_SLOT + 512 + 2048 + 256;
  if ((rc = ensure_smem(reinterpret_cast<const void*>(ppo_dw_tc_kernel<false>), smem2)) != FLYB_OK) return rc;
  if ((rc = ensure_smem(reinterpret_cast<const void*>(ppo_dw_tc_kernel<true>), smem2)) != FLYB_OK) return rc;
  const int A = net->num_act;
  const int sizes[12] = {256 * net->k0_pad, 256, 128 * 256, 128, 64 * 128, 64, A * 64, A, 64 * 128, 64, 64, 1};
  if (fuse && fuse->tail && 2 * w.ctas_per_role <= num_sms()) {
    // flyb_ppo_step: reduction + clip + Adam in this kernel's tail, behind grid barriers -> cooperative launch
    DwTail& t = d.tail;
    t.enabled = 1;
    for (int i = 0; i < 12; ++i) { t.off[i] = net->off[i]; t.size[i] = sizes[i]; }
    t.total = net->total; t.nparts = w.ctas_per_role; t.nloss = w.grid1; t.inv_rows = ga.inv_rows;
    t.loss_part = w.loss_part; t.loss = loss; t.grad = grad; t.sq_part = w.sq_part; t.step = fuse->step;
    t.params = fuse->params; t.exp_avg = fuse->exp_avg; t.exp_avg_sq = fuse->exp_avg_sq; t.grad_norm_out = fuse->grad_norm_out;
    t.adam = AdamArgs{fuse->lr, fuse->beta1, fuse->beta2, fuse->eps, fuse->max_norm, peers ? fuse->grad_scale : 1.0f};
    if (peers) {
      t.has_peers = 1;
      t.pv = make_peer_view(*peers);
    }
    t.wpack = (float*)const_cast<void*>(wpack);
    t.wpack_index = (const int4*)tc_wpack_index(net, wpack);
    void* kargs[] = {(void*)&d};
    rc = check_cuda(cudaLaunchCooperativeKernel(reinterpret_cast<const void*>(ppo_dw_tc_kernel<true>), dim3(2 * w.ctas_per_role),
                                                dim3(DW_THREADS), kargs, smem2, st),
                    "cudaLaunchCooperativeKernel(ppo_dw_tc_kernel)");
    if (rc != FLYB_OK) return rc;
    fuse->done = 1;
    return FLYB_OK;
  }
  ppo_dw_tc_kernel<false><<<2 * w.ctas_per_role, DW_THREADS, smem2, st>>>(d);
  FLYB_LAUNCH_CHECK("ppo_dw_tc_kernel");

  ReduceArgs r{};
  for (int i = 0; i < 12; ++i) { r.off[i] = net->off[i]; r.size[i] = sizes[i]; }
  r.total = net->total; r.nparts = w.ctas_per_role; r.nloss = w.grid1; r.inv_rows = ga.inv_rows;
  const int red_grid = w.red_grid;
  r.scale = 1.0f;
  const bool fused_exchange = fuse && peers && fuse->exchange;
  if (fuse && (!peers || fused_exchange)) {   // the caller goes straight on to Adam: the reduction also takes the squared norm
    r.sq_part = w.sq_part;             // (of the sum over ranks, scaled, when it does the exchange too) and counts the step
    r.step = fuse->step;
    fuse->sq_part = w.sq_part;
    fuse->sq_parts = red_grid;
  }
  // Opt-in (FLYB_ADAM_FUSED=1; measured SLOWER than the separate Adam kernel: 142.9 vs 138.7 us per optimiser step —
  // gathering 272 partials in every CTA and 64 threads per CTA doing Adam cost more than the launch they save): clip +
  // Adam inside the reduction, when the caller goes straight on to Adam and the whole grid can be resident at once (its
  // CTAs wait for each other's norm partials).
  const dim3 rb(RED_QUADS * RED_GROUPS);
  ReduceAdam ad{};
  bool adam_in = false;
  if (fuse && (!peers || fused_exchange) && fuse->params) {
    const char* env = getenv("FLYB_ADAM_FUSED");   // read per call: a test flips it
    static const int resident = [] {
      int per_sm0 = 0, per_sm2 = 0;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm0, ppo_reduce_tc_kernel<0, true>, RED_QUADS * RED_GROUPS, 0);
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm2, ppo_reduce_tc_kernel<2, true>, RED_QUADS * RED_GROUPS, 0);
      return std::min(per_sm0, per_sm2) * num_sms();
    }();
    adam_in = env && env[0] == '1' && red_grid <= resident;
  }
  if (adam_in) {
    ad.sq_tag = w.sq_tag;
    ad.params = fuse->params; ad.exp_avg = fuse->exp_avg; ad.exp_avg_sq = fuse->exp_avg_sq;
    ad.grad_norm_out = fuse->grad_norm_out;
    ad.wpack = (float*)const_cast<void*>(wpack);
    ad.wpack_index = (const int4*)tc_wpack_index(net, wpack);
    ad.peer_epoch = peers ? make_peer_view(*peers).epoch : nullptr;
    ad.adam = AdamArgs{fuse->lr, fuse->beta1, fuse->beta2, fuse->eps, fuse->max_norm, peers ? fuse->grad_scale : 1.0f};
  }
  if (fused_exchange) {
    r.scale = fuse->grad_scale;
    if (adam_in) ppo_reduce_tc_kernel<2, true><<<red_grid, rb, 0, st>>>(r, w.dw_part, w.loss_part, grad, loss, make_peer_view(*peers), ad);
    else ppo_reduce_tc_kernel<2, false><<<red_grid, rb, 0, st>>>(r, w.dw_part, w.loss_part, grad, loss, make_peer_view(*peers), ad);
    fuse->done = adam_in ? 3 : 2;   // 2: exchanged, only clip + Adam are left; 3: the whole step is done
  } else if (peers) {
    ppo_reduce_tc_kernel<1, false><<<red_grid, rb, 0, st>>>(r, w.dw_part, w.loss_part, grad, loss, make_peer_view(*peers), ad);
  } else {
    if (adam_in) ppo_reduce_tc_kernel<0, true><<<red_grid, rb, 0, st>>>(r, w.dw_part, w.loss_part, grad, loss, PeerView{}, ad);
    else ppo_reduce_tc_kernel<0, false><<<red_grid, rb, 0, st>>>(r, w.dw_part, w.loss_part, grad, loss, PeerView{}, ad);
    if (adam_in) fuse->done = 3;
  }
  FLYB_LAUNCH_CHECK("ppo_reduce_tc_kernel");
  return FLYB_OK;
}

// The fused exchange makes CTA c of ppo_reduce_tc_kernel<2, *> wait for the words CTA c of every peer's grid pushes.
// That cannot dead-lock only if, on every rank, the whole grid is resident at once (then every CTA pushes before it
// waits, whatever order the hardware dispatches them in).  With a grid larger than the GPU holds, two ranks that
// dispatch their CTAs in different orders could each fill the machine with CTAs waiting for peers' CTAs that are not
// scheduled yet.  True on B200 for both compiled shapes (272 / 370 CTAs against 148 SMs x several CTAs); the check keeps
// it true on a smaller part.  It cannot see SMs held by OTHER work (MPS clients, a concurrent stream of the caller's).
bool tc_ppo_exchange_resident(const FlybNetLayout* net) {
  static const int resident = [] {
    int per_sm = 0, per_sm_adam = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tcb::ppo_reduce_tc_kernel<2, false>, tcb::RED_QUADS * tcb::RED_GROUPS, 0) != cudaSuccess ||
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm_adam, tcb::ppo_reduce_tc_kernel<2, true>, tcb::RED_QUADS * tcb::RED_GROUPS, 0) != cudaSuccess) {
      cudaGetLastError();
      return 0;
    }
    return std::min(per_sm, per_sm_adam) * num_sms();
  }();
  return (net->total + tcb::RED_ELEMS - 1) / tcb::RED_ELEMS <= resident;
}

}  // namespace FLYB_TC_VNS

}  // namespace flyb
