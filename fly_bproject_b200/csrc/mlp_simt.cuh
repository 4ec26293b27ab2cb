// fp32 CUDA-core forward of the actor-critic MLP (reference ppo.py:10-38, 88-102) on a 64-row tile held in
// shared memory.  This is the exact-fp32 path (parity <= 1e-5 against the CPU oracle); activations never
// leave the SM between layers, weights stream L2 -> smem in 16-column k-chunks through a cp.async double buffer.
//
// CTA = 256 threads, tile = 64 rows.  Thread (rg = tid/32, cg = tid%32) owns rows rg*8..rg*8+7 and output
// columns {cg + 32 j}.  Per 4 k-values a thread issues 8 broadcast LDS.128 for x, NOUT/32 conflict-free LDS.128
// for w (row stride 20 floats: the 8 lanes of a quarter-warp cover all 32 banks) and 32*NOUT/32 FFMA.
#pragma once
#include "flyb_common.cuh"

namespace flyb {

constexpr int MLP_TM = 64;        // rows per tile
constexpr int MLP_THREADS = 256;
constexpr int MLP_KC = 16;        // k-chunk
constexpr int MLP_WLD = 20;       // smem row stride of a weight chunk (floats)
constexpr int MLP_H1 = 256, MLP_H2 = 128, MLP_H3 = 64;
constexpr int MLP_LD0 = MLP_H1 + 4;  // row stride of buf0 (holds h1, later m1 | mu | v1 side by side)

struct MlpSmem {
  float* buf0;   // [64][260]
  float* buf1;   // [64][ld1]   x, then h2
  float* wst;    // [2][256][20]
  int ld1;
};

__host__ __device__ inline int mlp_ld1(int k0_pad) { return (k0_pad > MLP_H2 ? k0_pad : MLP_H2) + 4; }
__host__ __device__ inline size_t mlp_smem_bytes(int k0_pad) {
  return sizeof(float) * ((size_t)MLP_TM * MLP_LD0 + (size_t)MLP_TM * mlp_ld1(k0_pad) + 2u * 256u * MLP_WLD);
}
FLYB_D MlpSmem mlp_carve(float* smem, int k0_pad) {
  MlpSmem s;
  s.ld1 = mlp_ld1(k0_pad);
  s.buf0 = smem;
  s.buf1 = s.buf0 + MLP_TM * MLP_LD0;
  s.wst = s.buf1 + MLP_TM * s.ld1;
  return s;
}

template <int NOUT>
FLYB_D void mlp_load_wchunk(float* ws, const float* __restrict__ Wg, int ldw, int k0, int nout_real) {
  // NOUT rows x 16 floats -> NOUT*4 16-byte pieces
  for (int f = threadIdx.x; f < NOUT * 4; f += MLP_THREADS) {
    const int n = f >> 2, q = f & 3;
    float* dst = ws + n * MLP_WLD + q * 4;
    if (n < nout_real) {
      cp_async16(dst, Wg + (size_t)n * ldw + k0 + q * 4);
    } else {
      *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

// Y[64][NOUT] = act(X[64][K] * W^T + b);  W is [nout_real][ldw] row-major in global memory, K % 16 == 0.
// gout (nullable): also store the first `g_rows` rows of Y to global memory, row stride nout_real.
template <int NOUT, bool ELU>
FLYB_D void mlp_layer(const float* __restrict__ Wg, int ldw, const float* __restrict__ bg, int K, int nout_real,
                      const float* Xs, int ldx, float* Ys, int ldy, float* wst, float* __restrict__ gout = nullptr,
                      int g_rows = 0) {
  constexpr int CPT = NOUT / 32;
  static_assert(NOUT % 32 == 0, "NOUT must be a multiple of 32");
  const int cg = threadIdx.x & 31, rg = threadIdx.x >> 5;
  float acc[8][CPT];
#pragma unroll
  for (int r = 0; r < 8; ++r)
#pragma unroll
    for (int j = 0; j < CPT; ++j) acc[r][j] = 0.0f;

  const int nchunks = K / MLP_KC;
  mlp_load_wchunk<NOUT>(wst, Wg, ldw, 0, nout_real);
  cp_async_commit();
  for (int c = 0; c < nchunks; ++c) {
    float* ws = wst + (c & 1) * (256 * MLP_WLD);
    if (c + 1 < nchunks) {
      mlp_load_wchunk<NOUT>(wst + ((c + 1) & 1) * (256 * MLP_WLD), Wg, ldw, (c + 1) * MLP_KC, nout_real);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const float* xrow = Xs + (rg * 8) * ldx + c * MLP_KC;
#pragma unroll
    for (int kq = 0; kq < MLP_KC / 4; ++kq) {
      float4 w[CPT];
#pragma unroll
      for (int j = 0; j < CPT; ++j) w[j] = *reinterpret_cast<const float4*>(ws + (cg + 32 * j) * MLP_WLD + kq * 4);
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 x = *reinterpret_cast<const float4*>(xrow + r * ldx + kq * 4);
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
          acc[r][j] = fmaf(x.x, w[j].x, acc[r][j]);
          acc[r][j] = fmaf(x.y, w[j].y, acc[r][j]);
          acc[r][j] = fmaf(x.z, w[j].z, acc[r][j]);
          acc[r][j] = fmaf(x.w, w[j].w, acc[r][j]);
        }
      }
    }
    __syncthreads();
  }
#pragma unroll
  for (int j = 0; j < CPT; ++j) {
    const int n = cg + 32 * j;
    const float b = (n < nout_real) ? __ldg(bg + n) : 0.0f;
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      float y = acc[r][j] + b;
      if (ELU) y = elu(y);
      Ys[(rg * 8 + r) * ldy + n] = y;
      if (gout != nullptr && n < nout_real && rg * 8 + r < g_rows) gout[(size_t)(rg * 8 + r) * nout_real + n] = y;
    }
  }
  __syncthreads();
}

// Backward-through-a-layer product on a tile:  Y[64][NOUT] (+)= X[64][K] * W,  W is [k_real][NOUT] row-major in
// global memory (the SAME nn.Linear weight, read along its rows), K % 16 == 0, rows k >= k_real read as zero.
// Accumulators are passed in/out so that two products can be summed (dh2 = dZm1*Wm1 + dZv1*Wv1).
template <int NOUT>
FLYB_D void mlp_load_wchunk_nt(float* ws, const float* __restrict__ Wg, int k0, int k_real) {
  // 16 rows x NOUT floats -> 16 * NOUT/4 16-byte pieces; smem row stride NOUT
  for (int f = threadIdx.x; f < MLP_KC * (NOUT / 4); f += MLP_THREADS) {
    const int kk = f / (NOUT / 4), q = f - kk * (NOUT / 4);
    float* dst = ws + kk * NOUT + q * 4;
    if (k0 + kk < k_real) {
      cp_async16(dst, Wg + (size_t)(k0 + kk) * NOUT + q * 4);
    } else {
      *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
}

template <int NOUT>
FLYB_D void mlp_layer_nt_acc(float (&acc)[8][NOUT / 32], const float* __restrict__ Wg, int K, int k_real,
                             const float* Xs, int ldx, float* wst) {
  constexpr int CPT = NOUT / 32;
  const int cg = threadIdx.x & 31, rg = threadIdx.x >> 5;
  const int nchunks = K / MLP_KC;
  mlp_load_wchunk_nt<NOUT>(wst, Wg, 0, k_real);
  cp_async_commit();
  for (int c = 0; c < nchunks; ++c) {
    const float* ws = wst + (c & 1) * (256 * MLP_WLD);
    if (c + 1 < nchunks) {
      mlp_load_wchunk_nt<NOUT>(wst + ((c + 1) & 1) * (256 * MLP_WLD), Wg, (c + 1) * MLP_KC, k_real);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const float* xrow = Xs + (rg * 8) * ldx + c * MLP_KC;
#pragma unroll
    for (int kq = 0; kq < MLP_KC / 4; ++kq) {
      float w[4][CPT];
#pragma unroll
      for (int u = 0; u < 4; ++u)
#pragma unroll
        for (int j = 0; j < CPT; ++j) w[u][j] = ws[(kq * 4 + u) * NOUT + cg + 32 * j];
#pragma unroll
      for (int r = 0; r < 8; ++r) {
        const float4 x = *reinterpret_cast<const float4*>(xrow + r * ldx + kq * 4);
#pragma unroll
        for (int j = 0; j < CPT; ++j) {
          acc[r][j] = fmaf(x.x, w[0][j], acc[r][j]);
          acc[r][j] = fmaf(x.y, w[1][j], acc[r][j]);
          acc[r][j] = fmaf(x.z, w[2][j], acc[r][j]);
          acc[r][j] = fmaf(x.w, w[3][j], acc[r][j]);
        }
      }
    }
    __syncthreads();
  }
}

// Stage a tile of observation rows (row-major, O floats each) into buf1, zero-padded to k0_pad columns.
FLYB_D void mlp_load_obs_tile(const MlpSmem& s, const float* __restrict__ obs, long long row0, int nrows, int O,
                              int k0_pad) {
  const float* src = obs + (size_t)row0 * O;
  const int n = nrows * O;
  if ((reinterpret_cast<uintptr_t>(src) & 15u) == 0) {
    const int n4 = n >> 2;
    for (int i = threadIdx.x; i < n4; i += MLP_THREADS) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(src) + i);
      const float vv[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int e = i * 4 + u;
        const int r = e / O, cidx = e - r * O;
        s.buf1[r * s.ld1 + cidx] = vv[u];
      }
    }
    for (int e = (n4 << 2) + threadIdx.x; e < n; e += MLP_THREADS) {
      const int r = e / O, cidx = e - r * O;
      s.buf1[r * s.ld1 + cidx] = __ldg(src + e);
    }
  } else {
    for (int e = threadIdx.x; e < n; e += MLP_THREADS) {
      const int r = e / O, cidx = e - r * O;
      s.buf1[r * s.ld1 + cidx] = __ldg(src + e);
    }
  }
  const int padc = k0_pad - O;
  for (int e = threadIdx.x; e < MLP_TM * padc; e += MLP_THREADS) {
    const int r = e / padc, cidx = O + (e - r * padc);
    s.buf1[r * s.ld1 + cidx] = 0.0f;
  }
  // rows beyond nrows: zero so that the arithmetic stays finite (results are never stored)
  for (int e = threadIdx.x; e < (MLP_TM - nrows) * O; e += MLP_THREADS) {
    const int r = nrows + e / O, cidx = e % O;
    s.buf1[r * s.ld1 + cidx] = 0.0f;
  }
  __syncthreads();
}

// Column offsets inside buf0 after the trunk: m1 [0,64), mu [64, 64+A_pad), v1 [128,192), value at col 192.
constexpr int MLP_COL_M1 = 0, MLP_COL_MU = 64, MLP_COL_V1 = 128, MLP_COL_VAL = 192;

// Optional global copies of the hidden activations of a tile (pointers already offset to the tile's first row).
struct MlpSave {
  float* h1 = nullptr;
  float* h2 = nullptr;
  float* m1 = nullptr;
  float* v1 = nullptr;
  int rows = 0;
};

// Full forward of one staged tile.  After return: (WANT_PI) m1 in buf0[:, 0:64], mu in buf0[:, 64 : 64+A];
// (WANT_V) v1 in buf0[:, 128:192], value in buf0[:, 192]; h2 in buf1[:, 0:128].
template <bool WANT_PI, bool WANT_V, int APAD>
FLYB_D void mlp_forward_tile(const MlpSmem& s, const FlybNetLayout& net, const float* __restrict__ params,
                             const MlpSave save = MlpSave()) {
  const float* p = params;
  // trunk (ppo.py:18-23)
  mlp_layer<MLP_H1, true>(p + net.off[0], net.k0_pad, p + net.off[1], net.k0_pad, MLP_H1, s.buf1, s.ld1, s.buf0,
                          MLP_LD0, s.wst, save.h1, save.rows);
  mlp_layer<MLP_H2, true>(p + net.off[2], MLP_H1, p + net.off[3], MLP_H1, MLP_H2, s.buf0, MLP_LD0, s.buf1, s.ld1,
                          s.wst, save.h2, save.rows);
  if (WANT_PI) {
    // mean head (ppo.py:26-31): ELU after BOTH layers
    mlp_layer<MLP_H3, true>(p + net.off[4], MLP_H2, p + net.off[5], MLP_H2, MLP_H3, s.buf1, s.ld1,
                            s.buf0 + MLP_COL_M1, MLP_LD0, s.wst, save.m1, save.rows);
    mlp_layer<APAD, true>(p + net.off[6], MLP_H3, p + net.off[7], MLP_H3, net.num_act, s.buf0 + MLP_COL_M1, MLP_LD0,
                          s.buf0 + MLP_COL_MU, MLP_LD0, s.wst);
  }
  if (WANT_V) {
    // value head (ppo.py:34-38): no activation on the output
    mlp_layer<MLP_H3, true>(p + net.off[8], MLP_H2, p + net.off[9], MLP_H2, MLP_H3, s.buf1, s.ld1,
                            s.buf0 + MLP_COL_V1, MLP_LD0, s.wst, save.v1, save.rows);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float w0 = __ldg(p + net.off[10] + lane), w1 = __ldg(p + net.off[10] + lane + 32);
    const float b = __ldg(p + net.off[11]);
    for (int r = warp; r < MLP_TM; r += MLP_THREADS / 32) {
      const float* v1 = s.buf0 + r * MLP_LD0 + MLP_COL_V1;
      float a = warp_sum(fmaf(v1[lane], w0, v1[lane + 32] * w1));
      if (lane == 0) s.buf0[r * MLP_LD0 + MLP_COL_VAL] = a + b;
    }
    __syncthreads();
  }
}

}  // namespace flyb
