// Stand-alone probe (test infrastructure, not product) of the MN-major SS form the weight-gradient kernel uses:
//   D[M=128 x N] = sum_r A[r][m] * B[r][n]      (the sample dimension r is the MMA's K)
// Both operands live in shared memory as "operand images": per 8-row k-step a block [hi | lo], each plane
// [F/4 feature groups][8 rows][4 floats] — the UMMA no-swizzle MN-major canonical layout (SBO = 128 B between
// feature groups of 4, rows of a k-step 16 B apart).  3xTF32, checked against a double-precision host product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o probe_umma_mn probe_umma_mn.cu && ./probe_umma_mn
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../fly_bproject_b200/csrc/tc_common.cuh"

using namespace flyb;
using namespace flyb::tc;

__global__ void __launch_bounds__(192) probe_kernel(const float* __restrict__ Aimg, const float* __restrict__ Bimg,
                                                    float* __restrict__ D, int FA, int N, int ksteps) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t bar_b, bar_d;
  __shared__ uint32_t tmem_base_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t a_bytes = (uint32_t)ksteps * FA * 64u, b_bytes = (uint32_t)ksteps * N * 64u;
  if (threadIdx.x == 0) {
    mbar_init(&bar_b, 1);
    mbar_init(&bar_d, 1);
    fence_mbar_init();
  }
  if (warp == 4) tmem_alloc(&tmem_base_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = tmem_base_slot;
  if (warp == 4 && lane == 0) {
    mbar_arrive_expect_tx(&bar_b, a_bytes + b_bytes);
    for (uint32_t off = 0; off < a_bytes; off += 8192u)
      tma_load_1d(smem + off, (const unsigned char*)Aimg + off, min(8192u, a_bytes - off), &bar_b);
    for (uint32_t off = 0; off < b_bytes; off += 8192u)
      tma_load_1d(smem + a_bytes + off, (const unsigned char*)Bimg + off, min(8192u, b_bytes - off), &bar_b);
    mbar_wait(&bar_b, 0);
    const uint32_t idesc = umma_idesc_tf32(N) | (1u << 15) | (1u << 16);   // A and B MN-major
    const uint32_t sa = smem_u32(smem), sb = sa + a_bytes;
    bool acc = false;
    for (int s = 0; s < ksteps; ++s) {
      const uint32_t ah = sa + (uint32_t)s * FA * 64u, al = ah + FA * 32u;
      const uint32_t bh = sb + (uint32_t)s * N * 64u, bl = bh + N * 32u;
      const uint64_t a_hi = umma_smem_desc(ah, FA * 64u, 128u), a_lo = umma_smem_desc(al, FA * 64u, 128u);
      const uint64_t b_hi = umma_smem_desc(bh, N * 64u, 128u), b_lo = umma_smem_desc(bl, N * 64u, 128u);
      umma_tf32_ss(tb, a_lo, b_hi, idesc, acc);
      umma_tf32_ss(tb, a_hi, b_lo, idesc, true);
      umma_tf32_ss(tb, a_hi, b_hi, idesc, true);
      acc = true;
    }
    umma_commit(&bar_d);
  }
  if (warp < 4) {
    mbar_wait(&bar_d, 0);
    tc_fence_after();
    const int row = warp * 32 + lane;
    for (int n0 = 0; n0 < N; n0 += 16) {
      uint32_t v[16];
      tmem_ld16(tmem_addr(tb, warp * 32, n0), v);
      tmem_wait_ld();
#pragma unroll
      for (int j = 0; j < 16; ++j) D[(size_t)row * N + n0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tb, 512);
}

static float host_tf32(float x) {
  uint32_t u = tf32_rna_bits(x);
  float r;
  memcpy(&r, &u, 4);
  return r;
}

// image of X[R rows][F feats]: element (r, f) of plane p at ((r/8) * 2 + p) * F*8 + (f/4)*32 + (r%8)*4 + f%4
static void make_image(const std::vector<float>& X, int R, int F, std::vector<float>& img) {
  img.assign((size_t)R * F * 2, 0.f);
  for (int r = 0; r < R; ++r)
    for (int f = 0; f < F; ++f) {
      const float x = X[(size_t)r * F + f];
      const float hi = host_tf32(x), lo = host_tf32(x - hi);
      const size_t base = (size_t)(r / 8) * 2 * F * 8 + (size_t)(f / 4) * 32 + (r % 8) * 4 + f % 4;
      img[base] = hi;
      img[base + (size_t)F * 8] = lo;
    }
}

// FA = features of the A image (64: the MMA's M = 128 over-reads, lanes >= 64 are ignored; or 128)
static int run_case(int FA, int N, int R) {
  std::vector<float> A((size_t)R * FA), B((size_t)R * N), Ai, Bi, D(128 * (size_t)N);
  srand(99 + FA + 3 * N + R);
  for (auto& x : A) x = (float)rand() / RAND_MAX * 2.f - 1.f;
  for (auto& x : B) x = ((float)rand() / RAND_MAX * 2.f - 1.f) * 0.3f;
  make_image(A, R, FA, Ai);
  make_image(B, R, N, Bi);
  float *dA, *dB, *dD;
  cudaMalloc(&dA, Ai.size() * 4 + 65536);
  cudaMalloc(&dB, Bi.size() * 4);
  cudaMalloc(&dD, D.size() * 4);
  cudaMemset(dA, 0, Ai.size() * 4 + 65536);
  cudaMemcpy(dA, Ai.data(), Ai.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, Bi.data(), Bi.size() * 4, cudaMemcpyHostToDevice);
  const size_t smem = (Ai.size() + Bi.size()) * 4 + 8192;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  cudaMemset(dD, 0xff, D.size() * 4);
  probe_kernel<<<1, 192, smem>>>(dA, dB, dD, FA, N, R / 8);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("FA=%d N=%d R=%d: CUDA error %s\n", FA, N, R, cudaGetErrorString(e));
    return 1;
  }
  cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
  double max_err = 0;
  for (int m = 0; m < FA && m < 128; ++m)
    for (int n = 0; n < N; ++n) {
      double ref = 0, mag = 0;
      for (int r = 0; r < R; ++r) {
        ref += (double)A[(size_t)r * FA + m] * B[(size_t)r * N + n];
        mag += fabs((double)A[(size_t)r * FA + m] * B[(size_t)r * N + n]);
      }
      const double err = fabs((double)D[(size_t)m * N + n] - ref) / (mag + 1e-30);
      if (!(err <= max_err)) max_err = err;
    }
  printf("MN-major FA=%3d N=%3d R=%3d: max |err| / sum|a b| = %.3e %s\n", FA, N, R, max_err,
         max_err <= 2e-6 ? "ok" : "FAIL");
  cudaFree(dA);
  cudaFree(dB);
  cudaFree(dD);
  return !(max_err <= 2e-6);
}

int main() {
  int bad = 0;
  bad |= run_case(128, 128, 32);
  bad |= run_case(128, 256, 16);
  bad |= run_case(128, 80, 32);
  bad |= run_case(128, 16, 32);
  bad |= run_case(64, 128, 32);
  printf(bad ? "PROBE FAILED\n" : "PROBE OK\n");
  return bad;
}
