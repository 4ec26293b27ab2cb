// Stand-alone probe of the tcgen05 building blocks in csrc/tc_common.cuh (test infrastructure, not product):
//   D[128 x N] = A[128 x K] * B[N x K]^T   with A staged in tensor memory (TS mode), B packed in shared memory
// via a 1-D bulk TMA copy, as one-pass TF32 and as 3xTF32, checked against a double-precision host product.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o probe_umma probe_umma.cu && ./probe_umma
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../fly_bproject_b200/csrc/tc_common.cuh"

using namespace flyb;
using namespace flyb::tc;

// packed B stream: [kstep][hi|lo][slab of N x 8]
__global__ void __launch_bounds__(192) probe_kernel(const float* __restrict__ A, const float* __restrict__ Bpacked,
                                                    float* __restrict__ D, int N, int K, int passes) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ uint64_t bar_b, bar_d;
  __shared__ uint32_t tmem_base_slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ksteps = K / 8;
  const uint32_t slab = (uint32_t)N * 32u;              // bytes of one (hi or lo) k-step slab
  const uint32_t total = (uint32_t)ksteps * 2u * slab;

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
    mbar_arrive_expect_tx(&bar_b, total);
    for (uint32_t off = 0; off < total; off += 8192u) {
      const uint32_t n = min(8192u, total - off);
      tma_load_1d(smem + off, (const unsigned char*)Bpacked + off, n, &bar_b);
    }
  }
  if (warp < 4) {   // A -> TMEM: hi at cols [0,K), lo at cols [K,2K)
    const int row = warp * 32 + lane;
    for (int k0 = 0; k0 < K; k0 += 8) {
      uint32_t hi[8], lo[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) split_tf32(A[(size_t)row * K + k0 + j], hi[j], lo[j]);
      tmem_st8(tmem_addr(tb, warp * 32, k0), hi);
      tmem_st8(tmem_addr(tb, warp * 32, K + k0), lo);
    }
    tmem_wait_st();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();

  const uint32_t dcol = 256;
  if (warp == 4 && lane == 0) {
    mbar_wait(&bar_b, 0);
    const uint32_t idesc = umma_idesc_tf32(N);
    const uint32_t sb = smem_u32(smem);
    bool acc = false;
    for (int s = 0; s < ksteps; ++s) {
      const uint64_t b_hi = umma_smem_desc(sb + (uint32_t)s * 2u * slab, (uint32_t)N * 16u, 128u);
      const uint64_t b_lo = umma_smem_desc(sb + (uint32_t)s * 2u * slab + slab, (uint32_t)N * 16u, 128u);
      const uint32_t a_hi = tb + (uint32_t)(s * 8), a_lo = tb + (uint32_t)(K + s * 8);
      if (passes == 3) {
        umma_tf32_ts(tb + dcol, a_lo, b_hi, idesc, acc);
        umma_tf32_ts(tb + dcol, a_hi, b_lo, idesc, true);
        umma_tf32_ts(tb + dcol, a_hi, b_hi, idesc, true);
      } else {
        umma_tf32_ts(tb + dcol, a_hi, b_hi, idesc, acc);
      }
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
      tmem_ld16(tmem_addr(tb, warp * 32, dcol + n0), v);
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

static int run_case(int N, int K) {
  std::vector<float> A(128 * K), B((size_t)N * K), Bp((size_t)(K / 8) * 2 * N * 8), D(128 * (size_t)N);
  srand(1234 + N * 7 + K);
  for (auto& x : A) x = (float)rand() / RAND_MAX * 2.f - 1.f;
  for (auto& x : B) x = ((float)rand() / RAND_MAX * 2.f - 1.f) * 0.3f;
  for (int s = 0; s < K / 8; ++s)
    for (int n = 0; n < N; ++n)
      for (int k8 = 0; k8 < 8; ++k8) {
        const float w = B[(size_t)n * K + s * 8 + k8];
        const float hi = host_tf32(w), lo = host_tf32(w - hi);
        const size_t base = (size_t)s * 2 * N * 8;
        Bp[base + packed_slab_offset(n, k8, N) / 4] = hi;
        Bp[base + (size_t)N * 8 + packed_slab_offset(n, k8, N) / 4] = lo;
      }
  float *dA, *dB, *dD;
  cudaMalloc(&dA, A.size() * 4);
  cudaMalloc(&dB, Bp.size() * 4);
  cudaMalloc(&dD, D.size() * 4);
  cudaMemcpy(dA, A.data(), A.size() * 4, cudaMemcpyHostToDevice);
  cudaMemcpy(dB, Bp.data(), Bp.size() * 4, cudaMemcpyHostToDevice);
  const size_t smem = Bp.size() * 4;
  cudaFuncSetAttribute(probe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int bad = 0;
  for (int passes : {1, 3}) {
    cudaMemset(dD, 0xff, D.size() * 4);
    probe_kernel<<<1, 192, smem>>>(dA, dB, dD, N, K, passes);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
      printf("N=%d K=%d passes=%d: CUDA error %s\n", N, K, passes, cudaGetErrorString(e));
      return 1;
    }
    cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
    double max_err = 0, max_ref = 0;
    for (int r = 0; r < 128; ++r)
      for (int n = 0; n < N; ++n) {
        double ref = 0, mag = 0;
        for (int k = 0; k < K; ++k) {
          ref += (double)A[(size_t)r * K + k] * B[(size_t)n * K + k];
          mag += fabs((double)A[(size_t)r * K + k] * B[(size_t)n * K + k]);
        }
        const double err = fabs((double)D[(size_t)r * N + n] - ref) / (mag + 1e-30);
        if (!(err <= max_err)) max_err = err;
        if (fabs(ref) > max_ref) max_ref = fabs(ref);
      }
    const double tol = passes == 3 ? 2e-6 : 2e-3;
    printf("N=%3d K=%3d passes=%d: max |err| / sum|a b| = %.3e (tol %.0e) %s\n", N, K, passes, max_err, tol,
           max_err <= tol ? "ok" : "FAIL");
    if (!(max_err <= tol)) bad = 1;
  }
  cudaFree(dA);
  cudaFree(dB);
  cudaFree(dD);
  return bad;
}

int main() {
  int bad = 0;
  bad |= run_case(64, 80);
  bad |= run_case(128, 64);
  bad |= run_case(32, 64);
  bad |= run_case(64, 128);
  bad |= run_case(256, 80);
  printf(bad ? "PROBE FAILED\n" : "PROBE OK\n");
  return bad;
}
