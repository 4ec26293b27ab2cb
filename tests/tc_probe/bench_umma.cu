// Micro-benchmark of tcgen05.mma kind::tf32 issue/throughput on one SM (test infrastructure, not product):
// cycles per MMA for N in {32,64,128,256}, A from TMEM (TS) or shared memory (SS), with a commit every `cevery`
// MMAs, accumulating into one D tile or alternating between two.  Operand contents are irrelevant (zeros).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o bench_umma bench_umma.cu && ./bench_umma
#include <cstdio>
#include <cstdlib>

#include "../../fly_bproject_b200/csrc/tc_common.cuh"

using namespace flyb;
using namespace flyb::tc;

template <bool SS>
__device__ __forceinline__ void mma_if(uint32_t leader, uint32_t d, uint32_t a_t, uint64_t a_d, uint64_t b_d,
                                       uint32_t idesc) {
  if (SS) {
    asm volatile(
        "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %4, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, 1;\n\t}" ::"r"(d),
        "l"(a_d), "l"(b_d), "r"(idesc), "r"(leader)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %4, 0;\n\t"
        "@q tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, 1;\n\t}" ::"r"(d),
        "r"(a_t), "l"(b_d), "r"(idesc), "r"(leader)
        : "memory");
  }
}

// CE = commit after every CE MMAs (0 = never inside the loop); ALT = alternate between two accumulators
template <int N, bool SS, int CE, bool ALT>
__global__ void __launch_bounds__(64) bench_kernel(int trips, long long* out) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar, bar2;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < (64 * 1024) / 4; i += 64) reinterpret_cast<uint32_t*>(smem)[i] = 0;
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    mbar_init(&bar2, 1);
    fence_mbar_init();
  }
  if (warp == 0) tmem_alloc(&slot, 512);
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot;
  if (warp == 0) {
    const uint32_t leader = elect_one() ? 1u : 0u;
    const uint32_t idesc = umma_idesc_tf32(N);
    const uint32_t sb = smem_u32(smem);
    const uint64_t b0 = umma_smem_desc(sb, N * 16, 128);
    const uint64_t a0 = umma_smem_desc(sb + 32768u, 2048, 128);
    const uint32_t bar_addr = smem_u32(&bar);
    long long t0 = clock64();
    for (int i = 0; i < trips; ++i) {
#pragma unroll
      for (int u = 0; u < 24; ++u) {
        const uint64_t b_d = b0 + (uint64_t)(((u & 7) * (N * 32) % 32768) >> 4);
        const uint64_t a_d = a0 + (uint64_t)(((u & 3) * 4096) >> 4);
        const uint32_t d = tb + 256 + ((ALT && (u & 1)) ? (uint32_t)(N % 256) : 0u);
        mma_if<SS>(leader, d, tb + (uint32_t)((u & 7) * 8), a_d, b_d, idesc);
        if (CE > 0 && (u + 1) % CE == 0)
          asm volatile(
              "{\n\t.reg .pred q;\n\tsetp.ne.b32 q, %1, 0;\n\t"
              "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}" ::"r"(bar_addr),
              "r"(leader)
              : "memory");
      }
    }
    long long t1 = clock64();
    if (leader)
      asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&bar2))
                   : "memory");
    for (int spin = 0; spin < (1 << 24) && !mbar_try_wait(&bar2, 0); ++spin) {
    }
    long long t2 = clock64();
    if (threadIdx.x == 0) {
      out[0] = t1 - t0;
      out[1] = t2 - t0;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tb, 512);
}

template <int N, bool SS, int CE, bool ALT>
static void run(long long* dout) {
  const int trips = 40;
  cudaFuncSetAttribute(bench_kernel<N, SS, CE, ALT>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  bench_kernel<N, SS, CE, ALT><<<1, 64, 65536>>>(trips, dout);
  cudaError_t e = cudaDeviceSynchronize();
  long long h[2] = {0, 0};
  cudaMemcpy(h, dout, sizeof(h), cudaMemcpyDeviceToHost);
  printf("N=%3d %s commit_every=%2d alt_d=%d : issue %6.1f cyc/MMA, drained %6.1f cyc/MMA  (floor %d) %s\n", N,
         SS ? "SS" : "TS", CE, (int)ALT, (double)h[0] / (trips * 24), (double)h[1] / (trips * 24), N / 2,
         e == cudaSuccess ? "" : cudaGetErrorString(e));
}

template <bool SS>
static void sweep(long long* dout) {
  run<32, SS, 0, false>(dout);
  run<64, SS, 0, false>(dout);
  run<128, SS, 0, false>(dout);
  run<256, SS, 0, false>(dout);
  run<32, SS, 6, false>(dout);
  run<64, SS, 3, false>(dout);
  run<128, SS, 3, false>(dout);
  run<128, SS, 1, false>(dout);
  run<64, SS, 0, true>(dout);
  run<128, SS, 0, true>(dout);
}

int main() {
  long long* dout;
  cudaMalloc(&dout, 16);
  sweep<false>(dout);
  sweep<true>(dout);
  return 0;
}
