// Decode probe (test infrastructure): which shared-memory byte offsets does an MN-major tf32 SS MMA read for A (lane m)?
// One MMA (K = 8); A holds a single 1.0 that is moved over every 16-byte chunk of a 64 KB window, B (K-major) is all ones.
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../fly_bproject_b200/csrc/tc_common.cuh"
using namespace flyb;
using namespace flyb::tc;

constexpr int IMG = 65536;    // bytes of the A window
constexpr int BIMG = 16384;

__global__ void __launch_bounds__(192) k(const float* __restrict__ Bimg, float* __restrict__ D, int N, uint32_t lbo,
                                         uint32_t sbo, uint32_t flags, uint32_t layout_type, int pos, int which) {
  extern __shared__ __align__(1024) unsigned char smem[];
  __shared__ uint64_t bar_d;
  __shared__ uint32_t slot;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // which = 0: probe A (single 1.0 in A window, B ones); which = 1: probe B (A window all ones, single 1.0 in... B window = A window role swap)
  for (int i = threadIdx.x; i < IMG / 4; i += 192) reinterpret_cast<float*>(smem)[i] = (i == pos) ? 1.0f : 0.0f;
  for (int i = threadIdx.x; i < BIMG / 4; i += 192) reinterpret_cast<float*>(smem + IMG)[i] = Bimg[i];
  fence_proxy_async();
  if (threadIdx.x == 0) { mbar_init(&bar_d, 1); fence_mbar_init(); }
  if (warp == 4) tmem_alloc(&slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tb = slot;
  if (warp == 4 && lane == 0) {
    const uint32_t idesc = umma_idesc_tf32(N) | flags;
    const uint32_t sa = smem_u32(smem), sb = sa + IMG;
    const uint64_t probe = umma_smem_desc(sa, lbo, sbo) | ((uint64_t)layout_type << 61);
    const uint64_t ones = umma_smem_desc(sb, 2048, 128);
    if (which == 0) umma_tf32_ss(tb, probe, ones, idesc, false);
    else umma_tf32_ss(tb, ones, probe, idesc, false);
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
      for (int j = 0; j < 16; ++j) D[(size_t)row * N + n0 + j] = __uint_as_float(v[j]);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 4) tmem_dealloc(tb, 512);
}

int main(int argc, char** argv) {
  const int N = 64;
  float *dB, *dD;
  cudaMalloc(&dB, BIMG); cudaMalloc(&dD, 128 * N * 4);
  cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, IMG + BIMG + 1024);
  std::vector<float> ones(BIMG / 4, 1.0f), D(128 * N);
  cudaMemcpy(dB, ones.data(), BIMG, cudaMemcpyHostToDevice);
  struct Cfg { uint32_t lbo, sbo, flags, lt; int which; } cfgs[] = {
      {2048, 512, 1u << 15, 0, 0}, {512, 2048, 1u << 15, 0, 0}, {2048, 512, 1u << 16, 0, 1},
      {2048, 512, 1u << 15, 2, 0},   // SWIZZLE_128B (layout_type 2)
      {2048, 512, 1u << 15, 6, 0},   // SWIZZLE_32B
  };
  for (auto c : cfgs) {
    printf("== which=%d flags=%x lbo=%u sbo=%u layout_type=%u : chunk(16B) offset -> lanes/cols (first float of each chunk probed)\n",
           c.which, c.flags, c.lbo, c.sbo, c.lt);
    int shown = 0;
    for (int pos = 0; pos < IMG / 4; pos += 4) {
      k<<<1, 192, IMG + BIMG + 1024>>>(dB, dD, N, c.lbo, c.sbo, c.flags, c.lt, pos, c.which);
      if (cudaDeviceSynchronize() != cudaSuccess) { printf("CUDA error at pos %d\n", pos); return 1; }
      cudaMemcpy(D.data(), dD, D.size() * 4, cudaMemcpyDeviceToHost);
      bool any = false;
      if (c.which == 0) { for (int m = 0; m < 128; ++m) if (D[m * N] != 0.f) { if (!any) printf("  %6d:", pos * 4); any = true; printf(" m%d", m); } }
      else { for (int n = 0; n < N; ++n) if (D[n] != 0.f) { if (!any) printf("  %6d:", pos * 4); any = true; printf(" n%d", n); } }
      if (any) { printf("\n"); if (++shown > 80) { printf("  ...\n"); break; } }
    }
  }
  return 0;
}
