// Shared helpers for the flyb200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/flyb200.h"

#define FLYB_HD __host__ __device__ __forceinline__
#define FLYB_D __device__ __forceinline__

namespace flyb {

void set_error(const char* fmt, ...);
int check_cuda(cudaError_t err, const char* what);
// opt a kernel in to `smem` bytes of dynamic shared memory (> 48 KB needs it); cached per kernel ADDRESS
int ensure_smem(const void* kernel, size_t smem);

#define FLYB_REQUIRE(cond, ...)      \
  do {                               \
    if (!(cond)) {                   \
      flyb::set_error(__VA_ARGS__);  \
      return FLYB_EINVAL;            \
    }                                \
  } while (0)

#define FLYB_LAUNCH_CHECK(what)                                \
  do {                                                         \
    int _e = flyb::check_cuda(cudaGetLastError(), what);       \
    if (_e != FLYB_OK) return _e;                              \
  } while (0)

inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }

// streaming (read-once) 128-bit global load / store: keep L1 for the gathers that do have reuse
FLYB_D float4 ldg_stream4(const float4* p) {
  float4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w)
               : "l"(p));
  return v;
}
FLYB_D void stg_stream4(float4* p, const float4& v) {
  asm volatile("st.global.L1::no_allocate.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

FLYB_D float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

FLYB_D void cp_async16(void* smem_dst, const void* gmem_src) {
  unsigned s = static_cast<unsigned>(__cvta_generic_to_shared(smem_dst));
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(s), "l"(gmem_src) : "memory");
}
FLYB_D void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
FLYB_D void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// torch's ELU forward (alpha = 1): x > 0 ? x : exp(x) - 1   (ATen elu kernel; the reference's nn.ELU, ppo.py:20-30)
FLYB_D float elu(float x) { return x > 0.0f ? x : expf(x) - 1.0f; }
// d elu / dx expressed with the OUTPUT y: y > 0 -> 1, else exp(x) = y + 1
FLYB_D float elu_grad_from_out(float y) { return y > 0.0f ? 1.0f : y + 1.0f; }

}  // namespace flyb
