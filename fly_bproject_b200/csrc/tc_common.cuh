// Blackwell (sm_100a) tensor-core plumbing used by the flyb200 MLP kernels: mbarriers, 1-D bulk TMA copies,
// TMEM allocation / tcgen05.ld / tcgen05.st, UMMA descriptors and the tcgen05.mma kind::tf32 issue wrappers.
//
// Numerics: every GEMM on this path is a "3xTF32" product.  An fp32 value x is split into
//   hi = rna_tf32(x),  lo = rna_tf32(x - hi)          (x - hi is exact in fp32)
// and  x*y  is accumulated in fp32 TMEM as  lo_x*hi_y + hi_x*lo_y + hi_x*hi_y  (the dropped lo*lo term is
// <= 2^-22 |x y|), which keeps the MLP within ~1e-6 of an fp32 evaluation — the parity bar of this repo is
// 1e-5 — at one third of the tensor pipe's TF32 rate instead of the CUDA-core FFMA rate.
//
// Operand layouts (no swizzle, K-major "interleaved" canonical layout):
//   B (weights, shared memory): core matrix = 8 rows x 16 bytes (4 tf32), stored as 128 contiguous bytes;
//       element (n, k) of an N x 8 k-step slab lives at byte ((k / 4) * N + n) * 16 + (k % 4) * 4, i.e.
//       LBO (between the two 16-byte K chunks) = 16 N bytes and SBO (between 8-row groups) = 128 bytes.
//   A (activations, tensor memory): lane = row of the 128-row tile, column = k (one 32-bit column per tf32).
//   D (accumulators, tensor memory): lane = row, column = n, fp32.
#pragma once
#include "flyb_common.cuh"

namespace flyb {
namespace tc {

FLYB_D uint32_t smem_u32(const void* p) { return static_cast<uint32_t>(__cvta_generic_to_shared(p)); }

// ---------------------------------------------------------------------------------------------- mbarrier
FLYB_D void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
FLYB_D void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
// make generic-proxy smem writes visible to the async proxy (TMA / tcgen05 reading smem)
FLYB_D void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

FLYB_D void mbar_arrive(uint64_t* bar) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.shared::cta.b64 st, [%0];\n\t}" ::"r"(smem_u32(bar)) : "memory");
}
FLYB_D void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
FLYB_D bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
FLYB_D void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}
// same, for warps whose wake-up latency is not critical: back off between polls so that the polling does not
// take issue slots / barrier-unit bandwidth from the warps doing the work
FLYB_D void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) __nanosleep(64);
}

// ---------------------------------------------------------------------------------------------- TMA (1-D bulk)
// global -> shared::cta, completion signalled on `bar` as transaction bytes.  16-byte aligned, size % 16 == 0.
FLYB_D void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)),
               "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}
// shared::cta -> global bulk store (bulk async-group completion)
FLYB_D void tma_store_1d(void* gmem_dst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gmem_dst), "r"(smem_u32(smem_src)),
               "r"(bytes)
               : "memory");
}
FLYB_D void tma_store_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
FLYB_D void tma_store_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N>
FLYB_D void tma_store_wait() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}

// ---------------------------------------------------------------------------------------------- TMEM
// One full warp allocates `cols` (power of two >= 32) columns; the base address lands in *smem_slot.
FLYB_D void tmem_alloc(uint32_t* smem_slot, uint32_t cols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_slot)), "r"(cols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
FLYB_D void tmem_dealloc(uint32_t taddr, uint32_t cols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols) : "memory");
}
FLYB_D void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
FLYB_D void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
FLYB_D void tmem_wait_ld() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
FLYB_D void tmem_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// warp-collective: lane i of warp w (w % 4 selects the 32-lane TMEM quarter) receives 16 consecutive columns
FLYB_D void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
}
FLYB_D void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
FLYB_D void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr)
               : "memory");
}
FLYB_D void tmem_st8(uint32_t taddr, const uint32_t (&v)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(v[0]),
               "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7])
               : "memory");
}

// TMEM address = (lane << 16) | column
FLYB_D uint32_t tmem_addr(uint32_t base, uint32_t lane, uint32_t col) { return base + (lane << 16) + col; }

// ---------------------------------------------------------------------------------------------- 3xTF32 split
FLYB_HD uint32_t tf32_rna_bits(float x) {
#ifdef __CUDA_ARCH__
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return r;
#else
  // round-to-nearest, ties away from zero, to 10 explicit mantissa bits (host mirror; finite inputs)
  uint32_t u;
  memcpy(&u, &x, 4);
  u += 0x1000u;
  u &= 0xffffe000u;
  return u;
#endif
}
FLYB_D void split_tf32(float x, uint32_t& hi, uint32_t& lo) {
  hi = tf32_rna_bits(x);
  lo = tf32_rna_bits(x - __uint_as_float(hi));
}

// ---------------------------------------------------------------------------------------------- UMMA descriptors
// kind::tf32, fp32 accumulate, A and B K-major, M = 128, N = n (multiple of 16, <= 256)
FLYB_HD uint32_t umma_idesc_tf32(int n) {
  return (1u << 4)                       // c_format = F32
         | (2u << 7)                     // a_format = TF32
         | (2u << 10)                    // b_format = TF32
         | ((uint32_t)(n >> 3) << 17)    // n_dim
         | ((uint32_t)(128 >> 4) << 24); // m_dim
}
// shared-memory matrix descriptor, SWIZZLE_NONE, K-major; lbo / sbo in bytes
FLYB_D uint64_t umma_smem_desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3fffu);
  d |= (uint64_t)((lbo >> 4) & 0x3fffu) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3fffu) << 32;
  d |= (uint64_t)1 << 46;   // descriptor version (Blackwell)
  return d;
}

// D[tmem] (+)= A[tmem] * B[smem]^T      (single thread)
FLYB_D void umma_tf32_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], [%1], %2, %3, p;\n\t}" ::"r"(d_tmem),
      "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}
// D[tmem] (+)= A[smem] * B[smem]^T      (single thread)
FLYB_D void umma_tf32_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, bool accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
      "l"(a_desc), "l"(b_desc), "r"(idesc), "r"((uint32_t)accumulate)
      : "memory");
}
// arrive on `bar` once every tcgen05.mma issued so far by this thread has completed
FLYB_D void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}

FLYB_D bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}

// Byte offset of element (n, k) inside one packed k-step slab (N rows x 8 k) — see the header comment.
FLYB_HD int packed_slab_offset(int n, int k8, int N) { return ((k8 >> 2) * N + n) * 16 + (k8 & 3) * 4; }

}  // namespace tc
}  // namespace flyb
