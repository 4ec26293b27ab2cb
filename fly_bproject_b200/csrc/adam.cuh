// clip_grad_norm_ + torch.optim.Adam.step arithmetic (reference ppo.py:198-199), shared by ppo_adam_kernel and the
// tail of the tensor-core weight-gradient kernel (flyb_ppo_step).  Operation order follows torch's single-tensor Adam.
#pragma once
#include "tc_common.cuh"

namespace flyb {

struct AdamArgs {
  double lr, beta1, beta2, eps, max_norm;
  float scale;
};

struct AdamCoef {
  float coef, step_size, bc2_sqrt, total_norm;
};

// from the squared gradient norm (after `scale`) and Adam's step number
FLYB_D AdamCoef adam_coefficients(const AdamArgs& a, float sq_norm, long long step) {
  AdamCoef c;
  c.total_norm = sqrtf(sq_norm);
  // torch.nn.utils.clip_grad_norm_: clip_coef = max_norm / (total_norm + 1e-6), clamped to 1.0
  c.coef = fminf((float)a.max_norm / (c.total_norm + 1e-6f), 1.0f);
  const double tstep = (double)step;
  const double bc1 = 1.0 - pow(a.beta1, tstep);
  const double bc2 = 1.0 - pow(a.beta2, tstep);
  c.step_size = (float)(a.lr / bc1);
  c.bc2_sqrt = (float)sqrt(bc2);
  return c;
}

// one parameter: p.grad after clipping is left in grad[i]; the tensor-core weight stream is kept in step in place
FLYB_D void adam_entry(int i, const AdamArgs& a, const AdamCoef& c, float* __restrict__ params, float* __restrict__ grad,
                       float* __restrict__ exp_avg, float* __restrict__ exp_avg_sq, float* __restrict__ wpack,
                       const int4* __restrict__ wpack_index) {
  const float w1 = (float)(1.0 - a.beta1), b2 = (float)a.beta2, w2 = (float)(1.0 - a.beta2), eps = (float)a.eps;
  const float g = grad[i] * a.scale * c.coef;
  grad[i] = g;                                      // p.grad after clip_grad_norm_
  float m = exp_avg[i], v = exp_avg_sq[i];
  m = m + w1 * (g - m);                             // exp_avg.lerp_(grad, 1 - beta1)
  v = v * b2;                                       // exp_avg_sq.mul_(beta2)
  v = v + (w2 * g) * g;                             //   .addcmul_(grad, grad, value = 1 - beta2)
  const float denom = sqrtf(v) / c.bc2_sqrt + eps;  // (exp_avg_sq.sqrt() / bias_correction2_sqrt).add_(eps)
  const float p_new = params[i] + (-c.step_size) * (m / denom);   // param.addcdiv_(exp_avg, denom, value = -step_size)
  params[i] = p_new;
  exp_avg[i] = m;
  exp_avg_sq[i] = v;
  if (wpack_index) {   // hi/lo TF32 copies of this parameter in the weight stream (forward + transposed backward operand)
    const int4 d = wpack_index[i];
    if (d.x >= 0) {
      const float hi = __uint_as_float(tc::tf32_rna_bits(p_new));
      const float lo = __uint_as_float(tc::tf32_rna_bits(p_new - hi));
      wpack[d.x] = hi;
      wpack[d.x + d.y] = lo;
      if (d.z >= 0) {
        wpack[d.z] = hi;
        wpack[d.z + d.w] = lo;
      }
    }
  }
}

// Grid-wide barrier for kernels launched cooperatively (all CTAs resident): `state` = {arrival count, generation},
// zero at first use and self-resetting.  Every thread of every CTA calls it.
FLYB_D void grid_barrier(unsigned* state, unsigned nblocks) {
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned gen;
    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(gen) : "l"(state + 1) : "memory");
    __threadfence();
    if (atomicAdd(state, 1u) == nblocks - 1u) {
      state[0] = 0u;
      __threadfence();
      asm volatile("st.release.gpu.global.u32 [%0], %1;" ::"l"(state + 1), "r"(gen + 1u) : "memory");
    } else {
      unsigned now;
      do {
        __nanosleep(32);
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(now) : "l"(state + 1) : "memory");
      } while (now == gen);
    }
    __threadfence();
  }
  __syncthreads();
}

}  // namespace flyb
