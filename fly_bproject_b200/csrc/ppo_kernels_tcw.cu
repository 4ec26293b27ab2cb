// The wide shape (K0 = 160, APAD = 48: the 42-joint set, O = 145, A = 42) of the tensor-core PPO step — forward/backward
// chain, weight gradients, reduction: the same source as ppo_kernels_tc.cu, compiled into namespace flyb::tcw.
#define FLYB_TC_WIDE 1
#include "ppo_kernels_tc.cu"
