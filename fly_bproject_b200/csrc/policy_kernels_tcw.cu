// The wide shape (K0 = 160, APAD = 48: num_obs <= 160, num_act <= 48 — the 42-joint set, O = 145, A = 42) of the
// tensor-core forward / act kernels: the same source as policy_kernels_tc.cu, compiled into namespace flyb::tcw.
#define FLYB_TC_WIDE 1
#include "policy_kernels_tc.cu"
