// Counter-based normal noise for the policy sample (reference ppo.py:217-218 draws from torch's global generator;
// here the stream is keyed by (seed, draw number, element index) so that a captured CUDA graph draws fresh noise).
#pragma once
#include "flyb_common.cuh"

namespace flyb {

// ---- Philox4x32-10 (Salmon et al. 2011), counter-based: one call yields four 32-bit words
FLYB_D uint4 philox4x32_10(uint4 ctr, uint2 key) {
  const unsigned M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
  for (int i = 0; i < 10; ++i) {
    const unsigned hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
    const unsigned hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
    ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
    key.x += W0;
    key.y += W1;
  }
  return ctr;
}

// Two standard normals (Box-Muller, cos and sin branch) from Philox counter `pair` of draw number `offset` under
// `seed`.  Fast-math log / sin / cos: this is sampling noise, not part of the parity contract (parity tests pass
// their noise explicitly).
FLYB_D void philox_normal_pair(unsigned long long seed, unsigned long long offset, unsigned long long pair,
                               float& n_even, float& n_odd) {
  const uint4 r = philox4x32_10(make_uint4((unsigned)pair, (unsigned)(pair >> 32), (unsigned)offset,
                                           (unsigned)(offset >> 32)),
                                make_uint2((unsigned)seed, (unsigned)(seed >> 32)));
  const float u1 = ((float)r.x + 1.0f) * 2.3283064365386963e-10f;  // (0, 1]
  const float u2 = (float)r.y * 2.3283064365386963e-10f;           // [0, 1)
  const float rad = sqrtf(-2.0f * __logf(u1));
  float sn, cs;
  __sincosf(6.283185307179586f * u2, &sn, &cs);
  n_even = rad * cs;
  n_odd = rad * sn;
}

// standard normal for element `idx` of the draw: element 2j takes the cos branch of pair j, 2j+1 the sin branch
FLYB_D float philox_normal(unsigned long long seed, unsigned long long offset, unsigned long long idx) {
  float a, b;
  philox_normal_pair(seed, offset, idx >> 1, a, b);
  return (idx & 1ull) ? b : a;
}

}  // namespace flyb
