// Error reporting, ABI version and the flat parameter layout of the actor-critic net.
#include <stdarg.h>
#include <string.h>

#include "flyb_common.cuh"

namespace flyb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int check_cuda(cudaError_t err, const char* what) {
  if (err == cudaSuccess) return FLYB_OK;
  set_error("%s: %s (%s)", what, cudaGetErrorString(err), cudaGetErrorName(err));
  return FLYB_ECUDA;
}

int ensure_smem(const void* kernel, size_t smem) {
  struct Entry { const void* k; size_t granted; };
  static Entry table[64];
  static int used = 0;
  for (int i = 0; i < used; ++i)
    if (table[i].k == kernel) {
      if (smem <= table[i].granted) return FLYB_OK;
      int rc = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                          "cudaFuncSetAttribute(max dynamic smem)");
      if (rc == FLYB_OK) table[i].granted = smem;
      return rc;
    }
  int rc = check_cuda(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem),
                      "cudaFuncSetAttribute(max dynamic smem)");
  if (rc == FLYB_OK && used < 64) table[used++] = Entry{kernel, smem};
  return rc;
}

}  // namespace flyb

extern "C" int flyb_abi_version(void) { return FLYB_ABI_VERSION; }
extern "C" const char* flyb_last_error(void) { return flyb::g_err; }

extern "C" int flyb_device_count(void) {
  int n = 0;
  int rc = flyb::check_cuda(cudaGetDeviceCount(&n), "cudaGetDeviceCount");
  return rc == FLYB_OK ? n : rc;
}

// Segment order = the reference's state_dict order (ppo.py:18-38); out-features x in-features, row-major.
extern "C" int flyb_net_layout(int32_t num_obs, int32_t num_act, FlybNetLayout* out) {
  FLYB_REQUIRE(out != nullptr, "flyb_net_layout: out is NULL");
  FLYB_REQUIRE(num_obs > 0 && num_obs <= 1024, "num_obs %d out of range", num_obs);
  FLYB_REQUIRE(num_act > 0 && num_act <= FLYB_MAX_DOF, "num_act %d out of range", num_act);
  memset(out, 0, sizeof(*out));
  out->num_obs = num_obs;
  out->num_act = num_act;
  out->k0_pad = (num_obs + 15) / 16 * 16;
  const int sizes[12] = {256 * out->k0_pad, 256, 128 * 256, 128, 64 * 128, 64, num_act * 64, num_act,
                         64 * 128,          64,  64,        1};
  int off = 0;
  for (int i = 0; i < 12; ++i) {
    out->off[i] = off;
    off += (sizes[i] + 3) / 4 * 4;
  }
  out->total = off;
  return FLYB_OK;
}
