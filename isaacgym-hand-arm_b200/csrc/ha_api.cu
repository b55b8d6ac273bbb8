// C-ABI plumbing shared by all task families: version, error string, struct sizes.
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/handarm_b200.h"
#include "ha_host.h"

namespace ha {
static thread_local char g_err[512] = "";

int set_error(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

bool pdl_enabled() {
  static int v = -1;
  if (v < 0) {
    const char* e = getenv("HA_PDL");
    v = (e && e[0] == '1') ? 1 : 0;      // opt-in: measured neutral on B200 (the post kernel fills the register file, so no co-residency)
  }
  return v == 1;
}

int device_sm_count() {
  static int count[kMaxDevices] = {};
  const int dev = current_device();
  if (count[dev] == 0) cudaDeviceGetAttribute(&count[dev], cudaDevAttrMultiProcessorCount, dev);
  return count[dev];
}

int check_cuda(cudaError_t e, const char* what) {
  if (e == cudaSuccess) return HA_OK;
  return set_error(HA_ERR_CUDA, "%s: %s (%s)", what, cudaGetErrorName(e), cudaGetErrorString(e));
}
}  // namespace ha

extern "C" {
int ha_abi_version(void) { return HA_ABI_VERSION; }
const char* ha_last_error(void) { return ha::g_err; }
size_t ha_sizeof_kuka_params(void) { return sizeof(HaKukaParams); }
size_t ha_sizeof_kuka_buffers(void) { return sizeof(HaKukaBuffers); }
size_t ha_sizeof_kuka_dynamic(void) { return sizeof(HaKukaDynamic); }
}
