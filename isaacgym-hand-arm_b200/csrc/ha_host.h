// host-side helpers shared by the C-ABI translation units
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ha {
int set_error(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline bool aligned8(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 7u) == 0; }
// true iff HA_PDL=1: kernels are launched with programmatic stream serialization (the launch of kernel k+1
// overlaps the drain of kernel k; works in eager streams and under stream capture into CUDA graphs)
bool pdl_enabled();
// Per-device launch state: function attributes (cudaFuncSetAttribute) and the SM count belong to a device, not to the process,
// so a process that drives several GPUs configures each one the first time it launches there.
constexpr int kMaxDevices = 64;
inline int current_device() {
  int d = 0;
  cudaGetDevice(&d);
  return (d >= 0 && d < kMaxDevices) ? d : 0;
}
int device_sm_count();     // SM count of the current device (cached per device)
struct PerDeviceFlag {     // "configured on this device yet?"
  bool done[kMaxDevices] = {};
  bool& here() { return done[current_device()]; }
};

template <typename... KArgs, typename... Args>
inline cudaError_t launch(void (*kernel)(KArgs...), int grid, int block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
}  // namespace ha

#define HA_REQUIRE(cond, ...)                                        \
  do {                                                               \
    if (!(cond)) return ha::set_error(HA_ERR_INVALID, __VA_ARGS__);  \
  } while (0)
#define HA_REQUIRE_ALIGNED16(ptr)                                                                  \
  do {                                                                                             \
    if (!ha::aligned16(ptr)) return ha::set_error(HA_ERR_ALIGN, #ptr " must be 16-byte aligned"); \
  } while (0)
