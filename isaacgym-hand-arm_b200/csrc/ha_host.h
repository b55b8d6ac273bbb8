// host-side helpers shared by the C-ABI translation units
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace ha {
int set_error(int code, const char* fmt, ...);
int check_cuda(cudaError_t e, const char* what);
inline bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15u) == 0; }
inline bool aligned8(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 7u) == 0; }
}  // namespace ha

#define HA_REQUIRE(cond, ...)                                        \
  do {                                                               \
    if (!(cond)) return ha::set_error(HA_ERR_INVALID, __VA_ARGS__);  \
  } while (0)
#define HA_REQUIRE_ALIGNED16(ptr)                                                                  \
  do {                                                                                             \
    if (!ha::aligned16(ptr)) return ha::set_error(HA_ERR_ALIGN, #ptr " must be 16-byte aligned"); \
  } while (0)
