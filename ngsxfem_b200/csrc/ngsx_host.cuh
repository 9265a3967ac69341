// ngsx_host.cuh -- host-side plumbing shared by the translation units of libngsx_b200.so:
// error string, launch counter, per-device constant data (quadrature rules, decomposition table).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <atomic>
#include <string>
#include "../../include/ngsx_abi.h"
#include "ngsx_device.cuh"

namespace ngsx {

void set_error(const std::string &msg);
int reset_err(int64_t *err, cudaStream_t st);      // {count, first} record of the degenerate-level-set check -> {0, INT64_MAX}
extern std::atomic<int64_t> g_launches;

#define NGSX_CUDA(call)                                                                              \
  do {                                                                                               \
    cudaError_t err__ = (call);                                                                      \
    if (err__ != cudaSuccess) {                                                                      \
      ngsx::set_error(std::string(#call) + ": " + cudaGetErrorString(err__));                        \
      return NGSX_ERR_CUDA;                                                                          \
    }                                                                                                \
  } while (0)

#define NGSX_LAUNCH_CHECK()                                                                          \
  do {                                                                                               \
    ngsx::g_launches.fetch_add(1, std::memory_order_relaxed);                                        \
    cudaError_t err__ = cudaGetLastError();                                                          \
    if (err__ != cudaSuccess) {                                                                      \
      ngsx::set_error(std::string("kernel launch: ") + cudaGetErrorString(err__));                   \
      return NGSX_ERR_CUDA;                                                                          \
    }                                                                                                \
  } while (0)

#define NGSX_REQUIRE(cond, msg)                                                                      \
  do {                                                                                               \
    if (!(cond)) { ngsx::set_error(msg); return NGSX_ERR_ARG; }                                      \
  } while (0)

struct DeviceData {
  double4 *rules = nullptr;       // all reference rules back to back: point | segm | trig | tet
  CutTable *cut_table = nullptr;
  int sm_count = 0;
};

// per-device constant data, uploaded on first use.  returns nullptr (error set) on failure.
const DeviceData *device_data();
// reference rule of a simplex of dimension sdim (0..3): device pointer and size; false if the order is not tabulated
bool rule_lookup(const DeviceData *dd, int sdim, int order, const double4 **ptr, int *n);

// one-shot packed-output request of ngsx_set_packed_output: returns the row-slot map (or nullptr) and clears it
const int32_t *take_packed_output();

inline cudaStream_t as_stream(void *s) { return reinterpret_cast<cudaStream_t>(s); }

}  // namespace ngsx
