// Shared helpers for libeeg_b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/eeg_b200.h"

#define EEG_NUM_SMS 148

void eeg_set_error(const char* fmt, ...);

#define EEG_CHECK_LAUNCH(name)                                            \
  do {                                                                    \
    cudaError_t _e = cudaGetLastError();                                  \
    if (_e != cudaSuccess) {                                              \
      eeg_set_error("%s: launch failed: %s", name, cudaGetErrorString(_e)); \
      return 2;                                                           \
    }                                                                     \
  } while (0)

#define EEG_REQUIRE(cond, ...)       \
  do {                               \
    if (!(cond)) {                   \
      eeg_set_error(__VA_ARGS__);    \
      return 1;                      \
    }                                \
  } while (0)

static inline int eeg_div_up(int64_t a, int64_t b) { return (int)((a + b - 1) / b); }

// cudaFuncSetAttribute applies per device: one "already opted in" flag per device ordinal, looked up with the device
// that is current at launch time (a process may drive several GPUs).
struct EegDevFlags {
  bool f[64] = {};
};
static inline bool& eeg_dev_flag(EegDevFlags& fl) {
  int d = 0;
  cudaGetDevice(&d);
  return fl.f[d & 63];
}

__device__ __forceinline__ float eeg_silu(float x) { return x / (1.0f + __expf(-x)); }
