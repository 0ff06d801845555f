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

__device__ __forceinline__ float eeg_silu(float x) { return x / (1.0f + __expf(-x)); }
