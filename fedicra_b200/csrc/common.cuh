// common.cuh -- shared helpers of libfedicra_sm100.so (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/fedicra_sm100.h"

#define FEDICRA_ABI_VERSION 4

// Record an error message for fedicra_last_error() and return `code` (always negative).
int fedicra_fail(int code, const char *fmt, ...);

#define FEDICRA_REQUIRE(cond, ...)                      \
    do {                                                \
        if (!(cond)) return fedicra_fail(-1, __VA_ARGS__); \
    } while (0)

// Check the launch that was just issued (asynchronous errors surface at the caller's next sync, as with any kernel).
#define FEDICRA_CHECK_LAUNCH(name)                                                             \
    do {                                                                                       \
        fedicra_count_launches(1);                                                             \
        cudaError_t e_ = cudaGetLastError();                                                   \
        if (e_ != cudaSuccess) return fedicra_fail(-2, "%s: %s", name, cudaGetErrorString(e_)); \
    } while (0)

// Telemetry: number of kernels this library has launched in the process (bench.py's `gpu_launches`).
void fedicra_count_launches(int n);

static inline size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

__device__ __forceinline__ float ex2_approx(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
