// Shared host-side helpers for the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/az_b200.h"

#define AZ_NUM_SMS_B200 148

// Records the CUDA error text for az_last_cuda_error() and returns AZ_ERR_CUDA.
int az_set_cuda_error(cudaError_t e, const char* where);

#define AZ_CUDA_TRY(expr)                                         \
    do {                                                          \
        cudaError_t _e = (expr);                                  \
        if (_e != cudaSuccess) return az_set_cuda_error(_e, #expr); \
    } while (0)

#define AZ_CHECK_LAUNCH(name)                                     \
    do {                                                          \
        cudaError_t _e = cudaGetLastError();                      \
        if (_e != cudaSuccess) return az_set_cuda_error(_e, name); \
    } while (0)

static inline bool az_valid_n(int n) { return n == 4 || n == 6 || n == 8; }

// Grid for a grid-stride elementwise kernel: enough blocks to cover `items`, capped at a multiple of the SM
// count (148 SMs x 8 resident 256-thread CTAs).
static inline int az_grid_for(int64_t items, int block) {
    int64_t need = (items + block - 1) / block;
    int64_t cap = (int64_t)AZ_NUM_SMS_B200 * 8;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}
