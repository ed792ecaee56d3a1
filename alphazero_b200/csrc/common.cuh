// Shared host-side helpers for the C-ABI translation units.
#pragma once
#include <cuda_runtime.h>
#include <stdlib.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/az_b200.h"

#define AZ_NUM_SMS_B200 148

// Records the CUDA error text for az_last_cuda_error() and returns AZ_ERR_CUDA.
int az_set_cuda_error(cudaError_t e, const char* where);
int az_set_error_message(int code, const char* msg);

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

// Programmatic dependent launch: the kernel may be scheduled while its predecessor in the stream (or graph) still
// runs, once every CTA of the predecessor has executed az_pdl_trigger() (or exited); it must call az_pdl_wait() before
// touching anything the predecessor writes.  Without a triggering predecessor this is an ordinary launch.
template <typename... KArgs, typename... Args>
static inline cudaError_t az_launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                        Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = stream;
    static const bool off = getenv("AZ_B200_NO_PDL") != nullptr;      // A/B switch for measurements
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = off ? 0 : 1;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}
#ifdef __CUDACC__
// waits until the predecessor grid has completed and its writes are visible (no-op for an ordinary launch)
__device__ __forceinline__ void az_pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
// lets a dependent launched with az_launch_pdl start its prologue on SM resources this grid no longer needs
__device__ __forceinline__ void az_pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
#endif

static inline bool az_valid_n(int n) { return n == 4 || n == 6 || n == 8; }

// Grid for a grid-stride elementwise kernel: enough blocks to cover `items`, capped at a multiple of the SM
// count (148 SMs x 8 resident 256-thread CTAs).
static inline int az_grid_for(int64_t items, int block) {
    int64_t need = (items + block - 1) / block;
    int64_t cap = (int64_t)AZ_NUM_SMS_B200 * 8;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}
