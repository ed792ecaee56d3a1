// az_set_cuda_error for the probe library (the product library has its own copy in board_kernels.cu).
#include <cuda_runtime.h>
#include <stdio.h>
#include "../../alphazero_b200/csrc/common.cuh"
static char g_err[256];
int az_set_cuda_error(cudaError_t e, const char* where) {
    snprintf(g_err, sizeof g_err, "%s: %s", where, cudaGetErrorString(e));
    return AZ_ERR_CUDA;
}
extern "C" const char* az_probes_last_error(void) { return g_err; }
