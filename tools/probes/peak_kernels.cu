// Measurement-only kernels: the integer-ALU issue rate of this GPU, the denominator of the board kernels' roofline
// (MEASURED_PEAKS.json only holds HBM bandwidth and bf16 throughput).  Not on any product path.
//
// Every thread runs 8 independent dependency chains of one instruction kind; with 8 resident 256-thread CTAs per SM
// the issue slots, not the latency, are the limit.  Reported unit: warp instructions per second.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/az_b200.h"
#include "../../alphazero_b200/csrc/common.cuh"
#include "az_probes.h"

template <int KIND>
__global__ void __launch_bounds__(256) k_peak_int32(uint32_t* __restrict__ out, int iters, uint32_t b, uint32_t c) {
    uint32_t a[8];
    #pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = (blockIdx.x * 256u + threadIdx.x) * 8u + j;
    for (int i = 0; i < iters; ++i) {
        #pragma unroll
        for (int u = 0; u < 4; ++u) {
            #pragma unroll
            for (int j = 0; j < 8; ++j) {
                // volatile asm: the chains must not be folded into closed forms
                if (KIND == 0) asm volatile("lop3.b32 %0, %0, %1, %2, 0x6a;" : "+r"(a[j]) : "r"(b), "r"(c));   // LOP3
                else if (KIND == 1) asm volatile("shf.l.wrap.b32 %0, %0, %1, %2;" : "+r"(a[j]) : "r"(b), "r"(c)); // SHF
                else if (KIND == 2) asm volatile("add.u32 %0, %0, %1;" : "+r"(a[j]) : "r"(b));                  // IADD3
                else {                                                      // the movegen step: x |= (x << s) & m
                    uint32_t t;
                    asm volatile("shf.l.wrap.b32 %0, %1, %2, %3;" : "=r"(t) : "r"(a[j]), "r"(b), "r"(c));
                    asm volatile("lop3.b32 %0, %0, %1, %2, 0xf8;" : "+r"(a[j]) : "r"(t), "r"(b));             // a | (t & b)
                }
            }
        }
        b += (uint32_t)i;      // keeps the chain from folding
    }
    uint32_t x = 0;
    #pragma unroll
    for (int j = 0; j < 8; ++j) x ^= a[j];
    out[blockIdx.x * 256u + threadIdx.x] = x;
}

extern "C" int az_peak_int32(int kind, int iters, uint32_t* out, int64_t out_len, int64_t* ops_per_thread,
                             void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    const int blocks = AZ_NUM_SMS_B200 * 8;
    if (kind < 0 || kind > 3 || iters < 1 || !out || out_len < (int64_t)blocks * 256) return AZ_ERR_ARG;
    switch (kind) {
        case 0: k_peak_int32<0><<<blocks, 256, 0, stream>>>(out, iters, 0x5a5a5a5au, 0x01234567u); break;
        case 1: k_peak_int32<1><<<blocks, 256, 0, stream>>>(out, iters, 0x5a5a5a5au, 0x01234567u); break;
        case 2: k_peak_int32<2><<<blocks, 256, 0, stream>>>(out, iters, 0x5a5a5a5au, 0x01234567u); break;
        default: k_peak_int32<3><<<blocks, 256, 0, stream>>>(out, iters, 0x5a5a5a5au, 0x01234567u); break;
    }
    AZ_CHECK_LAUNCH("k_peak_int32");
    if (ops_per_thread) *ops_per_thread = (int64_t)iters * 32 * (kind == 3 ? 2 : 1);
    return AZ_OK;
}
