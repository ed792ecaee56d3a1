// Measurement-only kernel: cycles per tcgen05.mma (kind::f16, bf16 operands from shared memory in the K-major
// no-swizzle layout the network trunk uses, fp32 accumulator in TMEM) as a function of N, of the number of
// accumulating K steps and of the number of accumulator regions in flight.  The numbers explain the trunk's design
// (DESIGN.md section 6); nothing on the product path calls this.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/az_b200.h"
#include "../../alphazero_b200/csrc/common.cuh"
#include "az_probes.h"

namespace probe {

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ uint64_t desc(uint32_t smem_addr, uint32_t lbo, uint32_t sbo) {
    uint64_t d = 0;
    d |= (uint64_t)((smem_addr >> 4) & 0x3FFFu);
    d |= (uint64_t)((lbo >> 4) & 0x3FFFu) << 16;
    d |= (uint64_t)((sbo >> 4) & 0x3FFFu) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}

__device__ __forceinline__ uint32_t idesc_bf16(uint32_t m, uint32_t n) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((m >> 4) << 24);
}

__device__ __forceinline__ void mma(uint32_t tmem_d, uint64_t a, uint64_t b, uint32_t idesc, uint32_t acc) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d), "l"(a), "l"(b), "r"(idesc), "r"(acc)
        : "memory");
}

}  // namespace probe

#define PROBE_MAX_KSTEPS 8
#define PROBE_A_BYTES (PROBE_MAX_KSTEPS * 2 * 128 * 16)          // [k chunk][128 rows][8] bf16
#define PROBE_B_BYTES (PROBE_MAX_KSTEPS * 2 * 256 * 16)          // [k chunk][256 rows][8] bf16

// flags: 1 = every MMA accumulates (no zero-initialising MMA at the start of a chain), 2 = tcgen05.commit after every
// chain (onto a barrier nobody waits for), 4 = warps 1-3 keep reading 96 accumulator columns (tcgen05.ld) meanwhile,
// 8 = (flags >> 8) dependent ALU instructions of "bookkeeping" between chains (is the MMA queue deep enough to hide them?)
template <int KS>
__global__ void __launch_bounds__(128) k_mma_probe(int m, int n, int reps, int regions, int flags,
                                                   long long* __restrict__ cycles) {
    extern __shared__ __align__(128) uint8_t smem[];
    __shared__ uint64_t bar_w[4];
    __shared__ uint64_t bar2;
    __shared__ volatile int done;
    __shared__ uint32_t tmem_ptr;
    uint8_t* a_smem = smem;
    uint8_t* b_smem = smem + PROBE_A_BYTES;
    for (int i = threadIdx.x; i < (PROBE_A_BYTES + PROBE_B_BYTES) / 16; i += blockDim.x)
        reinterpret_cast<uint4*>(smem)[i] = make_uint4(0x3c003c00u, 0x3c003c00u, 0u, 0u);
    const int warp = threadIdx.x >> 5;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(probe::smem_u32(&tmem_ptr)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    if (threadIdx.x == 0) {
        for (int i = 0; i < 4; ++i)
            asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(probe::smem_u32(&bar_w[i])), "r"(1));
        asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(probe::smem_u32(&bar2)), "r"(1 << 20));
        done = 0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = tmem_ptr;
    long long t0 = 0, t1 = 0;
    const int issuers = ((flags >> 4) & 3) + 1;      // warps 0 .. issuers-1 each issue `reps` chains (n <= 128 then)
    if ((threadIdx.x & 31) == 0 && warp < issuers) {
        uint64_t& bar = bar_w[warp];
        const uint32_t idesc = probe::idesc_bf16((uint32_t)m, (uint32_t)n);
        const uint32_t a_lbo = 128 * 16, b_lbo = 256 * 16;          // bytes between the two 8-element K chunks
        uint64_t ad[KS], bd[KS];                                     // descriptors precomputed: the timed loop is
        #pragma unroll                                               // nothing but the MMA instructions
        for (int ks = 0; ks < KS; ++ks) {
            ad[ks] = probe::desc(probe::smem_u32(a_smem) + ks * 2 * a_lbo, a_lbo, 128);
            bd[ks] = probe::desc(probe::smem_u32(b_smem) + ks * 2 * b_lbo, b_lbo, 128);
        }
        const uint32_t d_step = (regions == 2 && issuers == 1) ? 256u : 0u;
        const uint32_t bar2_addr = probe::smem_u32(&bar2);
        for (int pass = 0; pass < 2; ++pass) {                      // pass 0 warms up, pass 1 is timed
            t0 = clock64();
            uint32_t d = tmem_base + (issuers > 1 ? (uint32_t)warp * 128u : 0u);
            for (int r = 0; r < reps; ++r) {
                #pragma unroll
                for (int ks = 0; ks < KS; ++ks) probe::mma(d, ad[ks], bd[ks], idesc, (ks > 0 || (flags & 1)) ? 1u : 0u);
                if (flags & 2)
                    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar2_addr) : "memory");
                d ^= d_step;
                if (flags & 8) {                     // ~delay of bookkeeping between chains: a dependent ALU chain
                    uint32_t x = (uint32_t)r;
                    #pragma unroll 1
                    for (int i = 0; i < (flags >> 8); ++i) asm volatile("lop3.b32 %0, %0, %0, %0, 0x6a;" : "+r"(x));
                    if (x == 0xdeadbeefu) d ^= 1u;
                }
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(probe::smem_u32(&bar)) : "memory");
            asm volatile(
                "{\n\t.reg .pred p;\n\tPW:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra PD;\n\tbra PW;\n\tPD:\n\t}" ::"r"(probe::smem_u32(&bar)), "r"(pass) : "memory");
            t1 = clock64();
        }
        if (warp == 0) cycles[blockIdx.x] = t1 - t0;
        done = 1;
    } else if ((flags & 4) && warp > 0 && (threadIdx.x & 31) >= 0 && issuers == 1) {
        uint32_t sink = 0;
        while (!done) {
            uint32_t r[32];
            #pragma unroll
            for (int blk = 0; blk < 3; ++blk) {
                const uint32_t taddr = tmem_base + ((uint32_t)(warp * 32) << 16) + 256u + blk * 32u;
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                    : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
                      "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
                      "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
                      "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
                    : "r"(taddr));
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                #pragma unroll
                for (int i = 0; i < 32; ++i) sink ^= r[i];
            }
        }
        if (sink == 0x12345678u) cycles[blockIdx.x] = 0;
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

extern "C" int az_peak_mma(int m, int n, int k_steps, int reps, int regions, int flags, int64_t* cycles,
                           void* stream) {
    if ((m != 64 && m != 128) || n < 16 || n > 256 || (n % 16) || k_steps < 1 || k_steps > PROBE_MAX_KSTEPS || reps < 1 ||
        (regions != 1 && regions != 2) || !cycles)
        return AZ_ERR_ARG;
    const int bytes = PROBE_A_BYTES + PROBE_B_BYTES;
#define PROBE_CASE(KS)                                                                                            \
    case KS:                                                                                                      \
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_mma_probe<KS>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));   \
        k_mma_probe<KS><<<AZ_NUM_SMS_B200, 128, bytes, (cudaStream_t)stream>>>(m, n, reps, regions, flags,        \
                                                                               (long long*)cycles);               \
        break;
    switch (k_steps) {
        PROBE_CASE(1) PROBE_CASE(2) PROBE_CASE(3) PROBE_CASE(4) PROBE_CASE(6) PROBE_CASE(8)
        default: return AZ_ERR_ARG;
    }
    AZ_CHECK_LAUNCH("k_mma_probe");
    return AZ_OK;
}
