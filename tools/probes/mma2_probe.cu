// Measurement / semantics probe (not part of the product library): one tcgen05.mma.cta_group::2 tile on a CTA pair.
// D[256 x N] = A[256 x 64] * B[N x 64]^T with M = 256: CTA r of the pair stages ITS 128 rows of A and `its` N/2 rows of
// B (K-major no-swizzle), the leader issues K/16 MMAs, both CTAs dump their 128 TMEM lanes x N columns.  The Python side
// (tools/mma2_probe.py) compares the dumps with the float64 product to pin down which half of B / which rows of D live
// in which CTA before the product kernel relies on it.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../../alphazero_b200/csrc/umma.cuh"

#define P2_K 64

__global__ void __launch_bounds__(128, 1) k_mma2_probe(const __nv_bfloat16* __restrict__ A, const __nv_bfloat16* __restrict__ B,
                                                        float* __restrict__ D, int N, int swap_b, int shift) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint32_t rank;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    const int nh = N / 2;
    // A: [8 planes][8 zero rows | 128 rows | 8 zero rows][8] bf16; the descriptor may start `shift` rows (16 bytes each)
    // off the first data row: does a K-major no-swizzle operand need more than 16-byte alignment of its start?
    constexpr int A_PLANE = (128 + 16) * 16;
    uint8_t* a_s = smem;
    uint8_t* b_s = smem + 8 * A_PLANE;                // [8 planes][nh rows][8]
    const uint32_t bar = smem_u32(smem + 8 * A_PLANE + 8 * 16 * 128);
    uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(smem + 8 * A_PLANE + 8 * 16 * 128 + 16);
    for (int i = tid; i < 8 * A_PLANE / 16; i += 128) reinterpret_cast<uint4*>(a_s)[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();

    // operands: thread = row
    {
        const __nv_bfloat16* src = A + (size_t)(rank * 128 + tid) * P2_K;
        for (int pl = 0; pl < 8; ++pl)
            *reinterpret_cast<uint4*>(a_s + pl * A_PLANE + (8 + tid) * 16) = *reinterpret_cast<const uint4*>(src + pl * 8);
        const int half = swap_b ? 1 - (int)rank : (int)rank;
        if (tid < nh) {
            const __nv_bfloat16* bs = B + (size_t)(half * nh + tid) * P2_K;
            for (int pl = 0; pl < 8; ++pl)
                *reinterpret_cast<uint4*>(b_s + pl * nh * 16 + tid * 16) = *reinterpret_cast<const uint4*>(bs + pl * 8);
        }
    }
    if (tid == 0) {
        mbar_init(bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr)), "r"(256));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr;

    if (rank == 0 && warp == 1) {
        if (elect_one()) {
            const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (((uint32_t)N >> 3) << 17) | ((256u >> 4) << 24);
            for (int k = 0; k < P2_K / 16; ++k) {
                const uint64_t ad = umma_desc(smem_u32(a_s) + (uint32_t)(k * 2 * A_PLANE + (8 + shift) * 16), (uint32_t)A_PLANE, 128u);
                const uint64_t bd = umma_desc(smem_u32(b_s) + (uint32_t)(k * 32 * nh), (uint32_t)(16 * nh), 128u);
                asm volatile(
                    "{\n\t"
                    ".reg .pred p;\n\t"
                    "setp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
                    "}" ::"r"(tmem_base), "l"(ad), "l"(bd), "r"(idesc), "r"(k ? 1u : 0u) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                         ::"r"(bar), "h"((uint16_t)3) : "memory");
        }
        __syncwarp();
    }
    // both CTAs: wait for the commit on their own barrier, then dump lanes 32 w .. 32 w + 31, N columns
    {
        uint32_t done = 0;
        long long t0 = clock64();
        while (!done) {
            asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                         : "=r"(done) : "r"(bar), "r"(0u) : "memory");
            if (clock64() - t0 > 2000000000LL) __trap();
        }
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        for (int c0 = 0; c0 < N; c0 += 16) {
            uint32_t v[16];
            tmem_ld16_issue(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
            for (int i = 0; i < 16; ++i) D[((size_t)rank * 128 + tid) * N + c0 + i] = __uint_as_float(v[i]);
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256));
}

extern "C" int az_probe_mma2(const void* A, const void* B, void* D, int N, int swap_b, int shift, void* stream) {
    if (N < 32 || N > 256 || (N % 32) || shift < -8 || shift > 8) return -1;
    const int smem_bytes = 8 * (128 + 16) * 16 + 8 * 16 * 128 + 64;
    cudaFuncSetAttribute(k_mma2_probe, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(2); cfg.blockDim = dim3(128); cfg.dynamicSmemBytes = smem_bytes; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_mma2_probe, (const __nv_bfloat16*)A, (const __nv_bfloat16*)B, (float*)D, N, swap_b, shift);
    if (e != cudaSuccess) return -2;
    return cudaGetLastError() == cudaSuccess ? 0 : -3;
}
