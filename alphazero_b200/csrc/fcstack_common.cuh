// Helpers shared by the two fully-connected-stack kernels (fcstack_kernels.cu, fcstack2_kernels.cu): cluster barrier,
// watchdog mbarrier waits, remote arrives, bulk copies, TMEM loads (sm_100a inline PTX).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "umma.cuh"

__device__ __forceinline__ void fs_cluster_sync() {
    asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

// mbarrier waits of this kernel: a wait that lasts longer than ~2 s of SM clocks is a protocol bug (a peer CTA that
// never sends / never releases) -- trap, so that the launch fails with an error instead of hanging the device.
#define FS_WAIT_LIMIT (4000000000LL)
__device__ __forceinline__ void fs_mbar_wait(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (;;) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        const long long t = clock64();
        if (t0 == 0) t0 = t;
        else if (t - t0 > FS_WAIT_LIMIT) __trap();
    }
}
// the same with cluster-scope acquire: pairs with the remote release-arrive of another CTA's epilogue
__device__ __forceinline__ void fs_mbar_wait_cluster(uint32_t bar, uint32_t parity) {
    uint32_t done = 0;
    long long t0 = 0;
    for (;;) {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.acquire.cluster.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}" : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        const long long t = clock64();
        if (t0 == 0) t0 = t;
        else if (t - t0 > FS_WAIT_LIMIT) __trap();
    }
}

// arrive (release, cluster scope) on the barrier at the same shared-memory offset in CTA `rank` of the cluster
__device__ __forceinline__ void fs_mbar_arrive_remote(uint32_t bar, uint32_t rank) {
    asm volatile(
        "{\n\t"
        ".reg .b32 ra;\n\t"
        "mapa.shared::cluster.u32 ra, %0, %1;\n\t"
        "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t"
        "}" ::"r"(bar), "r"(rank) : "memory");
}

__device__ __forceinline__ void fs_expect_tx(uint32_t bar, uint32_t bytes) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
}

// global -> shared bulk copy completing on `bar`; with more than one destination the bytes (and the completion) go to
// the same offsets in every CTA of `mask`
__device__ __forceinline__ void fs_bulk_copy(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar, uint32_t mask,
                                             bool multicast) {
    if (multicast)
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1], %2, [%3], %4;"
                     ::"r"(dst), "l"(src), "r"(bytes), "r"(bar), "h"((uint16_t)mask) : "memory");
    else
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

__device__ __forceinline__ void fs_commit_multicast(uint32_t bar, uint32_t mask) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)mask) : "memory");
}

// wait for the tcgen05.ld's issued so far; the registers are read-write operands so that no use is scheduled before it
__device__ __forceinline__ void fs_tmem_wait64(uint32_t (&v)[64]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
        : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
          "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]),
          "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]),
          "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31]),
          "+r"(v[32]), "+r"(v[33]), "+r"(v[34]), "+r"(v[35]), "+r"(v[36]), "+r"(v[37]), "+r"(v[38]), "+r"(v[39]),
          "+r"(v[40]), "+r"(v[41]), "+r"(v[42]), "+r"(v[43]), "+r"(v[44]), "+r"(v[45]), "+r"(v[46]), "+r"(v[47]),
          "+r"(v[48]), "+r"(v[49]), "+r"(v[50]), "+r"(v[51]), "+r"(v[52]), "+r"(v[53]), "+r"(v[54]), "+r"(v[55]),
          "+r"(v[56]), "+r"(v[57]), "+r"(v[58]), "+r"(v[59]), "+r"(v[60]), "+r"(v[61]), "+r"(v[62]), "+r"(v[63])
        :: "memory");
}

__device__ __forceinline__ void fs_tmem_ld16(uint32_t taddr, uint32_t* r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}


__device__ __forceinline__ void fs_tmem_wait32(uint32_t (&v)[32]) {
    asm volatile("tcgen05.wait::ld.sync.aligned;"
        : "+r"(v[0]), "+r"(v[1]), "+r"(v[2]), "+r"(v[3]), "+r"(v[4]), "+r"(v[5]), "+r"(v[6]), "+r"(v[7]),
          "+r"(v[8]), "+r"(v[9]), "+r"(v[10]), "+r"(v[11]), "+r"(v[12]), "+r"(v[13]), "+r"(v[14]), "+r"(v[15]),
          "+r"(v[16]), "+r"(v[17]), "+r"(v[18]), "+r"(v[19]), "+r"(v[20]), "+r"(v[21]), "+r"(v[22]), "+r"(v[23]),
          "+r"(v[24]), "+r"(v[25]), "+r"(v[26]), "+r"(v[27]), "+r"(v[28]), "+r"(v[29]), "+r"(v[30]), "+r"(v[31])
        :: "memory");
}

// address of the same shared-memory offset in CTA `rank` of the cluster (distributed shared memory)
__device__ __forceinline__ uint32_t fs_mapa(uint32_t addr, uint32_t rank) {
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
    return r;
}
__device__ __forceinline__ void fs_st_cluster_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared::cluster.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}

// ------------------------------------------------------------------------------------------------ host side

// Shapes of the three layers (az_net_fcstack_plan).  variant 1: k_net_fcstack (any cluster shape); variant 2:
// k_net_fcstack2 (clusters of 4 column slices, one row tile): Nc[0] is then the width of one of fc1's two passes and
// Nc[2] the padded count of ALL head outputs (the heads are split over K, not over their outputs); variant 3:
// k_net_fcstack3 (the same schedule on CTA pairs, RT = 2: each CTA stages half of its column slice's weight columns).
struct FcPlan {
    int variant;
    int kch[3], Nc[3], K[3];
    int hpc, n3c;
    uint32_t w_off[3], b_off[3];
    int64_t blob_bytes;
};

int az_fcstack2_launch(const FcPlan& pl, const void* x, const void* wblob, void* h1, void* logits, int ld, int n_out,
                       int64_t G, long long* debug, int late_trigger, bool pdl, void* stream);
int az_fcstack3_launch(const FcPlan& pl, const void* x, const void* wblob, void* h1, void* xch, void* logits, int ld,
                       int n_out, int64_t G, long long* debug, int late_trigger, bool pdl, void* stream);
