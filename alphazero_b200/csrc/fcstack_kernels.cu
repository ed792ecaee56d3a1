// The fully connected stack of the policy-value networks -- fc1 + BN + ReLU -> fc2 + BN + ReLU -> fc_probs || fc_value
// (alphazero/games/othello.py:352-362 layers, :374-382 forward; connect4.py:381-390, :403-411) -- as ONE tcgen05
// kernel for sm_100a.  Replaces the three library GEMMs the first version of the fused evaluator launched.
//
// Work split.  A thread-block cluster of RT x NS CTAs owns RT row tiles of 128 positions; CTA (r, c) computes, for row
// tile r, the c-th slice of the output columns of every layer (fc1: N1/NS columns, fc2: N2/NS, heads: a padded share
// of the A + 1 outputs).  With 4096 positions, RT = 2 and NS = 4 that is 16 clusters = 128 CTAs, one per SM.
//
// Operands.  Everything the tensor core reads is staged in shared memory in the K-major no-swizzle layout
// ([8-wide K plane][row][8] bf16: 16 bytes per row per plane), and every stage is ONE contiguous block of global
// memory in exactly that layout, so a stage is filled by plain bulk async copies (cp.async.bulk, no tensor maps):
//   * activations ("packed A"): [row tile][K chunk][K plane][128 rows][8]; the conv trunk writes its features in this
//     layout, and the epilogues of fc1 / fc2 write h1 / h2 in it (global scratch that never leaves the L2);
//   * weights ("packed B", built once on the host): [column slice c][K chunk][K plane][N/NS rows][8].
// The A block of a stage is needed by the NS CTAs of a row tile and the B block by the RT CTAs of a column slice: each
// CTA issues 1/NS of the A block and 1/RT of the B block with .multicast::cluster to the whole group, so every byte
// leaves the L2 once per cluster.  A stage's `full` barrier collects the bytes from all issuers; its `empty` barrier
// collects one tcgen05.commit (multicast) from every CTA the issuer sends to.
//
// Layer chaining.  fc1's accumulator (128 x N1/NS fp32) sits in TMEM columns [0, 256), fc2's in [256, 384), the heads'
// in [384, 512): no accumulator is reused, so the MMA warp never waits for an epilogue.  The epilogue warps turn a
// finished accumulator into bf16 blocks of the next layer's K-chunk width (bias + ReLU, BatchNorm folded on the
// host), store them as packed A blocks and signal a per-block `ready` barrier in all NS CTAs of the row tile
// (remote mbarrier arrive); the producer of the next layer waits for the blocks it is about to fetch.  The heads'
// epilogue adds the bias and writes the bf16 logits row that the tree step (az_tree.logits) or az_net_heads reads.
//
// Warp roles (19 warps): 0-15 epilogue (warp % 4 = TMEM lane quarter, warp / 4 = which 64-column block of the layer's
// output), 16 / 17 producers of the weight / activation blocks (one lane each: a bulk copy costs its issuing thread ~370
// cycles), 18 MMA issue.  What the cycle stamps of az_net_fcstack_debug showed, and what was tried and dropped
// (multicast, a cp.async loader path beside the bulk-copy unit, cluster-scope acquires in the producer), is in
// profiles/r2_fcstack_experiments.txt.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "fcstack_common.cuh"

// Operand ring: 192 KB cut into 4 slots of 48 KB for fc1 (A: 64 K x 128 rows = 16 KB, B: 64 K x 256 columns = 32 KB) and
// into 3 slots of 64 KB for fc2 / the heads, whose stages hold TWO 64-wide K chunks (A 32 KB + B up to 32 KB): half as
// many stage hand-overs (barrier waits, commits, bulk-copy issues -- ~370 cycles per copy for the issuing thread) per
// tcgen05.mma.  The two sets have their own barriers; fc2's first fill waits until fc1's last MMA has completed.
#define FS_STAGES 4
#define FS_A_BYTES 16384
#define FS_B_BYTES 32768
#define FS_STAGE_BYTES (FS_A_BYTES + FS_B_BYTES)
#define FS_STAGES2 3
#define FS_A2_BYTES 32768
#define FS_STAGE2_BYTES 65536
#define FS_EPI_WARPS 16                       // warps 0-15: epilogue (warp % 4 = TMEM lane quarter, warp / 4 = output block)
#define FS_PRODUCER_WARP FS_EPI_WARPS         // warp 16: bulk copies of the B (weight) blocks, one lane
#define FS_PRODUCER_A_WARP (FS_EPI_WARPS + 1) // warp 17: bulk copies of the A (activation) blocks, one lane.  Issuing a
                                              // bulk copy costs the issuing thread ~370 cycles (measured): one thread for
                                              // both copies of a stage made the producer the bottleneck of every layer
#define FS_MMA_WARP (FS_EPI_WARPS + 2)        // warps 18, 19: tcgen05.mma issue (one elected lane each); 18 allocates TMEM.
                                              // One thread issues a tcgen05.mma every ~97 cycles (measured), which made
                                              // the issue the longest single item of a CTA: the two warps take alternate
                                              // stages and accumulate into their OWN accumulator (the epilogue adds the two)
#define FS_MMA_WARPS 2
#define FS_THREADS ((FS_EPI_WARPS + 2 + FS_MMA_WARPS) * 32)
#define FS_MAX_CHUNKS 16                      // K chunks per layer
#define FS_BIAS_FLOATS 512                    // 256 + 128 + 128
#define FB_FULL 0
#define FB_EMPTY (FB_FULL + FS_STAGES)
#define FB_FULL2 (FB_EMPTY + FS_STAGES)
#define FB_EMPTY2 (FB_FULL2 + FS_STAGES2)
#define FB_ACC (FB_EMPTY2 + FS_STAGES2)
#define FB_TFREE (FB_ACC + 3)                 // fc1's accumulators have been read: fc2's may overwrite their TMEM columns
#define FB_READY (FB_TFREE + 1)               // [2][FS_MAX_CHUNKS]: block kc of h1 / h2 of this row tile is in global memory
#define FB_COUNT (FB_READY + 2 * FS_MAX_CHUNKS)
#define FS_SM_BIAS (FS_STAGES * FS_STAGE_BYTES)
#define FS_SM_BARS (FS_SM_BIAS + FS_BIAS_FLOATS * 4)
#define FS_SM_TMEM_PTR (FS_SM_BARS + 8 * 56)
#define FS_SM_TOTAL (FS_SM_TMEM_PTR + 16)
static_assert(FB_COUNT <= 56, "barrier area");
static_assert(FS_STAGES2 * FS_STAGE2_BYTES <= FS_STAGES * FS_STAGE_BYTES, "both slot sets share the ring");
static_assert(FS_SM_TOTAL <= 232448, "shared memory per CTA on sm_100");

struct FcStackParams {
    const uint8_t* x;          // layer-1 input, packed A: [tiles][K0/kch0][kch0/8][128][8] bf16
    const uint8_t* wblob;      // packed B of the three layers, then their float32 biases (az_net_fcstack_plan)
    uint8_t* h1;               // scratch, packed A: [tiles][K1/kch1][kch1/8][128][8]
    uint8_t* h2;               // scratch, packed A: [tiles][K2/kch2][kch2/8][128][8]
    __nv_bfloat16* logits;     // [G][ld]
    long long G;
    int ld, n_out;             // row pitch of logits; real outputs (action_size + 1)
    int K[3], Nc[3], kch[3];   // per layer: K, output columns per CTA, K chunk (16 / 32 / 64)
    int RT, NS, hpc;           // cluster shape; real head outputs per column slice
    int mc_a, mc_b;            // 1: the A / B block of a stage is fetched once per cluster and multicast; 0: every CTA
                               //    fetches its own copy (the bytes INTO an SM are the same either way)
    uint32_t w_off[3], b_off[3];
    int late_trigger;          // 1: let the dependent kernel (the tree step) launch only when this grid has finished
    long long* debug;          // optional [n_ctas][32] cycle stamps (az_net_fcstack_debug), see tools/fcstack_probe.py
};

// Stage geometry of layer L, the same in every role: K chunks per stage, stage-iterations, slot size / count, offset of
// the B block inside a slot and the first full barrier of the layer's slot set (its empty barriers follow).
struct FsGeom {
    int chunks, iters, slots;
    uint32_t slot_bytes, b_off, bar_full, bar_empty;
};
__device__ __forceinline__ FsGeom fs_geom(const FcStackParams& p, int L, uint32_t bars) {
    FsGeom g;
    const int nch = p.K[L] / p.kch[L];
    if (L == 0) {
        g.chunks = 1; g.slots = FS_STAGES; g.slot_bytes = FS_STAGE_BYTES; g.b_off = FS_A_BYTES;
        g.bar_full = bars + 8 * FB_FULL; g.bar_empty = bars + 8 * FB_EMPTY;
    } else {
        g.chunks = (p.kch[L] == 64 && (nch & 1) == 0) ? 2 : 1;
        g.slots = FS_STAGES2; g.slot_bytes = FS_STAGE2_BYTES; g.b_off = FS_A2_BYTES;
        g.bar_full = bars + 8 * FB_FULL2; g.bar_empty = bars + 8 * FB_EMPTY2;
    }
    g.iters = nch / g.chunks;
    return g;
}

__global__ void __launch_bounds__(FS_THREADS, 1) k_net_fcstack(FcStackParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int RT = p.RT, NS = p.NS;
    long long* dbg = p.debug ? p.debug + (size_t)blockIdx.x * 32 : nullptr;
    const long long t_entry = dbg ? clock64() : 0;
    if (dbg && tid == 0) { unsigned long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt)); dbg[25] = (long long)gt; }
    uint32_t rank, cluster_id;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    asm("mov.u32 %0, %%clusterid.x;" : "=r"(cluster_id));
    const int r = (int)rank / NS, c = (int)rank % NS;
    const int m = (int)cluster_id * RT + r;                   // row tile of this CTA
    const uint32_t a_mask = ((1u << NS) - 1u) << (r * NS);    // the CTAs of row tile r
    uint32_t b_mask = 0u;                                     // the CTAs of column slice c
    for (int rr = 0; rr < RT; ++rr) b_mask |= 1u << (rr * NS + c);
    const bool mc_a = p.mc_a && NS > 1, mc_b = p.mc_b && RT > 1, mc_any = mc_a || mc_b;
    const uint32_t all_mask = (1u << rank) | (mc_a ? a_mask : 0u) | (mc_b ? b_mask : 0u);   // CTAs this one sends stage data to

    const uint32_t bars = smem_u32(smem + FS_SM_BARS);
    const uint32_t bar_acc0 = bars + 8 * FB_ACC;
    const uint32_t bar_ready0 = bars + 8 * FB_READY;
    float* bias_s = reinterpret_cast<float*>(smem + FS_SM_BIAS);
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + FS_SM_TMEM_PTR);
    const int bias_off[3] = {0, 256, 384};
    // TMEM columns of the two accumulators (one per MMA warp) of the three layers: fc1 fills all 512 columns, fc2 reuses
    // [0, 256) once fc1's epilogue has read them (FB_TFREE), the heads sit beside fc2's
    const uint32_t col_off[3][2] = {{0u, 256u}, {0u, 128u}, {256u, 288u}};
    const uint32_t bar_tfree = bars + 8 * FB_TFREE;

    // ---- setup: barriers, biases of this column slice (weights: not produced by the preceding kernel), TMEM
    if (tid < FB_COUNT) {      // one barrier per thread
        uint32_t count = 1u;
        if (tid >= FB_ACC && tid < FB_ACC + 3) {                                  // acc[L]: one commit per MMA warp with work
            const int L = tid - FB_ACC, nch = p.K[L] / p.kch[L];
            const int iters = (L > 0 && p.kch[L] == 64 && (nch & 1) == 0) ? nch / 2 : nch;
            count = iters >= 2 ? 2u : 1u;
        }
        if (tid == FB_TFREE) count = FS_EPI_WARPS;
        if (tid < FB_EMPTY || (tid >= FB_FULL2 && tid < FB_EMPTY2)) count = 1u;   // full[s]: the producer's expect_tx
        else if (tid < FB_ACC) count = (uint32_t)__popc(all_mask);                // empty[s]: one commit per receiving CTA
        else if (tid >= FB_READY) count = 4u;                                     // ready[block]: the four quarter warps
        mbar_init(bars + 8 * tid, count);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == FS_MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    fs_cluster_sync();            // every CTA's barriers exist before any multicast copy / remote arrive targets them
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr_smem;
    az_pdl_wait();                // the layer-1 input is the preceding kernel's output
    if (!p.late_trigger) az_pdl_trigger();
    if (dbg && tid == 0) dbg[1] = clock64() - t_entry;

    if (warp == FS_PRODUCER_WARP) {
        // ================================================================== producer of the weight blocks (one lane)
        if (lane == 0) {
            uint32_t it = 0;
            long long w_empty = 0, w_ready = 0;
#pragma unroll
            for (int L = 0; L < 3; ++L) {
                const FsGeom g = fs_geom(p, L, bars);
                if (L == 1) { it = 0; fs_mbar_wait(bar_acc0, 0u); }      // the 64 KB slots overlay fc1's: fc1 must be done
                const uint32_t a_bytes = (uint32_t)(g.chunks * p.kch[L]) * 256u;
                const uint32_t b_bytes = (uint32_t)(g.chunks * p.kch[L]) * (uint32_t)p.Nc[L] * 2u;
                // with multicast this CTA fetches part r of the B block for its whole column group
                const uint32_t b_part = mc_b ? b_bytes / (uint32_t)RT : b_bytes, b_skip = mc_b ? (uint32_t)r * b_part : 0u;
                const uint8_t* b_src = p.wblob + p.w_off[L] + (size_t)c * g.iters * b_bytes + b_skip;
                for (int q = 0; q < g.iters; ++q, ++it) {
                    const uint32_t s = it % (uint32_t)g.slots, ph = (it / (uint32_t)g.slots) & 1u;
                    const long long c0 = dbg ? clock64() : 0;
                    fs_mbar_wait(g.bar_empty + 8 * s, ph ^ 1u);
                    const long long c1 = dbg ? clock64() : 0;
                    const uint32_t full = g.bar_full + 8 * s;
                    fs_expect_tx(full, a_bytes + b_bytes);      // the A bytes come from the other producer warp
                    fs_bulk_copy(smem_u32(smem) + s * g.slot_bytes + g.b_off + b_skip, b_src + (size_t)q * b_bytes, b_part,
                                 full, b_mask, mc_b);
                    if (dbg) { w_empty += c1 - c0; w_ready += clock64() - c1; if (q == 0) dbg[26 + L] = c0 - t_entry; }
                }
            }
            if (dbg) { dbg[20] = w_empty; dbg[21] = w_ready; }
        }
        __syncwarp();
    } else if (warp == FS_PRODUCER_A_WARP) {
        // ================================================================== producer of the activation blocks (one lane)
        if (lane == 0) {
            uint32_t it = 0;
#pragma unroll
            for (int L = 0; L < 3; ++L) {
                const FsGeom g = fs_geom(p, L, bars);
                if (L == 1) { it = 0; fs_mbar_wait(bar_acc0, 0u); }
                const uint32_t a_bytes = (uint32_t)(g.chunks * p.kch[L]) * 256u;
                const uint32_t a_part = mc_a ? a_bytes / (uint32_t)NS : a_bytes, a_skip = mc_a ? (uint32_t)c * a_part : 0u;
                const uint8_t* a_src = (L == 0 ? p.x : (L == 1 ? p.h1 : p.h2)) + (size_t)m * g.iters * a_bytes + a_skip;
                for (int q = 0; q < g.iters; ++q, ++it) {
                    const uint32_t s = it % (uint32_t)g.slots, ph = (it / (uint32_t)g.slots) & 1u;
                    fs_mbar_wait(g.bar_empty + 8 * s, ph ^ 1u);
                    // h1 / h2 blocks in ascending K order (all blocks of column slice 0, then slice 1, ...).  The writers
                    // made their stores visible at cluster scope before the (release) arrive that completes a block's
                    // barrier, and the copy below reads the L2, so observing the barrier is enough: a cluster-scope
                    // acquire + proxy fence here cost ~540 cycles per stage (profiles/r2_fcstack_experiments.txt).
                    if (L > 0)
                        for (int i = 0; i < g.chunks; ++i)
                            fs_mbar_wait(bar_ready0 + 8 * ((L - 1) * FS_MAX_CHUNKS + q * g.chunks + i), 0u);
                    fs_bulk_copy(smem_u32(smem) + s * g.slot_bytes + a_skip, a_src + (size_t)q * a_bytes, a_part,
                                 g.bar_full + 8 * s, a_mask, mc_a);
                }
            }
        }
        __syncwarp();
    } else if (warp >= FS_MMA_WARP) {
        // ================================================================== MMA issue: warp w takes the stages q with q % 2 == w
        const int w = warp - FS_MMA_WARP;
        uint32_t it0 = 0;
        long long w_full = 0, t_issue = 0;
#pragma unroll
        for (int L = 0; L < 3; ++L) {
            const FsGeom g = fs_geom(p, L, bars);
            if (L == 1) {
                it0 = 0;
                fs_mbar_wait(bar_tfree, 0u);                  // fc1's accumulators have been read out of [0, 256)
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            }
            const int ksteps = g.chunks * (p.kch[L] >> 4);
            const uint32_t nc = (uint32_t)p.Nc[L];
            const uint32_t idesc = umma_idesc_bf16_m128(nc);
            const uint32_t d_tmem = tmem_base + (w ? col_off[L][1] : col_off[L][0]);
            for (int q = w; q < g.iters; q += FS_MMA_WARPS) {
                const uint32_t it = it0 + (uint32_t)q;
                const uint32_t s = it % (uint32_t)g.slots, ph = (it / (uint32_t)g.slots) & 1u;
                const long long c0 = dbg ? clock64() : 0;
                fs_mbar_wait(g.bar_full + 8 * s, ph);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const long long c1 = dbg ? clock64() : 0;
                if (dbg && lane == 0 && w == 0) { w_full += c1 - c0; if (q == 0) dbg[2 + L * 6] = c1 - t_entry; }
                if (elect_one()) {
                    const uint32_t sa = smem_u32(smem) + s * g.slot_bytes, sb = sa + g.b_off;
                    for (int k = 0; k < ksteps; ++k) {
                        const uint64_t ad = umma_desc(sa + (uint32_t)k * 4096u, 2048u, 128u);
                        const uint64_t bd = umma_desc(sb + (uint32_t)k * 32u * nc, 16u * nc, 128u);
                        umma_bf16(d_tmem, ad, bd, idesc, (q >= FS_MMA_WARPS || k) ? 1u : 0u);
                    }
                    if (mc_any) fs_commit_multicast(g.bar_empty + 8 * s, all_mask);
                    else umma_commit(g.bar_empty + 8 * s);
                    if (q + FS_MMA_WARPS >= g.iters) umma_commit(bar_acc0 + 8 * L);     // this warp's last stage of the layer
                }
                __syncwarp();
                if (dbg && lane == 0 && w == 0) { const long long c2 = clock64(); t_issue += c2 - c1; if (q + FS_MMA_WARPS >= g.iters) dbg[3 + L * 6] = c2 - t_entry; }
            }
            it0 += (uint32_t)g.iters;
        }
        if (dbg && lane == 0 && w == 0) { dbg[22] = w_full; dbg[23] = t_issue; }
    } else {
        // ================================================================== epilogue
        const int quarter = warp & 3, blk = warp >> 2;        // blk: which 64-wide output block of a layer this warp converts
        const int row = quarter * 32 + lane;                  // TMEM lane = row of the tile
        const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        {   // biases of this column slice -> shared memory, while the first layer's operands stream in
            const int n0 = p.Nc[0], n1 = p.Nc[1], n2 = p.Nc[2];
            for (int i = tid; i < n0 + n1 + n2; i += FS_EPI_WARPS * 32) {
                const int L = i < n0 ? 0 : (i < n0 + n1 ? 1 : 2);
                const int j = i - (L == 0 ? 0 : (L == 1 ? n0 : n0 + n1));
                const int nc = L == 0 ? n0 : (L == 1 ? n1 : n2);
                const uint32_t off = L == 0 ? p.b_off[0] : (L == 1 ? p.b_off[1] : p.b_off[2]);
                bias_s[(L == 0 ? 0 : (L == 1 ? 256 : 384)) + j] = reinterpret_cast<const float*>(p.wblob + off)[c * nc + j];
            }
            asm volatile("bar.sync 1, %0;" ::"r"(FS_EPI_WARPS * 32) : "memory");
        }
        uint32_t v[64];
#pragma unroll
        for (int i = 0; i < 64; ++i) v[i] = 0u;
#pragma unroll
        for (int L = 0; L < 3; ++L) {
            fs_mbar_wait(bar_acc0 + 8 * L, 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (dbg && tid == 0) dbg[4 + L * 6] = clock64() - t_entry;
            if (L < 2) {
                const int bw = p.kch[L + 1];                  // block width = the next layer's K chunk
                const int nblk = p.Nc[L] / bw, nchn = p.K[L + 1] / bw;
                uint8_t* hbase = (L == 0 ? p.h1 : p.h2) + (size_t)m * nchn * ((size_t)bw * 256) + (size_t)row * 16;
                const bool two = fs_geom(p, L, bars).iters >= 2;        // both MMA warps accumulated: add their halves
                for (int j = blk; j < nblk; j += FS_EPI_WARPS / 4) {
                    const int kcn = c * nblk + j;             // this block's K chunk in the next layer
                    uint8_t* dst = hbase + (size_t)kcn * ((size_t)bw * 256);
#pragma unroll
                    for (int hb = 0; hb < 2; ++hb) {          // 32 columns at a time: 32 + 32 accumulator registers
                        if (hb * 32 < bw) {
                            const uint32_t col = (uint32_t)(j * bw + hb * 32);
                            fs_tmem_ld16(t_lane + col_off[L][0] + col, &v[0]);
                            fs_tmem_ld16(t_lane + col_off[L][0] + col + 16u, &v[16]);
                            if (two) {
                                fs_tmem_ld16(t_lane + col_off[L][1] + col, &v[32]);
                                fs_tmem_ld16(t_lane + col_off[L][1] + col + 16u, &v[48]);
                            }
                            fs_tmem_wait64(v);
                            if (dbg && tid == 0 && hb == 0) dbg[29 + L] = clock64() - t_entry;
                            const float* bs = bias_s + bias_off[L] + j * bw + hb * 32;
#pragma unroll
                            for (int pl = 0; pl < 4; ++pl) {
                                const float4 b0 = *reinterpret_cast<const float4*>(bs + pl * 8);
                                const float4 b1 = *reinterpret_cast<const float4*>(bs + pl * 8 + 4);
                                float x[8];
#pragma unroll
                                for (int e = 0; e < 8; ++e)
                                    x[e] = __uint_as_float(v[pl * 8 + e]) + (two ? __uint_as_float(v[32 + pl * 8 + e]) : 0.f);
                                uint4 o;
                                o.x = pack_relu_bf16x2(x[0] + b0.x, x[1] + b0.y);
                                o.y = pack_relu_bf16x2(x[2] + b0.z, x[3] + b0.w);
                                o.z = pack_relu_bf16x2(x[4] + b1.x, x[5] + b1.y);
                                o.w = pack_relu_bf16x2(x[6] + b1.z, x[7] + b1.w);
                                *reinterpret_cast<uint4*>(dst + (hb * 4 + pl) * 2048) = o;
                            }
                        }
                    }
                    if (dbg && tid == 0) dbg[6 + L * 6] = clock64() - t_entry;
                    // The block is read by bulk copies (async proxy) of all CTAs of the row tile: every storing lane's
                    // proxy fence, then the release of the remote arrive -- cumulative over the warp's stores through
                    // __syncwarp -- publishes it.  (A fence.acq_rel.cluster per lane on top of that cost ~1400 cycles per
                    // block and is redundant.)
                    asm volatile("fence.proxy.async.global;" ::: "memory");
                    __syncwarp();
                    if (lane < NS) fs_mbar_arrive_remote(bar_ready0 + 8 * (L * FS_MAX_CHUNKS + kcn), (uint32_t)(r * NS + lane));
                    if (dbg && tid == 0) dbg[7 + L * 6] = clock64() - t_entry;
                }
                if (L == 0) {     // this warp no longer reads fc1's accumulators
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) mbar_arrive(bar_tfree);
                }
            } else if (blk == 0) {
                // heads: bias, bf16, then through shared memory so that the global stores are row-contiguous 4-byte words
                // (lane = row stores of 2 bytes at a 144-byte pitch cost one L1 wavefront per lane: 3 800 cycles measured)
                const int ncols = p.Nc[2], hpc = p.hpc, pw = hpc >> 1;      // hpc is even (fc_plan)
                const bool two = fs_geom(p, 2, bars).iters >= 2;
#pragma unroll
                for (int qq = 0; qq < 2; ++qq)
                    if (qq * 16 < ncols) {
                        fs_tmem_ld16(t_lane + col_off[2][0] + (uint32_t)(qq * 16), &v[qq * 16]);
                        if (two) fs_tmem_ld16(t_lane + col_off[2][1] + (uint32_t)(qq * 16), &v[32 + qq * 16]);
                    }
                fs_tmem_wait64(v);
                if (two) {
#pragma unroll
                    for (int e = 0; e < 32; ++e) v[e] = __float_as_uint(__uint_as_float(v[e]) + __uint_as_float(v[32 + e]));
                }
                if (dbg && tid == 0) dbg[31] = clock64() - t_entry;
                uint32_t* stage = reinterpret_cast<uint32_t*>(smem);         // stage 0 is free: every MMA has completed
                const float* bs = bias_s + bias_off[2];
#pragma unroll
                for (int k = 0; k < 16; ++k)
                    if (k < pw) {
                        __nv_bfloat162 two = __floats2bfloat162_rn(__uint_as_float(v[2 * k]) + bs[2 * k],
                                                                   __uint_as_float(v[2 * k + 1]) + bs[2 * k + 1]);
                        stage[row * 17 + k] = *reinterpret_cast<uint32_t*>(&two);   // odd pitch: no bank conflicts
                    }
                asm volatile("bar.sync 2, 128;" ::: "memory");
                const long long row0 = (long long)m * 128;
                for (int idx = tid; idx < 128 * pw; idx += 128) {
                    const int rr = idx / pw, k = idx - rr * pw;
                    const int h = c * hpc + 2 * k;
                    if (row0 + rr < p.G && h < p.n_out) {
                        __nv_bfloat16* out = p.logits + (row0 + rr) * p.ld + h;
                        const uint32_t w = stage[rr * 17 + k];
                        if (h + 1 < p.n_out) *reinterpret_cast<uint32_t*>(out) = w;
                        else *out = __ushort_as_bfloat16((unsigned short)(w & 0xffffu));
                    }
                }
            }
            if (dbg && tid == 0) dbg[5 + L * 6] = clock64() - t_entry;
        }
    }

    // ---- teardown: no CTA leaves while a peer may still signal its barriers or copy into its shared memory
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    fs_cluster_sync();
    if (dbg && tid == 0) {
        dbg[24] = clock64() - t_entry;
        unsigned long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt)); dbg[19] = (long long)gt;
    }
    if (warp == FS_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

// ------------------------------------------------------------------------------------------------ C ABI

static int fc_chunk(int k) { return k >= 64 ? 64 : k; }

// Shapes of the three layers for a network K1 -> N1 -> N2 -> n_out with NS column slices; false when the kernel has
// no mapping for them.
static bool fc_plan2(int K1, int N1, int N2, int n_out, FcPlan* pl);

static bool fc_plan(int K1, int N1, int N2, int n_out, int RT, int NS, FcPlan* pl) {
    if (NS < 1 || NS > 8 || RT < 1 || RT * NS > 8 || K1 < 16 || N1 < 16 || N2 < 16 || n_out < 1) return false;
    if ((N1 % NS) || (N2 % NS)) return false;
    static const bool v1_only = getenv("AZ_B200_FCSTACK_V1") != nullptr;
    if (RT == 1 && NS == 4 && !v1_only && fc_plan2(K1, N1, N2, n_out, pl)) return true;
    // RT = 2: the same shapes on CTA pairs, if the partial sums of a row tile fit its share of the h2 scratch
    if (RT == 2 && NS == 4 && !v1_only && fc_plan2(K1, N1, N2, n_out, pl)) {
        if (12 * (pl->Nc[2] / 16) * 2048 + 256 <= 128 * N2 * 2) { pl->variant = 3; return true; }
    }
    pl->variant = 1;
    pl->hpc = (((n_out + NS - 1) / NS) + 1) & ~1;      // even: the heads' epilogue stores bf16 pairs
    pl->n3c = ((pl->hpc + 15) / 16) * 16;
    pl->K[0] = K1; pl->K[1] = N1; pl->K[2] = N2;
    pl->Nc[0] = N1 / NS; pl->Nc[1] = N2 / NS; pl->Nc[2] = pl->n3c;
    pl->kch[0] = fc_chunk(K1); pl->kch[1] = fc_chunk(pl->Nc[0]); pl->kch[2] = fc_chunk(pl->Nc[1]);
    const int max_nc[3] = {256, 128, 32};
    int64_t off = 0;
    for (int L = 0; L < 3; ++L) {
        const int kch = pl->kch[L];
        if ((kch % 16) || (pl->K[L] % kch) || pl->K[L] / kch > FS_MAX_CHUNKS) return false;
        if ((pl->Nc[L] % 16) || pl->Nc[L] > max_nc[L]) return false;
        if (L > 0 && (pl->Nc[L - 1] % kch)) return false;        // whole blocks per producing CTA
        if ((kch * 256 / NS) % 16) return false;                 // bulk-copy granularity of an A part
        pl->w_off[L] = (uint32_t)off;
        off += (int64_t)NS * pl->Nc[L] * pl->K[L] * 2;
    }
    for (int L = 0; L < 3; ++L) { pl->b_off[L] = (uint32_t)off; off += (int64_t)NS * pl->Nc[L] * 4; }
    pl->blob_bytes = off;
    return true;
}

// Variant 2 (k_net_fcstack2): clusters of 4 column slices, fc1 in two passes of PW columns, the heads split over K.
// Weights: fc1 [8 pass-slices][K1/64][8][PW][8], fc2 [4][N1/64][8][Nc2][8], heads [4 K slices][Nc2/8][n3][8]; biases
// in natural order (the heads' zero padded to n3).
static bool fc_plan2(int K1, int N1, int N2, int n_out, FcPlan* pl) {
    const int NS = 4;
    if ((K1 % 64) || K1 > 512 || (N1 % (2 * NS * 64)) || N1 / NS > 256 || (N2 % (NS * 32)) || N2 / NS > 128) return false;
    const int n3 = ((n_out + 15) / 16) * 16;
    if (n3 > 80) return false;
    pl->variant = 2;
    pl->K[0] = K1; pl->K[1] = N1; pl->K[2] = N2;
    pl->Nc[0] = N1 / NS / 2; pl->Nc[1] = N2 / NS; pl->Nc[2] = n3;
    pl->kch[0] = 64; pl->kch[1] = 64; pl->kch[2] = pl->Nc[1];
    pl->hpc = n_out; pl->n3c = n3;
    int64_t off = 0;
    pl->w_off[0] = (uint32_t)off; off += (int64_t)N1 * K1 * 2;
    pl->w_off[1] = (uint32_t)off; off += (int64_t)N2 * N1 * 2;
    pl->w_off[2] = (uint32_t)off; off += (int64_t)n3 * N2 * 2;
    pl->b_off[0] = (uint32_t)off; off += (int64_t)N1 * 4;
    pl->b_off[1] = (uint32_t)off; off += (int64_t)N2 * 4;
    pl->b_off[2] = (uint32_t)off; off += (int64_t)n3 * 4;
    pl->blob_bytes = off;
    return true;
}

// out[0..2] K chunk per layer, out[3..5] columns per CTA per layer, out[6] real head outputs per slice,
// out[7..9] byte offsets of the packed weights, out[10..12] of the float32 biases, out[13] blob bytes, out[14] the
// kernel variant the blob must be packed for (1: k_net_fcstack, 2: k_net_fcstack2 -- see FcPlan)
extern "C" int az_net_fcstack_plan(int K1, int N1, int N2, int n_out, int RT, int NS, int64_t* out) {
    FcPlan pl;
    if (!out || !fc_plan(K1, N1, N2, n_out, RT, NS, &pl)) return AZ_ERR_ARG;
    for (int L = 0; L < 3; ++L) {
        out[L] = pl.kch[L]; out[3 + L] = pl.Nc[L]; out[7 + L] = pl.w_off[L]; out[10 + L] = pl.b_off[L];
    }
    out[6] = pl.hpc; out[13] = pl.blob_bytes; out[14] = pl.variant;
    return AZ_OK;
}

static long long* g_fcstack_debug = nullptr;
// Measurement hook: launches after this call write per-CTA cycle stamps to buf ([grid][32] int64, device); NULL = off.
extern "C" int az_net_fcstack_debug(void* buf) { g_fcstack_debug = (long long*)buf; return AZ_OK; }

extern "C" int az_net_fcstack(const void* x, const void* wblob, void* h1, void* h2, void* logits, int ld, int n_out,
                              int64_t G, int K1, int N1, int N2, int RT, int NS, void* stream) {
    FcPlan pl;
    if (!fc_plan(K1, N1, N2, n_out, RT, NS, &pl) || ld < n_out || G < 0 ||
        (G > 0 && (!x || !wblob || !h1 || !logits)))
        return AZ_ERR_ARG;
    if (pl.variant == 1 && G > 0 && !h2) return AZ_ERR_ARG;
    if (pl.variant >= 2 && (ld & 1)) return AZ_ERR_ARG;      // the logits rows are stored as 4-byte words
    if (pl.variant == 3 && G > 0 && !h2) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    static const int late_trigger = getenv("AZ_B200_FCSTACK_LATE_TRIGGER") ? atoi(getenv("AZ_B200_FCSTACK_LATE_TRIGGER")) : 0;
    static const bool no_pdl = getenv("AZ_B200_NO_PDL") != nullptr;
    if (pl.variant == 3)
        return az_fcstack3_launch(pl, x, wblob, h1, h2, logits, ld, n_out, G, g_fcstack_debug, late_trigger, !no_pdl, stream);
    if (pl.variant == 2)
        return az_fcstack2_launch(pl, x, wblob, h1, logits, ld, n_out, G, g_fcstack_debug, late_trigger, !no_pdl, stream);
    static bool attr_set = false;
    if (!attr_set) {
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_fcstack, cudaFuncAttributeMaxDynamicSharedMemorySize, FS_SM_TOTAL));
        attr_set = true;
    }
    FcStackParams p;
    p.x = (const uint8_t*)x; p.wblob = (const uint8_t*)wblob; p.h1 = (uint8_t*)h1; p.h2 = (uint8_t*)h2;
    p.logits = (__nv_bfloat16*)logits; p.G = G; p.ld = ld; p.n_out = n_out;
    for (int L = 0; L < 3; ++L) {
        p.K[L] = pl.K[L]; p.Nc[L] = pl.Nc[L]; p.kch[L] = pl.kch[L]; p.w_off[L] = pl.w_off[L]; p.b_off[L] = pl.b_off[L];
    }
    p.RT = RT; p.NS = NS; p.hpc = pl.hpc; p.debug = g_fcstack_debug;
    p.late_trigger = late_trigger;
    static const int mc_flags = getenv("AZ_B200_FCSTACK_MC") ? atoi(getenv("AZ_B200_FCSTACK_MC")) : 0;   // bit 0: A, bit 1: B
    p.mc_a = mc_flags & 1; p.mc_b = (mc_flags >> 1) & 1;
    const int64_t tiles = (G + 127) / 128, clusters = (tiles + RT - 1) / RT;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(clusters * RT * NS)); cfg.blockDim = dim3(FS_THREADS);
    cfg.dynamicSmemBytes = FS_SM_TOTAL; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)(RT * NS); attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = no_pdl ? 1 : 2;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_net_fcstack, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_fcstack");
    AZ_CHECK_LAUNCH("k_net_fcstack");
    return AZ_OK;
}

// ------------------------------------------------------------------------------------------------ no-allocation variant

// az_net_fcstack with the h1 / h2 scratch in a file-scope device array instead of caller-allocated buffers, for callers
// that cannot allocate device memory (code under CUDA graph capture, serving loops, harnesses that forbid it).  No
// counterpart in the reference.  The array is sized for
// AZ_FCSTACK_STATIC_MAX_ROWS positions of OthelloNet's stack (N1 + N2 = 1536 columns per row): 6 MB that every
// GPU which loads the library holds until the process exits.
#define AZ_FCSTACK_STATIC_MAX_ROWS 2048
#define AZ_FCSTACK_STATIC_BYTES ((size_t)AZ_FCSTACK_STATIC_MAX_ROWS * 1536 * 2)
static __device__ __align__(256) uint8_t g_fcstack_static_scratch[AZ_FCSTACK_STATIC_BYTES];

// In host code the array's name is only a handle: its device address comes from cudaGetSymbolAddress, asked once per
// device with that device current.
static int fcstack_static_scratch(uint8_t** out) {
    static uint8_t* addr[64] = {};
    int dev = 0;
    AZ_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64) return az_set_error_message(AZ_ERR_ARG, "az_net_fcstack_static: device ordinal beyond 63");
    if (!addr[dev]) {
        void* ptr = nullptr;
        AZ_CUDA_TRY(cudaGetSymbolAddress(&ptr, g_fcstack_static_scratch));
        addr[dev] = (uint8_t*)ptr;
    }
    *out = addr[dev];
    return AZ_OK;
}

extern "C" int az_net_fcstack_static(const void* x, const void* wblob, void* logits, int ld, int n_out, int64_t G,
                                     int K1, int N1, int N2, int RT, int NS, void* stream) {
    if (G < 0 || RT < 1 || N1 < 1 || N2 < 1) return AZ_ERR_ARG;
    const int64_t tiles = ((G + 127) / 128 + RT - 1) / RT * RT;              // whole clusters, like the caller's x
    const int64_t h1_bytes = tiles * 128 * (int64_t)N1 * 2, h2_bytes = tiles * 128 * (int64_t)N2 * 2;
    if (h1_bytes + h2_bytes > (int64_t)AZ_FCSTACK_STATIC_BYTES) {
        char msg[320];
        snprintf(msg, sizeof(msg),
                 "az_net_fcstack_static: %lld positions of a %d -> %d -> %d stack need %lld bytes of scratch, the static array "
                 "holds %lld (%d positions at N1 + N2 = 1536); use az_net_fcstack with caller-allocated h1 / h2",
                 (long long)G, K1, N1, N2, (long long)(h1_bytes + h2_bytes), (long long)AZ_FCSTACK_STATIC_BYTES,
                 AZ_FCSTACK_STATIC_MAX_ROWS);
        return az_set_error_message(AZ_ERR_CAPACITY, msg);
    }
    if (G == 0) return AZ_OK;
    uint8_t* scratch = nullptr;
    const int rc = fcstack_static_scratch(&scratch);
    if (rc != AZ_OK) return rc;
    // h1 / h2 blocks are written in this launch before they are read; the only words that must start at zero are the
    // counters of the pair kernel (variant 3: the last 256 bytes of every row tile's h2 region), and an earlier call of
    // another shape may have left activations there: clear them in stream order (a memset node, no allocation)
    FcPlan pl;
    if (fc_plan(K1, N1, N2, n_out, RT, NS, &pl) && pl.variant == 3) {
        const size_t tile_bytes = (size_t)128 * N2 * 2;
        AZ_CUDA_TRY(cudaMemset2DAsync(scratch + h1_bytes + tile_bytes - 256, tile_bytes, 0, 256, (size_t)tiles,
                                      (cudaStream_t)stream));
    }
    return az_net_fcstack(x, wblob, scratch, scratch + h1_bytes, logits, ld, n_out, G, K1, N1, N2, RT, NS, stream);
}
