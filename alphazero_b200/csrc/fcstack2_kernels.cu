// The fully connected stack of the policy-value networks, second kernel (see fcstack_kernels.cu for the layers and
// the operand layouts; alphazero/games/othello.py:352-362, :374-382).  Same work, same cluster of 4 column slices per
// row tile of 128 positions, but the layer boundaries -- 10 400 of the first kernel's 35 400 cycles
// (profiles/r2_fcstack_experiments.txt) -- are taken off the critical path:
//
//   fc1   The CTA's 2 x PW output columns are computed in two passes of PW columns, each streaming the row tile's input
//         and the pass's weights.  The epilogue of pass 0 -- TMEM -> bias / ReLU / bf16 -> h1 blocks in global scratch
//         (L2) -> remote `ready` arrives in the four CTAs of the row tile -- runs UNDER the MMAs of pass 1.  (Keeping the
//         input resident in shared memory instead of streaming it twice was tried first: it leaves 96 KB of ring, and a
//         stage's round trip -- empty -> copy issued -> landed -> MMAs -> commit -- is ~3 000 cycles whatever its size,
//         so the bytes in flight set the rate: 96 KB gave 26 B/clk into the SM against 64 B/clk with 192 KB.)
//   fc2   starts on the pass-0 blocks of h1 (own CTA's first) the moment fc1's last MMA is issued, and reaches the
//         pass-1 blocks after their epilogue has long finished.
//   heads split over K instead of over the outputs: every CTA multiplies ITS OWN 128 columns of h2 -- written by its fc2
//         epilogue straight into shared memory as the A operand, no L2 round trip, no peer to wait for -- with the
//         matching K slice of the heads' weights for ALL A + 1 outputs.  The four fp32 partial sums of a row meet in the
//         CTA that owns the row's quarter of the tile (rows 32 q .. 32 q + 31 -> CTA q): st.shared::cluster into its
//         exchange buffer, one cluster barrier (which is also the kernel's teardown barrier), then the owner adds the
//         three partials to its registers, adds the bias and writes the bf16 logits row-contiguously.  h2 never leaves
//         the SM.
//
// One operand ring for the whole kernel: 3 slots of 64 KB (A blocks at +0, B blocks at +32 KB; two 64-wide K chunks per
// stage), stages 0 .. n1-1 fc1, then fc2, then one stage that carries the heads' weights; h2 and the staged logits rows
// take the two slots after it.  The exchange buffer has its own 30 KB: a peer may finish its heads while this CTA's fc2
// is still reading the ring.
//
// TMEM columns: fc1 pass 0 [0, PW), pass 1 [PW, 2 PW), fc2 [256, 256 + Nc2), heads [384, 384 + n3): nothing is reused, the
// MMA warp never waits for an epilogue.
//
// Warps: 0-15 epilogue (warp % 4 = TMEM lane quarter, warp / 4 = 32-column group), 16 weight producer, 17 activation
// producer (one lane each: issuing a bulk copy costs its thread ~370 cycles), 18 MMA issue.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "fcstack_common.cuh"

#define F2_NS 4
#define F2_SLOT 65536u                         // A block(s) at +0 (<= 32 KB), B block(s) at +32 KB
#define F2_B_OFF 32768u
#define F2_SLOTS 3
#define F2_XCHG_OFF (F2_SLOTS * F2_SLOT)       // [3 senders][5 column groups][4][32 lanes][16 B] fp32 partial logits = 30 KB
#define F2_XCHG_BYTES (3 * 5 * 2048)
#define F2_EPI_WARPS 16
#define F2_PRODUCER_B_WARP 16
#define F2_PRODUCER_A_WARP 17
#define F2_MMA_WARP 18
#define F2_THREADS (19 * 32)
#define F2_MAX_N3 80
#define B2_FULL 0                              // [3]  stage landed (A + B bytes)
#define B2_EMPTY 3                             // [3]  stage consumed (tcgen05.commit)
#define B2_ACC 6                               // [4]  fc1 pass 0, pass 1, fc2, heads accumulator complete
#define B2_H2 10                               // h2 is in shared memory (16 epilogue warps)
#define B2_READY 11                            // [16] block kc of h1 of this row tile is in global memory
#define B2_COUNT 27
#define F2_SM_BIAS (F2_XCHG_OFF + F2_XCHG_BYTES)
#define F2_SM_BARS (F2_SM_BIAS + 512 * 4)
#define F2_SM_TMEM_PTR (F2_SM_BARS + 8 * 32)
#define F2_SM_TOTAL (F2_SM_TMEM_PTR + 16)
static_assert(F2_SM_TOTAL <= 232448, "shared memory per CTA on sm_100");

struct Fc2Params {
    const uint8_t* x;          // layer-1 input, packed A: [tiles][K1/64][8][128][8] bf16
    const uint8_t* wblob;      // packed weights of the three layers, then their float32 biases (az_net_fcstack_plan, variant 2)
    uint8_t* h1;               // scratch, packed A: [tiles][N1/64][8][128][8]
    __nv_bfloat16* logits;     // [G][ld]
    long long G;
    int ld, n_out;
    int nkx, cps;              // K1 / 64; 64-wide K chunks per fc1 stage (2 when nkx is even)
    int PW, Nc2, n3;           // fc1 columns per pass, fc2 columns per CTA, padded head outputs
    int nch2, cps2;            // N1 / 64 = K chunks of fc2; chunks per fc2 stage (2 when a pass yields 2 blocks per CTA)
    uint32_t w_off[3], b_off[3];
    int late_trigger;
    long long* debug;          // optional [n_ctas][32] cycle stamps
};

// K chunk of fc2's q-th chunk in CTA c: the pass-0 blocks of h1 first, the CTA's own before its peers'.  With two
// blocks per pass (bpp = 2) chunks 2i and 2i + 1 are neighbours in h1 and in the packed weights: one stage, one copy.
__device__ __forceinline__ int f2_kc(int q, int c, int bpp) {
    const int per = F2_NS * bpp, h = q / per, rem = q - h * per;
    const int cs = ((rem / bpp) + c) & (F2_NS - 1), j = rem % bpp;
    return cs * 2 * bpp + h * bpp + j;
}

__global__ void __launch_bounds__(F2_THREADS, 1) k_net_fcstack2(Fc2Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    long long* dbg = p.debug ? p.debug + (size_t)blockIdx.x * 32 : nullptr;
    const long long t_entry = dbg ? clock64() : 0;
    if (dbg && tid == 0) { unsigned long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt)); dbg[25] = (long long)gt; }
    uint32_t rank, cluster_id;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    asm("mov.u32 %0, %%clusterid.x;" : "=r"(cluster_id));
    const int c = (int)rank, m = (int)cluster_id;             // column slice, row tile
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = smem_u32(smem + F2_SM_BARS);
    float* bias_s = reinterpret_cast<float*>(smem + F2_SM_BIAS);          // fc1 [0, 256), fc2 [256, 384), heads [384, 512)
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + F2_SM_TMEM_PTR);
    const int PW = p.PW, Nc2 = p.Nc2, n3 = p.n3, bpp = PW >> 6;
    const int n1p = p.nkx / p.cps, n1 = 2 * n1p;              // fc1 stages per pass / in total
    const int n2 = p.nch2 / p.cps2;                           // fc2 stages
    // One ring for the whole kernel: stages 0 .. n1-1 fc1, n1 .. n1+n2-1 fc2, then stage n1+n2 = the heads' weights;
    // the slot after that takes h2 (the heads' A operand), the one after the staged logits rows.
    const uint32_t it_heads = (uint32_t)(n1 + n2);
    const uint32_t slot_hb = it_heads % F2_SLOTS, slot_h2 = (it_heads + 1u) % F2_SLOTS, slot_out = (it_heads + 2u) % F2_SLOTS;

    // ---- setup
    if (tid < B2_COUNT) {
        uint32_t count = 1u;
        if (tid == B2_H2) count = F2_EPI_WARPS;
        if (tid >= B2_READY) count = 8u;                      // 4 lane quarters x 2 halves of a 64-column block
        mbar_init(bars + 8 * tid, count);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == F2_MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    fs_cluster_sync();            // every CTA's barriers exist before a peer's remote arrive targets them
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr_smem;
    az_pdl_wait();                // X is the preceding kernel's output
    if (!p.late_trigger) az_pdl_trigger();
    if (dbg && tid == 0) dbg[1] = clock64() - t_entry;

    uint32_t v[32];               // epilogue registers; the heads' partial sums stay in them across the cluster barrier
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = 0u;
    const int quarter = warp & 3, blk = warp >> 2;
    const int row = quarter * 32 + lane;                      // TMEM lane = row of the tile

    if (warp == F2_PRODUCER_B_WARP) {
        // ================================================================== weights (one lane)
        if (lane == 0) {
            const uint32_t a1 = (uint32_t)p.cps * 16384u, b1 = (uint32_t)(p.cps * 64 * PW * 2);
            const uint32_t a2 = (uint32_t)p.cps2 * 16384u, b2 = (uint32_t)(64 * Nc2 * 2), b2s = (uint32_t)p.cps2 * b2;
            for (uint32_t it = 0; it <= it_heads; ++it) {
                const uint32_t s = it % F2_SLOTS, ph = (it / F2_SLOTS) & 1u;
                fs_mbar_wait(bars + 8 * (B2_EMPTY + s), ph ^ 1u);
                const uint32_t full = bars + 8 * (B2_FULL + s), dst = sbase + s * F2_SLOT + F2_B_OFF;
                if (it < (uint32_t)n1) {                     // the A bytes of a stage come from the other producer warp
                    const int h = (int)it / n1p, q = (int)it - h * n1p;
                    fs_expect_tx(full, a1 + b1);
                    fs_bulk_copy(dst, p.wblob + p.w_off[0] + ((size_t)(c * 2 + h) * n1p + q) * b1, b1, full, 0u, false);
                } else if (it < it_heads) {
                    const int kc = f2_kc(((int)it - n1) * p.cps2, c, bpp);
                    fs_expect_tx(full, a2 + b2s);
                    fs_bulk_copy(dst, p.wblob + p.w_off[1] + ((size_t)c * p.nch2 + kc) * b2, b2s, full, 0u, false);
                } else {                                     // this CTA's K slice of the heads' weights
                    const uint32_t hb_bytes = (uint32_t)(Nc2 * n3 * 2);
                    fs_expect_tx(full, hb_bytes);
                    fs_bulk_copy(dst, p.wblob + p.w_off[2] + (size_t)c * hb_bytes, hb_bytes, full, 0u, false);
                }
            }
        }
        __syncwarp();
    } else if (warp == F2_PRODUCER_A_WARP) {
        // ================================================================== activations (one lane)
        if (lane == 0) {
            const uint32_t a1 = (uint32_t)p.cps * 16384u, a2 = (uint32_t)p.cps2 * 16384u;
            const uint8_t* xsrc = p.x + (size_t)m * p.nkx * 16384u;
            const uint8_t* hsrc = p.h1 + (size_t)m * p.nch2 * 16384u;
            for (uint32_t it = 0; it < it_heads; ++it) {
                const uint32_t s = it % F2_SLOTS, ph = (it / F2_SLOTS) & 1u;
                fs_mbar_wait(bars + 8 * (B2_EMPTY + s), ph ^ 1u);
                const uint32_t full = bars + 8 * (B2_FULL + s), dst = sbase + s * F2_SLOT;
                if (it < (uint32_t)n1) {
                    const int q = (int)it % n1p;             // both passes read the whole row tile of X
                    fs_bulk_copy(dst, xsrc + (size_t)q * a1, a1, full, 0u, false);
                } else {
                    const int kc = f2_kc(((int)it - n1) * p.cps2, c, bpp);
                    // The writers made their stores visible at cluster scope before the (release) arrive that completes the
                    // block's barrier, and the copy reads the L2: observing the barrier is enough (see fcstack_kernels.cu).
                    for (int i = 0; i < p.cps2; ++i) fs_mbar_wait(bars + 8 * (B2_READY + kc + i), 0u);
                    fs_bulk_copy(dst, hsrc + (size_t)kc * 16384u, a2, full, 0u, false);
                }
            }
        }
        __syncwarp();
    } else if (warp == F2_MMA_WARP) {
        // ================================================================== MMA issue
        const uint32_t idesc1 = umma_idesc_bf16_m128((uint32_t)PW), idesc2 = umma_idesc_bf16_m128((uint32_t)Nc2);
        for (uint32_t it = 0; it < it_heads; ++it) {
            const uint32_t s = it % F2_SLOTS, ph = (it / F2_SLOTS) & 1u;
            fs_mbar_wait(bars + 8 * (B2_FULL + s), ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const bool first = it < (uint32_t)n1;
            const int h = first ? (int)it / n1p : 2;          // accumulator: fc1 pass 0 / 1, fc2
            const int q = first ? (int)it - h * n1p : (int)it - n1;
            const bool last = first ? q == n1p - 1 : q == n2 - 1;
            if (dbg && lane == 0) {
                if (it == 0) dbg[2] = clock64() - t_entry;
                if (it == (uint32_t)n1) dbg[9] = clock64() - t_entry;
                if (it == (uint32_t)(n1 + n2 / 2)) dbg[10] = clock64() - t_entry;
            }
            if (elect_one()) {
                const uint32_t sa = sbase + s * F2_SLOT, sb = sa + F2_B_OFF;
                const int ks = (first ? p.cps : p.cps2) * 4;
                const uint32_t nc = (uint32_t)(first ? PW : Nc2);
                const uint32_t d = tmem_base + (first ? (uint32_t)(h * PW) : 256u);
                for (int k = 0; k < ks; ++k)
                    umma_bf16(d, umma_desc(sa + (uint32_t)k * 4096u, 2048u, 128u),
                              umma_desc(sb + (uint32_t)k * 32u * nc, 16u * nc, 128u), first ? idesc1 : idesc2, (q || k) ? 1u : 0u);
                umma_commit(bars + 8 * (B2_EMPTY + s));
                if (last) umma_commit(bars + 8 * (B2_ACC + h));
            }
            __syncwarp();
            if (dbg && lane == 0 && last) dbg[h == 2 ? 11 : 3 + h] = clock64() - t_entry;   // [3] pass 0, [4] pass 1, [11] fc2 issued
        }
        fs_mbar_wait(bars + 8 * (B2_FULL + slot_hb), (it_heads / F2_SLOTS) & 1u);
        fs_mbar_wait(bars + 8 * B2_H2, 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (dbg && lane == 0) dbg[14] = clock64() - t_entry;
        if (elect_one()) {
            const uint32_t idesc3 = umma_idesc_bf16_m128((uint32_t)n3);
            const uint32_t sa = sbase + slot_h2 * F2_SLOT, sb = sbase + slot_hb * F2_SLOT + F2_B_OFF;
            for (int k = 0; k < (Nc2 >> 4); ++k)
                umma_bf16(tmem_base + 384u, umma_desc(sa + (uint32_t)k * 4096u, 2048u, 128u),
                          umma_desc(sb + (uint32_t)(k * 32 * n3), (uint32_t)(16 * n3), 128u), idesc3, k ? 1u : 0u);
            umma_commit(bars + 8 * (B2_ACC + 3));
        }
        __syncwarp();
        if (dbg && lane == 0) dbg[15] = clock64() - t_entry;
    } else {
        // ================================================================== epilogue
        const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        {   // biases -> shared memory while the first operands stream in
            const float* b1 = reinterpret_cast<const float*>(p.wblob + p.b_off[0]) + c * 2 * PW;
            const float* b2 = reinterpret_cast<const float*>(p.wblob + p.b_off[1]) + c * Nc2;
            const float* b3 = reinterpret_cast<const float*>(p.wblob + p.b_off[2]);
            for (int i = tid; i < 2 * PW; i += F2_EPI_WARPS * 32) bias_s[i] = b1[i];
            for (int i = tid; i < Nc2; i += F2_EPI_WARPS * 32) bias_s[256 + i] = b2[i];
            for (int i = tid; i < n3; i += F2_EPI_WARPS * 32) bias_s[384 + i] = b3[i];
            asm volatile("bar.sync 1, %0;" ::"r"(F2_EPI_WARPS * 32) : "memory");
        }
        // ---- fc1, pass h: 32 columns per warp -> bias / ReLU / bf16 -> half a 64-wide block of h1 -> ready
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            fs_mbar_wait(bars + 8 * (B2_ACC + h), 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (dbg && tid == 0) dbg[5 + h * 2] = clock64() - t_entry;                          // [5], [7]: accumulator seen
            for (int g = blk; g < (PW >> 5); g += F2_EPI_WARPS / 4) {
                const int j = g >> 1, hb = g & 1;
                const uint32_t col = (uint32_t)(h * PW + g * 32);
                fs_tmem_ld16(t_lane + col, &v[0]);
                fs_tmem_ld16(t_lane + col + 16u, &v[16]);
                fs_tmem_wait32(v);
                const int kc = c * 2 * bpp + h * bpp + j;
                uint8_t* dst = p.h1 + ((size_t)m * p.nch2 + kc) * 16384u + (size_t)row * 16 + (size_t)hb * 4 * 2048;
                const float* bs = bias_s + h * PW + g * 32;
#pragma unroll
                for (int pl = 0; pl < 4; ++pl) {
                    const float4 b0 = *reinterpret_cast<const float4*>(bs + pl * 8);
                    const float4 b1 = *reinterpret_cast<const float4*>(bs + pl * 8 + 4);
                    uint4 o;
                    o.x = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 0]) + b0.x, __uint_as_float(v[pl * 8 + 1]) + b0.y);
                    o.y = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 2]) + b0.z, __uint_as_float(v[pl * 8 + 3]) + b0.w);
                    o.z = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 4]) + b1.x, __uint_as_float(v[pl * 8 + 5]) + b1.y);
                    o.w = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 6]) + b1.z, __uint_as_float(v[pl * 8 + 7]) + b1.w);
                    *reinterpret_cast<uint4*>(dst + pl * 2048) = o;
                }
                // read by bulk copies (async proxy) of all four CTAs: every storing lane's proxy fence, then the release
                // of the remote arrive -- cumulative over the warp's stores through __syncwarp -- publishes it.  (Staging
                // the block in shared memory and writing it with one bulk store was tried: its wait_group takes ~7 000
                // cycles against ~3 500 for this path, profiles/r2_fcstack_experiments.txt.)
                asm volatile("fence.proxy.async.global;" ::: "memory");
                __syncwarp();
                if (lane < F2_NS) fs_mbar_arrive_remote(bars + 8 * (B2_READY + kc), (uint32_t)lane);
            }
            if (dbg && tid == 0) dbg[6 + h * 2] = clock64() - t_entry;                          // [6], [8]: blocks published
        }
        // ---- fc2: 32 columns per warp -> this CTA's 128 x Nc2 slice of h2 as the heads' A operand
        fs_mbar_wait(bars + 8 * (B2_ACC + 2), 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (dbg && tid == 0) dbg[12] = clock64() - t_entry;
        for (int g = blk; g < (Nc2 >> 5); g += F2_EPI_WARPS / 4) {
            fs_tmem_ld16(t_lane + 256u + (uint32_t)(g * 32), &v[0]);
            fs_tmem_ld16(t_lane + 256u + (uint32_t)(g * 32) + 16u, &v[16]);
            fs_tmem_wait32(v);
            uint8_t* dst = smem + (size_t)slot_h2 * F2_SLOT + (size_t)(g * 4) * 2048 + (size_t)row * 16;
            const float* bs = bias_s + 256 + g * 32;
#pragma unroll
            for (int pl = 0; pl < 4; ++pl) {
                const float4 b0 = *reinterpret_cast<const float4*>(bs + pl * 8);
                const float4 b1 = *reinterpret_cast<const float4*>(bs + pl * 8 + 4);
                uint4 o;
                o.x = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 0]) + b0.x, __uint_as_float(v[pl * 8 + 1]) + b0.y);
                o.y = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 2]) + b0.z, __uint_as_float(v[pl * 8 + 3]) + b0.w);
                o.z = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 4]) + b1.x, __uint_as_float(v[pl * 8 + 5]) + b1.y);
                o.w = pack_relu_bf16x2(__uint_as_float(v[pl * 8 + 6]) + b1.z, __uint_as_float(v[pl * 8 + 7]) + b1.w);
                *reinterpret_cast<uint4*>(dst + pl * 2048) = o;
            }
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the tensor core reads shared memory through the async proxy
        __syncwarp();
        if (lane == 0) mbar_arrive(bars + 8 * B2_H2);
        if (dbg && tid == 0) dbg[13] = clock64() - t_entry;
        // ---- heads: this CTA's partial sums (its K slice) of all n3 outputs; 16-column groups blk and blk + 4
        fs_mbar_wait(bars + 8 * (B2_ACC + 3), 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (dbg && tid == 0) dbg[16] = clock64() - t_entry;
        const int ngrp = n3 >> 4;
        if (blk < ngrp) fs_tmem_ld16(t_lane + 384u + (uint32_t)(blk * 16), &v[0]);
        if (blk + 4 < ngrp) fs_tmem_ld16(t_lane + 384u + (uint32_t)((blk + 4) * 16), &v[16]);
        fs_tmem_wait32(v);
        if (quarter != c) {       // rows of quarter q are summed by CTA q
            const uint32_t si = (uint32_t)((c - quarter + F2_NS) & (F2_NS - 1)) - 1u;
            const uint32_t peer = fs_mapa(sbase + F2_XCHG_OFF, (uint32_t)quarter);
#pragma unroll
            for (int gi = 0; gi < 2; ++gi) {
                const int g = blk + 4 * gi;
                if (g < ngrp) {
                    // [sender][group][i][lane][16 B]: every store instruction of the warp writes 512 contiguous bytes
                    const uint32_t a = peer + (si * 5u + (uint32_t)g) * 2048u + (uint32_t)lane * 16u;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        fs_st_cluster_v4(a + 512u * i, v[gi * 16 + 4 * i], v[gi * 16 + 4 * i + 1], v[gi * 16 + 4 * i + 2],
                                         v[gi * 16 + 4 * i + 3]);
                }
            }
        }
        if (dbg && tid == 0) dbg[17] = clock64() - t_entry;
    }

    // ---- the partial sums have been sent: one cluster barrier publishes them; it is also the teardown barrier (no CTA
    // leaves while a peer may still signal its barriers or write into its shared memory)
    __syncwarp();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    fs_cluster_sync();
    if (dbg && tid == 0) dbg[18] = clock64() - t_entry;

    if (warp < F2_EPI_WARPS && quarter == c) {
        // ---- owner of rows 32 c .. 32 c + 31: add the three peers' partial sums and the bias, bf16, then through shared
        // memory so that the global stores are row-contiguous 4-byte words
        const int ngrp = n3 >> 4;
        uint32_t* stage = reinterpret_cast<uint32_t*>(smem + (size_t)slot_out * F2_SLOT);
        const int pitch = (F2_MAX_N3 >> 1) + 1;               // words per staged row, odd: no bank conflicts
#pragma unroll
        for (int gi = 0; gi < 2; ++gi) {
            const int g = blk + 4 * gi;
            if (g < ngrp) {
                float acc[16];
#pragma unroll
                for (int e = 0; e < 16; ++e) acc[e] = __uint_as_float(v[gi * 16 + e]) + bias_s[384 + g * 16 + e];
#pragma unroll
                for (int si = 0; si < 3; ++si) {
                    const float4* src = reinterpret_cast<const float4*>(smem + F2_XCHG_OFF + (size_t)(si * 5 + g) * 2048 + lane * 16);
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const float4 t = src[i * 32];
                        acc[4 * i] += t.x; acc[4 * i + 1] += t.y; acc[4 * i + 2] += t.z; acc[4 * i + 3] += t.w;
                    }
                }
#pragma unroll
                for (int i = 0; i < 8; ++i) stage[lane * pitch + g * 8 + i] = pack_bf16x2(acc[2 * i], acc[2 * i + 1]);
            }
        }
        asm volatile("bar.sync 2, 128;" ::: "memory");
        const int pw = (p.n_out + 1) >> 1;
        const long long row0 = (long long)m * 128 + c * 32;
        const int t128 = blk * 32 + lane;
        for (int idx = t128; idx < 32 * pw; idx += 128) {
            const int rr = idx / pw, k = idx - rr * pw, hcol = 2 * k;
            if (row0 + rr < p.G) {
                __nv_bfloat16* out = p.logits + (row0 + rr) * p.ld + hcol;
                const uint32_t w = stage[rr * pitch + k];
                if (hcol + 1 < p.n_out) *reinterpret_cast<uint32_t*>(out) = w;
                else *out = __ushort_as_bfloat16((unsigned short)(w & 0xffffu));
            }
        }
    }
    if (dbg && tid == 0) {
        dbg[24] = clock64() - t_entry;
        unsigned long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt)); dbg[19] = (long long)gt;
    }
    if (warp == F2_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

int az_fcstack2_launch(const FcPlan& pl, const void* x, const void* wblob, void* h1, void* logits, int ld, int n_out,
                       int64_t G, long long* debug, int late_trigger, bool pdl, void* stream) {
    static bool attr_set = false;
    if (!attr_set) {
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_fcstack2, cudaFuncAttributeMaxDynamicSharedMemorySize, F2_SM_TOTAL));
        attr_set = true;
    }
    Fc2Params p;
    p.x = (const uint8_t*)x; p.wblob = (const uint8_t*)wblob; p.h1 = (uint8_t*)h1; p.logits = (__nv_bfloat16*)logits;
    p.G = G; p.ld = ld; p.n_out = n_out;
    p.nkx = pl.K[0] / 64; p.cps = (p.nkx & 1) ? 1 : 2;
    p.PW = pl.Nc[0]; p.Nc2 = pl.Nc[1]; p.n3 = pl.Nc[2]; p.nch2 = pl.K[1] / 64;
    p.cps2 = p.PW == 128 ? 2 : 1;
    for (int L = 0; L < 3; ++L) { p.w_off[L] = pl.w_off[L]; p.b_off[L] = pl.b_off[L]; }
    p.late_trigger = late_trigger; p.debug = debug;
    const int64_t tiles = (G + 127) / 128;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(tiles * F2_NS)); cfg.blockDim = dim3(F2_THREADS);
    cfg.dynamicSmemBytes = F2_SM_TOTAL; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = F2_NS; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_net_fcstack2, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_fcstack2");
    AZ_CHECK_LAUNCH("k_net_fcstack2");
    return AZ_OK;
}
