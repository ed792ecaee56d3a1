// The fully connected stack of the policy-value networks, third kernel: k_net_fcstack2's schedule on CTA PAIRS
// (tcgen05.mma.cta_group::2, M = 256).  See fcstack_kernels.cu for the layers and the operand layouts
// (alphazero/games/othello.py:352-362, :374-382) and fcstack2_kernels.cu for the two-pass fc1 / K-split heads schedule.
//
// Why.  The second kernel's main loops run at the rate its operands arrive: every SM receives ~1 MB per launch (its 128
// rows of X twice, of h1 once, and its column slice of the weights) at 52 - 64 B/clk, which is what an SM can take
// (profiles/r2_fcstack_experiments.txt).  With M = 256 MMAs two CTAs that hold DIFFERENT row tiles and the SAME column
// slice share the weights: each stages only half of the slice's columns and the tensor cores of both SMs read both
// halves.  Weight bytes per SM halve (1 044 -> 778 KB in all), and one thread issues the MMAs of two SMs.
//
// Work split.  A group of 8 CTAs owns 2 row tiles (256 positions); cluster = pair p of the group = column slice p,
// rank = row tile within the group.  (A cluster of all 8 would give the exchange below cluster barriers and
// distributed shared memory, but only 15 clusters of 8 are resident on this part at once, and 4096 positions need 16.)
// So the four pairs of a group talk through global memory (L2) and flags:
//   * h1 blocks: as before st.global in the packed operand layout, then one red.release.gpu per storing warp on the
//     block's counter; the activation producers of the 4 CTAs of that row tile poll it (ld.acquire.gpu) before their
//     bulk copy of the block;
//   * the heads' fp32 partial sums (the heads are split over K): every CTA writes the partials of the rows it does not
//     own to the scratch, coalesced, and bumps the owner's counter; the owner (row quarter q of a tile -> the CTA of
//     column slice q) polls, adds the three partials to its registers, adds the bias and writes the bf16 logits.
// Counters live at the end of each row tile's region of the `h2` scratch (h2 itself never leaves the SM), must be zero
// before the first launch and are zeroed again by the last CTA of a group to finish.
//
// Pair protocol.  Both CTAs run the same producers (own rows of A, own half of B) into their own ring (4 slots of
// 48 KB: A 32 KB + B 16 KB, two 64-wide K chunks per stage).  Rank 1's control warp relays "my stage has landed" to a
// second barrier set in rank 0; rank 0's control warp issues the MMAs once both halves are in place and commits to the
// empty / accumulator barriers of BOTH CTAs (tcgen05.commit ... multicast::cluster).  Semantics of the pair MMA (CTA r
// supplies rows 128 r .. of A and columns r N/2 .. of B; its TMEM receives rows 128 r .. x all N columns) were pinned with
// tools/mma2_probe.py (profiles/r2_mma2_probe.txt).
//
// TMEM columns (each CTA, its own 128 rows): fc1 pass 0 [0, PW), pass 1 [PW, 2 PW), fc2 [256, 256 + Nc2), heads [384, ..).
// Warps: 0-15 epilogue, 16 weight producer, 17 activation producer, 18 control (rank 0: MMA issue, rank 1: relay).
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "fcstack_common.cuh"

#define F3_NS 4
#define F3_SLOT 49152u                         // A block(s) at +0 (32 KB), this CTA's half of the B block(s) at +32 KB (16 KB)
#define F3_B_OFF 32768u
#define F3_SLOTS 4
#define F3_EPI_WARPS 16
#define F3_PRODUCER_B_WARP 16
#define F3_PRODUCER_A_WARP 17
#define F3_CTRL_WARP 18
#define F3_THREADS (19 * 32)
#define F3_MAX_N3 80
#define B3_FULL 0                              // [4]  this CTA's stage has landed (A + B bytes)
#define B3_EMPTY 4                             // [4]  stage consumed (tcgen05.commit from rank 0, both CTAs)
#define B3_PFULL 8                             // [4]  rank 0 only: the peer's stage has landed (relayed by rank 1)
#define B3_ACC 12                              // [4]  fc1 pass 0, pass 1, fc2, heads accumulator complete
#define B3_H2 16                               // h2 is in shared memory (16 epilogue warps)
#define B3_COUNT 17
#define F3_SM_BIAS (F3_SLOTS * F3_SLOT)
#define F3_SM_BARS (F3_SM_BIAS + 512 * 4)
#define F3_SM_TMEM_PTR (F3_SM_BARS + 8 * 24)
#define F3_SM_TOTAL (F3_SM_TMEM_PTR + 16)
static_assert(F3_SM_TOTAL <= 232448, "shared memory per CTA on sm_100");
// scratch of a row tile (its region of `h2`): partial sums [owner quarter 4][sender 3][16-column group 5][2048 B], then
// 64 counters in the last 256 bytes: [0, 16) h1 block kc stored (8 warps), [16, 20) partials for owner quarter q stored
// (3 senders x 4 warps), [20] CTAs of the group that have finished (first tile of the group only)
#define F3_FLAG_BYTES 256

struct Fc3Params {
    const uint8_t* x;          // layer-1 input, packed A: [tiles][K1/64][8][128][8] bf16, tiles even
    const uint8_t* wblob;      // packed weights (az_net_fcstack_plan, variant 3), then the float32 biases
    uint8_t* h1;               // scratch, packed A: [tiles][N1/64][8][128][8]
    uint8_t* xch;              // scratch (`h2` of the C ABI): per row tile partial sums + counters, see above
    __nv_bfloat16* logits;     // [G][ld]
    long long G;
    long long tile_bytes;      // bytes of `xch` per row tile (128 * N2 * 2)
    int ld, n_out;
    int nkx, cps;              // K1 / 64; 64-wide K chunks per fc1 stage (2 when nkx is even)
    int PW, Nc2, n3;           // fc1 columns per pass, fc2 columns per column slice, padded head outputs
    int nch2, cps2;            // N1 / 64 = K chunks of fc2; chunks per fc2 stage
    uint32_t w_off[3], b_off[3];
    int late_trigger;
    long long* debug;          // optional [n_ctas][32] cycle stamps
};

__device__ __forceinline__ int f3_kc(int q, int c, int bpp) {      // = f2_kc of fcstack2_kernels.cu
    const int per = F3_NS * bpp, h = q / per, rem = q - h * per;
    const int cs = ((rem / bpp) + c) & (F3_NS - 1), j = rem % bpp;
    return cs * 2 * bpp + h * bpp + j;
}

__device__ __forceinline__ void f3_flag_add(uint32_t* f) {
    asm volatile("red.release.gpu.global.add.u32 [%0], 1;" ::"l"(f) : "memory");
}
__device__ __forceinline__ void f3_flag_wait(const uint32_t* f, uint32_t want) {
    long long t0 = 0;
    for (;;) {
        uint32_t v;
        asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if (v >= want) return;
        const long long t = clock64();
        if (t0 == 0) t0 = t;
        else if (t - t0 > FS_WAIT_LIMIT) __trap();     // a sibling CTA that never arrives: fail the launch, do not hang
    }
}

__device__ __forceinline__ void f3_umma2(uint32_t tmem_d, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void f3_commit2(uint32_t bar) {      // arrives on `bar` in both CTAs of the pair
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ uint32_t f3_idesc(uint32_t n) {      // kind::f16, D fp32, A / B bf16 K-major, M = 256
    return (1u << 4) | (1u << 7) | (1u << 10) | ((n >> 3) << 17) | ((256u >> 4) << 24);
}

__global__ void __launch_bounds__(F3_THREADS, 1) k_net_fcstack3(Fc3Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    long long* dbg = p.debug ? p.debug + (size_t)blockIdx.x * 32 : nullptr;
    const long long t_entry = dbg ? clock64() : 0;
    if (dbg && tid == 0) { unsigned long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt)); dbg[25] = (long long)gt; }
    uint32_t rank, cluster_id;
    asm("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
    asm("mov.u32 %0, %%clusterid.x;" : "=r"(cluster_id));
    const int c = (int)cluster_id & (F3_NS - 1), grp = (int)cluster_id >> 2;      // column slice, group of two row tiles
    const int m = grp * 2 + (int)rank;                                            // row tile of this CTA
    const bool leader = rank == 0;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = smem_u32(smem + F3_SM_BARS);
    float* bias_s = reinterpret_cast<float*>(smem + F3_SM_BIAS);          // fc1 [0, 256), fc2 [256, 384), heads [384, 512)
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + F3_SM_TMEM_PTR);
    const int PW = p.PW, Nc2 = p.Nc2, n3 = p.n3, bpp = PW >> 6, ngrp = n3 >> 4;
    const int n1p = p.nkx / p.cps, n1 = 2 * n1p;              // fc1 stages per pass / in total
    const int n2 = p.nch2 / p.cps2;                           // fc2 stages
    const uint32_t it_heads = (uint32_t)(n1 + n2);            // the ring's last stage carries the heads' weights
    const uint32_t slot_hb = it_heads % F3_SLOTS, slot_out = (it_heads + 1u) % F3_SLOTS;
    uint8_t* xch_m = p.xch + (size_t)m * p.tile_bytes;        // partial sums for the rows of this tile
    uint32_t* flags_m = reinterpret_cast<uint32_t*>(xch_m + p.tile_bytes - F3_FLAG_BYTES);
    uint32_t* flags_g = reinterpret_cast<uint32_t*>(p.xch + (size_t)(grp * 2 + 1) * p.tile_bytes - F3_FLAG_BYTES);

    // ---- setup
    if (tid < B3_COUNT) {
        mbar_init(bars + 8 * tid, tid == B3_H2 ? F3_EPI_WARPS : 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == F3_CTRL_WARP) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    fs_cluster_sync();            // the peer's barriers exist before a commit / relay targets them
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr_smem;
    az_pdl_wait();                // X is the preceding kernel's output; the counters were zeroed by the launch before
    if (!p.late_trigger) az_pdl_trigger();
    if (dbg && tid == 0) dbg[1] = clock64() - t_entry;

    uint32_t v[32];               // epilogue registers; the owner's partial sums stay in them while the peers' arrive
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = 0u;
    const int quarter = warp & 3, blk = warp >> 2;
    const int row = quarter * 32 + lane;                      // TMEM lane = row of the tile

    if (warp == F3_PRODUCER_B_WARP) {
        // ================================================================== this CTA's half of the weights (one lane)
        if (lane == 0) {
            const uint32_t a1 = (uint32_t)p.cps * 16384u, b1 = (uint32_t)(p.cps * 64 * (PW >> 1) * 2);
            const uint32_t a2 = (uint32_t)p.cps2 * 16384u, b2 = (uint32_t)(64 * (Nc2 >> 1) * 2), b2s = (uint32_t)p.cps2 * b2;
            for (uint32_t it = 0; it <= it_heads; ++it) {
                const uint32_t s = it % F3_SLOTS, ph = (it / F3_SLOTS) & 1u;
                fs_mbar_wait(bars + 8 * (B3_EMPTY + s), ph ^ 1u);
                const uint32_t full = bars + 8 * (B3_FULL + s), dst = sbase + s * F3_SLOT + F3_B_OFF;
                if (it < (uint32_t)n1) {                     // the A bytes of a stage come from the other producer warp
                    const int h = (int)it / n1p, q = (int)it - h * n1p;
                    fs_expect_tx(full, a1 + b1);
                    fs_bulk_copy(dst, p.wblob + p.w_off[0] + ((size_t)((c * 2 + h) * 2 + (int)rank) * n1p + q) * b1, b1, full, 0u, false);
                } else if (it < it_heads) {
                    const int kc = f3_kc(((int)it - n1) * p.cps2, c, bpp);
                    fs_expect_tx(full, a2 + b2s);
                    fs_bulk_copy(dst, p.wblob + p.w_off[1] + ((size_t)(c * 2 + (int)rank) * p.nch2 + kc) * b2, b2s, full, 0u, false);
                } else {                                     // this CTA's column half of the heads' K slice c
                    const uint32_t hb = (uint32_t)(Nc2 * (n3 >> 1) * 2);
                    fs_expect_tx(full, hb);
                    fs_bulk_copy(dst, p.wblob + p.w_off[2] + (size_t)(c * 2 + (int)rank) * hb, hb, full, 0u, false);
                }
            }
        }
        __syncwarp();
    } else if (warp == F3_PRODUCER_A_WARP) {
        // ================================================================== this CTA's rows of the activations (one lane)
        if (lane == 0) {
            const uint32_t a1 = (uint32_t)p.cps * 16384u, a2 = (uint32_t)p.cps2 * 16384u;
            const uint8_t* xsrc = p.x + (size_t)m * p.nkx * 16384u;
            const uint8_t* hsrc = p.h1 + (size_t)m * p.nch2 * 16384u;
            for (uint32_t it = 0; it < it_heads; ++it) {
                const uint32_t s = it % F3_SLOTS, ph = (it / F3_SLOTS) & 1u;
                fs_mbar_wait(bars + 8 * (B3_EMPTY + s), ph ^ 1u);
                const uint32_t full = bars + 8 * (B3_FULL + s), dst = sbase + s * F3_SLOT;
                if (it < (uint32_t)n1) {
                    const int q = (int)it % n1p;             // both passes read the whole row tile of X
                    fs_bulk_copy(dst, xsrc + (size_t)q * a1, a1, full, 0u, false);
                } else {
                    const int kc = f3_kc(((int)it - n1) * p.cps2, c, bpp);
                    // the 8 warps that store a block release at gpu scope after their stores (and proxy fences): once the
                    // counter reads 8 the block is in the L2, which is what the bulk copy reads
                    for (int i = 0; i < p.cps2; ++i) f3_flag_wait(flags_m + kc + i, 8u);
                    fs_bulk_copy(dst, hsrc + (size_t)kc * 16384u, a2, full, 0u, false);
                }
            }
        }
        __syncwarp();
    } else if (warp == F3_CTRL_WARP && !leader) {
        // ================================================================== rank 1: relay "my stage has landed" to rank 0
        for (uint32_t it = 0; it <= it_heads; ++it) {
            const uint32_t s = it % F3_SLOTS, ph = (it / F3_SLOTS) & 1u;
            fs_mbar_wait(bars + 8 * (B3_FULL + s), ph);
            if (it == it_heads) fs_mbar_wait(bars + 8 * B3_H2, 0u);       // the heads' stage: weights landed and h2 written
            if (lane == 0) fs_mbar_arrive_remote(bars + 8 * (B3_PFULL + s), 0u);
            __syncwarp();
        }
    } else if (warp == F3_CTRL_WARP) {
        // ================================================================== rank 0: MMA issue for both CTAs
        const uint32_t idesc1 = f3_idesc((uint32_t)PW), idesc2 = f3_idesc((uint32_t)Nc2);
        long long w_full = 0, w_pfull = 0, t_issue = 0;
        for (uint32_t it = 0; it < it_heads; ++it) {
            const uint32_t s = it % F3_SLOTS, ph = (it / F3_SLOTS) & 1u;
            const long long c0 = dbg ? clock64() : 0;
            fs_mbar_wait(bars + 8 * (B3_FULL + s), ph);
            const long long c1 = dbg ? clock64() : 0;
            // (a cluster-scope acquire here costs ~500 cycles per stage even when the barrier is long complete; what the
            // barrier guards -- the peer's stage -- was written and is read through the async proxy, nothing generic)
            fs_mbar_wait(bars + 8 * (B3_PFULL + s), ph);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const long long c2 = dbg ? clock64() : 0;
            w_full += c1 - c0; w_pfull += c2 - c1;
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
                const uint32_t sa = sbase + s * F3_SLOT, sb = sa + F3_B_OFF;
                const int ks = (first ? p.cps : p.cps2) * 4;
                const uint32_t nh = (uint32_t)(first ? PW : Nc2) >> 1;        // B columns staged per CTA
                const uint32_t d = tmem_base + (first ? (uint32_t)(h * PW) : 256u);
                // descriptors of K step k = those of step 0 + k * (bytes per step >> 4) in the start-address field
                uint64_t ad = umma_desc(sa, 2048u, 128u), bd = umma_desc(sb, 16u * nh, 128u);
                const uint64_t bstep = (uint64_t)(2u * nh);
                for (int k = 0; k < ks; ++k, ad += 256u, bd += bstep)
                    f3_umma2(d, ad, bd, first ? idesc1 : idesc2, (q || k) ? 1u : 0u);
                f3_commit2(bars + 8 * (B3_EMPTY + s));
                if (last) f3_commit2(bars + 8 * (B3_ACC + h));
            }
            __syncwarp();
            if (dbg) t_issue += clock64() - c2;
            if (dbg && lane == 0 && last) dbg[h == 2 ? 11 : 3 + h] = clock64() - t_entry;   // [3] pass 0, [4] pass 1, [11] fc2 issued
        }
        if (dbg && lane == 0) { dbg[20] = w_full; dbg[21] = w_pfull; dbg[22] = t_issue; }
        fs_mbar_wait(bars + 8 * (B3_FULL + slot_hb), (it_heads / F3_SLOTS) & 1u);
        fs_mbar_wait(bars + 8 * B3_H2, 0u);
        fs_mbar_wait(bars + 8 * (B3_PFULL + slot_hb), (it_heads / F3_SLOTS) & 1u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (dbg && lane == 0) dbg[14] = clock64() - t_entry;
        if (elect_one()) {
            const uint32_t nh = (uint32_t)n3 >> 1;
            uint64_t ad = umma_desc(sbase + slot_hb * F3_SLOT, 2048u, 128u);
            uint64_t bd = umma_desc(sbase + slot_hb * F3_SLOT + F3_B_OFF, 16u * nh, 128u);
            const uint32_t idesc3 = f3_idesc((uint32_t)n3);
            for (int k = 0; k < (Nc2 >> 4); ++k, ad += 256u, bd += (uint64_t)(2u * nh))
                f3_umma2(tmem_base + 384u, ad, bd, idesc3, k ? 1u : 0u);
            f3_commit2(bars + 8 * (B3_ACC + 3));
        }
        __syncwarp();
        if (dbg && lane == 0) dbg[15] = clock64() - t_entry;
    } else {
        // ================================================================== epilogue (own 128 rows, in this CTA's TMEM)
        const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16);
        {   // biases -> shared memory while the first operands stream in
            const float* b1 = reinterpret_cast<const float*>(p.wblob + p.b_off[0]) + c * 2 * PW;
            const float* b2 = reinterpret_cast<const float*>(p.wblob + p.b_off[1]) + c * Nc2;
            const float* b3 = reinterpret_cast<const float*>(p.wblob + p.b_off[2]);
            for (int i = tid; i < 2 * PW; i += F3_EPI_WARPS * 32) bias_s[i] = b1[i];
            for (int i = tid; i < Nc2; i += F3_EPI_WARPS * 32) bias_s[256 + i] = b2[i];
            for (int i = tid; i < n3; i += F3_EPI_WARPS * 32) bias_s[384 + i] = b3[i];
            asm volatile("bar.sync 1, %0;" ::"r"(F3_EPI_WARPS * 32) : "memory");
        }
        // ---- fc1, pass h: 32 columns per warp -> bias / ReLU / bf16 -> half a 64-wide block of h1 -> counter
#pragma unroll 1
        for (int h = 0; h < 2; ++h) {
            fs_mbar_wait(bars + 8 * (B3_ACC + h), 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            if (dbg && tid == 0) dbg[5 + h * 2] = clock64() - t_entry;                          // [5], [7]: accumulator seen
            for (int g = blk; g < (PW >> 5); g += F3_EPI_WARPS / 4) {
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
                // read by bulk copies (async proxy) of the four CTAs of this row tile, which sit in four clusters: every
                // storing lane's proxy fence, then one gpu-scope release on the block's counter (cumulative over the
                // warp's stores through __syncwarp)
                asm volatile("fence.proxy.async.global;" ::: "memory");
                __syncwarp();
                if (lane == 0) f3_flag_add(flags_m + kc);
            }
            if (dbg && tid == 0) dbg[6 + h * 2] = clock64() - t_entry;                          // [6], [8]: blocks published
        }
        // ---- fc2: 32 columns per warp -> this CTA's 128 x Nc2 slice of h2 as the heads' A operand (slot_hb, A area)
        fs_mbar_wait(bars + 8 * (B3_ACC + 2), 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (dbg && tid == 0) dbg[12] = clock64() - t_entry;
        for (int g = blk; g < (Nc2 >> 5); g += F3_EPI_WARPS / 4) {
            fs_tmem_ld16(t_lane + 256u + (uint32_t)(g * 32), &v[0]);
            fs_tmem_ld16(t_lane + 256u + (uint32_t)(g * 32) + 16u, &v[16]);
            fs_tmem_wait32(v);
            uint8_t* dst = smem + (size_t)slot_hb * F3_SLOT + (size_t)(g * 4) * 2048 + (size_t)row * 16;
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
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // the tensor cores read shared memory through the async proxy
        __syncwarp();
        if (lane == 0) mbar_arrive(bars + 8 * B3_H2);
        if (dbg && tid == 0) dbg[13] = clock64() - t_entry;
        // ---- heads: this CTA's partial sums (its K slice) of all n3 outputs; 16-column groups blk and blk + 4
        fs_mbar_wait(bars + 8 * (B3_ACC + 3), 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (dbg && tid == 0) dbg[16] = clock64() - t_entry;
        if (blk < ngrp) fs_tmem_ld16(t_lane + 384u + (uint32_t)(blk * 16), &v[0]);
        if (blk + 4 < ngrp) fs_tmem_ld16(t_lane + 384u + (uint32_t)((blk + 4) * 16), &v[16]);
        fs_tmem_wait32(v);
        if (quarter != c) {       // rows of quarter q are summed by the CTA of column slice q (same row tile, another cluster)
            const uint32_t si = (uint32_t)((c - quarter + F3_NS) & (F3_NS - 1)) - 1u;
#pragma unroll
            for (int gi = 0; gi < 2; ++gi) {
                const int g = blk + 4 * gi;
                if (g < ngrp) {
                    // [owner][sender][group][i][lane][16 B]: every store instruction of the warp writes 512 contiguous bytes
                    uint8_t* a = xch_m + ((size_t)((quarter * 3 + (int)si) * ngrp + g)) * 2048 + lane * 16;
#pragma unroll
                    for (int i = 0; i < 4; ++i)
                        *reinterpret_cast<uint4*>(a + 512 * i) = make_uint4(v[gi * 16 + 4 * i], v[gi * 16 + 4 * i + 1],
                                                                             v[gi * 16 + 4 * i + 2], v[gi * 16 + 4 * i + 3]);
                }
            }
            __syncwarp();
            if (lane == 0) f3_flag_add(flags_m + 16 + quarter);
            if (dbg && tid == 0) dbg[17] = clock64() - t_entry;
        } else {
            // ---- owner of rows 32 c .. 32 c + 31 of this tile: the three peers' partial sums + the bias, bf16, then
            // through shared memory so that the global stores are row-contiguous 4-byte words
            if (dbg && lane == 0 && blk == 0) dbg[17] = clock64() - t_entry;
            f3_flag_wait(flags_m + 16 + c, 12u);              // 3 senders x 4 warps
            if (dbg && lane == 0 && blk == 0) dbg[18] = clock64() - t_entry;
            uint32_t* stage = reinterpret_cast<uint32_t*>(smem + (size_t)slot_out * F3_SLOT);
            const int pitch = (F3_MAX_N3 >> 1) + 1;           // words per staged row, odd: no bank conflicts
#pragma unroll
            for (int gi = 0; gi < 2; ++gi) {
                const int g = blk + 4 * gi;
                if (g < ngrp) {
                    float acc[16];
#pragma unroll
                    for (int e = 0; e < 16; ++e) acc[e] = __uint_as_float(v[gi * 16 + e]) + bias_s[384 + g * 16 + e];
                    float4 t[12];                             // all twelve loads in flight before the first add
#pragma unroll
                    for (int si = 0; si < 3; ++si)
#pragma unroll
                        for (int i = 0; i < 4; ++i)
                            t[si * 4 + i] = __ldcg(reinterpret_cast<const float4*>(
                                xch_m + ((size_t)((c * 3 + si) * ngrp + g)) * 2048 + lane * 16 + 512 * i));
#pragma unroll
                    for (int si = 0; si < 3; ++si)
#pragma unroll
                        for (int i = 0; i < 4; ++i) {
                            acc[4 * i] += t[si * 4 + i].x; acc[4 * i + 1] += t[si * 4 + i].y;
                            acc[4 * i + 2] += t[si * 4 + i].z; acc[4 * i + 3] += t[si * 4 + i].w;
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
            if (dbg && lane == 0 && blk == 0) dbg[26] = clock64() - t_entry;
        }
    }

    // ---- teardown.  The last CTA of the group to get here zeroes the group's counters for the next launch (every CTA
    // has finished polling by then); the pair leaves together (TMEM of a cta_group::2 allocation, relay / commit targets).
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
        uint32_t old;
        // relaxed: every poll of this CTA has RETURNED by now (its value was branched on, bar.sync above), and every
        // counter increment of the group has been observed by some poll, so nothing can still read or add to a counter
        // when the eighth arrival zeroes them -- and a gpu-scope release here would wait for the logits stores
        asm volatile("atom.relaxed.gpu.global.add.u32 %0, [%1], 1;" : "=r"(old) : "l"(flags_g + 20) : "memory");
        if (old == 2u * F3_NS - 1u) {
            uint32_t* f1 = reinterpret_cast<uint32_t*>(p.xch + (size_t)(grp * 2 + 2) * p.tile_bytes - F3_FLAG_BYTES);
            for (int i = 0; i < 21; ++i) { flags_g[i] = 0u; f1[i] = 0u; }
        }
    }
    if (dbg && tid == 0) dbg[23] = clock64() - t_entry;
    asm volatile("barrier.cluster.arrive.relaxed.aligned;\n\tbarrier.cluster.wait.aligned;" ::: "memory");
    if (dbg && tid == 0) {
        dbg[24] = clock64() - t_entry;
        unsigned long long gt; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(gt)); dbg[19] = (long long)gt;
    }
    if (warp == F3_CTRL_WARP)
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
}

int az_fcstack3_launch(const FcPlan& pl, const void* x, const void* wblob, void* h1, void* xch, void* logits, int ld,
                       int n_out, int64_t G, long long* debug, int late_trigger, bool pdl, void* stream) {
    static bool attr_set = false;
    if (!attr_set) {
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_fcstack3, cudaFuncAttributeMaxDynamicSharedMemorySize, F3_SM_TOTAL));
        attr_set = true;
    }
    Fc3Params p;
    p.x = (const uint8_t*)x; p.wblob = (const uint8_t*)wblob; p.h1 = (uint8_t*)h1; p.xch = (uint8_t*)xch;
    p.logits = (__nv_bfloat16*)logits; p.G = G; p.ld = ld; p.n_out = n_out;
    p.tile_bytes = 128LL * pl.K[2] * 2;
    p.nkx = pl.K[0] / 64; p.cps = (p.nkx & 1) ? 1 : 2;
    p.PW = pl.Nc[0]; p.Nc2 = pl.Nc[1]; p.n3 = pl.Nc[2]; p.nch2 = pl.K[1] / 64;
    p.cps2 = p.PW == 128 ? 2 : 1;
    for (int L = 0; L < 3; ++L) { p.w_off[L] = pl.w_off[L]; p.b_off[L] = pl.b_off[L]; }
    p.late_trigger = late_trigger; p.debug = debug;
    const int64_t groups = ((G + 127) / 128 + 1) / 2;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(groups * 2 * F3_NS)); cfg.blockDim = dim3(F3_THREADS);
    cfg.dynamicSmemBytes = F3_SM_TOTAL; cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_net_fcstack3, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_fcstack3");
    AZ_CHECK_LAUNCH("k_net_fcstack3");
    return AZ_OK;
}
