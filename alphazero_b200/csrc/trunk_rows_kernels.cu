// Convolution trunk of OthelloNet(n = 8) for batches of 1024 positions and more: tile = one board ROW, one TMEM
// accumulator per OUTPUT tile, no cross-lane traffic in the epilogue.  (alphazero/games/othello.py:366-374: conv1..conv4 +
// BatchNorm + ReLU, pads 1, 1, 0, 0; same weight blob, same outputs as k_net_trunk<8, 8> in net_kernels.cu, which keeps
// the smaller batches and the other board sizes.)
//
// A tile (128 TMEM lanes = 128 rows of the A operand) is ONE BOARD ROW of 14 boards, lane = board * 9 + column (8 cells +
// 1 zero pad column).  The kernel COLUMN is a +-16-byte shift of the A operand (a K-major no-swizzle operand may start
// at any 16-byte row: tools/mma2_probe.py); the kernel ROW is the N dimension, N = 96 = (kernel row kh, c_out), so block
// kh of what the MMAs on INPUT tile r' produce belongs to OUTPUT tile r' - kh + 1 at the same lane: nothing ever has to
// move between lanes (the 9-tap formulation of k_net_trunk pays 64 shuffles per tile for that).  An earlier kernel
// with this layout kept one accumulator per input tile and let an epilogue item add a block from each of three: a stage
// was live for three items and everything serialised on a ring of five (profiles/r2_trunk_experiments.txt, item 5).
// Here the ACCUMULATORS ARE PER OUTPUT TILE and the tensor core does the scatter: the outputs of a layer sit side by
// side in TMEM in DESCENDING tile order, 32 columns each,
//     O[r] at column 32 * (7 - r) of the layer's region,
// so the three blocks of input tile r' -- for O[r'+1], O[r'], O[r'-1] -- are 96 CONSECUTIVE columns starting at
// 32 * (6 - r'), and ONE N = 96 MMA with that D address accumulates all three where they belong (fewer blocks where
// the layer does not produce the tile: conv2's padding rows, the rows conv3 / conv4 lose at the rim).  Successive tile
// jobs overlap in two of their three blocks; MMAs of one thread execute in order, which is all the accumulation needs.
// The first contribution to an output tile must not accumulate: the first MMA of a job is split into an N = 32 MMA
// that initialises the block of the tile the job reaches first and one that accumulates into the others (7 MMAs per
// job instead of 6).  Two regions of 8 x 32 columns alternate between the layers = all 512 TMEM columns; a region is
// reused two layers later, when the jobs of the layer in between have already waited for every tile the old one
// produced (conv2 of the next group, which follows conv4 without such a layer, waits for conv4's four items).
// The epilogue item of an output tile is then what conv1's always was: one 32-column TMEM read, bias, ReLU, bf16, four
// 16-byte stores IN PLACE (the activation buffer is converted layer by layer), one barrier arrival -- a third of the
// TMEM traffic of the other formulations, no adds, no shuffles, no accumulator hand-back.
//
// One thread cannot issue a tcgen05.mma more often than every ~65 cycles (profiles/r1_trunk_experiments.txt), and this
// kernel is a stream of ~200 small MMAs per group: TWO issuing warps, warp h owning the accumulators of the output tiles
// 4 h .. 4 h + 3 (disjoint: no ordering between the warps is needed, the results do not depend on timing).  The input
// tiles next to the border (3, 4) are jobs of both warps, each with the kernel rows that land on its own side.  Both
// go from the border outwards (4, 3, 2, 1, 0 and 3, 4, 5, 6, 7), and the items of a layer in the same order, so that
// what the next layer needs first is complete first.  Three epilogue teams take the items round robin; two producer
// warps build conv1's im2col operand of the next group.  Per CTA: 512 threads, 221 KB of shared memory, persistent.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "trunk_rows.cuh"

#define TR_NB 14                  // boards per group
#define TR_PITCH 9                // lanes per board in a tile
#define TR_TILES 8                // tiles per layer = board rows
#define TR_ROWS (TR_TILES * 128)
#define TR_MARGIN 8               // zero rows before / after a buffer (the +-1 row shifts of the first / last tile)
#define TR_PLANE ((TR_ROWS + 2 * TR_MARGIN) * 16)      // one 8-channel chunk plane: 16 640 B (= LBO of the A operand)
#define TR_BUF (4 * TR_PLANE)     // activation buffer: 66 560 B
#define TR_IM2COL (2 * TR_PLANE)  // conv1's A operand (K = 16: 9 taps + zeros): 33 280 B
#define TR_TEAMS 3
#define TR_EPI_THREADS (128 * TR_TEAMS)
#define TR_MMA_WARP (TR_EPI_THREADS / 32)
#define TR_MMA_WARPS 2            // issuing warp h accumulates the output tiles 4 h .. 4 h + 3 of every layer
#define TR_PRODUCERS 64
#define TR_THREADS (TR_EPI_THREADS + 32 * TR_MMA_WARPS + TR_PRODUCERS)
#define TR_HJOBS 18               // MMA tile jobs per group and issuing warp: 4 + 5 + 5 + 4
#define TR_ITEMS 26               // epilogue items per group = output tiles that are used: 8 + 8 + 6 + 4
// shared memory map
#define TR_SM_BUF(i) ((i) * TR_BUF)
#define TR_SM_IM2COL (2 * TR_BUF)
#define TR_SM_W (TR_SM_IM2COL + TR_IM2COL)             // conv2..4 (the blob's own section): [layer 3][kw 3][ci chunk 4][(kh, co) 96][8] bf16
#define TR_W_DC_BYTES (4 * 96 * 16)                    // one (layer, kernel column): 6 144 B
#define TR_SM_W1 (TR_SM_W + 9 * TR_W_DC_BYTES)         // conv1: [k chunk 2][co 32][8] bf16
#define TR_SM_BIAS (TR_SM_W1 + 1024)                   // float [4][32]
#define TR_SM_BARS (TR_SM_BIAS + 512)
#define TRB_JOB 0                                       // [2][18] the MMAs of tile job j of issuing warp h have completed
#define TRB_READY 36                                    // [2][8] output tile r of the group in buffer b is in shared memory
#define TRB_IM2COL 52                                   // conv1's A operand of the next group is in place
#define TRB_I2FREE 53                                   // conv1's MMAs of the current group have read it
#define TRB_O4DONE 54                                   // conv4's accumulators of the group have been read (16 warps)
#define TRB_W1 55                                      // conv1's weights and the biases have landed (bulk copy)
#define TRB_W 56                                       // conv2..conv4's weights have landed
#define TRB_COUNT 57
#define TR_SM_ITEMS (TR_SM_BARS + 8 * 58)              // uint32 [26]: layer, tile, last job, TMEM column of each epilogue item
#define TR_CJOBS 14                                    // jobs of conv2..conv4 per group and issuing warp
#define TR_SM_JOBS (TR_SM_ITEMS + 4 * 28)              // uint4 [2][14]: what the issuing thread needs to know about a job
#define TR_SM_TMEM_PTR (TR_SM_JOBS + 16 * 2 * TR_CJOBS)
#define TR_SM_TOTAL (TR_SM_TMEM_PTR + 16)
static_assert(TR_SM_TOTAL <= 232448, "shared memory per CTA on sm_100");
#define TR_WAIT_LIMIT (4000000000LL)

struct TrunkRowsParams {
    const uint64_t* own;       // [G] side-relative bitboards of the leaves
    const uint64_t* opp;
    const uint8_t* wblob;      // the blob of az_net_trunk (fused_net.pack_trunk_weights)
    __nv_bfloat16* features;   // [G][16 * 32] position-major, or the packed A operand layout of az_net_fcstack
    __nv_bfloat16* debug;      // optional [n_groups][1024][32] dump of the activations after `debug_layer` (row = tile * 128 + lane)
    int G;
    int debug_layer;           // 0 = off, 1..3 = conv1..conv3
    int features_packed;
    int late_trigger;
};

// a wait that outlasts ~2 s of SM clocks is a protocol bug: trap, so that the launch fails instead of hanging the device
__device__ __forceinline__ void tr_wait(uint32_t bar, uint32_t parity) {
    long long t0 = 0;
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        const long long t = clock64();
        if (t0 == 0) t0 = t;
        else if (t - t0 > TR_WAIT_LIMIT) __trap();
    }
}
// (Polling by one lane per warp does not pay: behind a branch it costs the issuing warps their uniform registers --
// 2 x slower --, predicated and voted it adds ~60 cycles to the ~150 a poll of a long-completed phase takes anyway.
// A poll issued early does not hide it either: its predicate is consumed where it is produced.)
// (Nor do shared-memory counters -- red.release by the writers, ld.acquire by the issuing warps -- instead of the
// "tile ready" mbarriers: 20.2 against 19.6 us on the same GPU.)
// arrivals that do not wait for the barrier's state word to come back
__device__ __forceinline__ void tr_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// tcgen05.mma with K-major no-swizzle descriptors given by their low words (start address >> 4 | LBO >> 4 << 16; the
// high word -- SBO = 128 bytes, descriptor version 1 -- is the same for every operand of this kernel): moving an
// operand by d bytes adds d >> 4 to the low word (shared memory addresses stay below 2^18), which is all the issuing
// thread computes per instruction
__device__ __forceinline__ void tr_mma(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        ".reg .b64 da, db;\n\t"
        "mov.b64 da, {%1, %5};\n\t"
        "mov.b64 db, {%2, %5};\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %3, p;\n\t"
        "}" ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "r"(0x4008u) : "memory");
}
__device__ __forceinline__ uint32_t tr_desc_lo(uint32_t smem_addr, uint32_t lbo) {
    return ((smem_addr >> 4) & 0x3FFFu) | ((lbo >> 4) << 16);
}

// Per layer L = 0..3 (conv1..conv4): first MMA tile, number of MMA tiles, job offset inside a group; first / last
// output tile that is used.
__device__ __forceinline__ int tr_mlo(int L) { return L == 3 ? 1 : 0; }
__device__ __forceinline__ int tr_mcnt(int L) { return L == 3 ? 6 : 8; }
__device__ __forceinline__ int tr_olo(int L) { return L < 2 ? 0 : L - 1; }          // conv3: 1, conv4: 2
__device__ __forceinline__ int tr_ohi(int L) { return L < 2 ? 7 : 8 - L; }          // conv3: 6, conv4: 5
// Issuing warp h owns the accumulators of the output tiles 4 h .. 4 h + 3: the output tiles it produces in layer L, the
// input tiles that contribute to them (its tile jobs: the two tiles next to the border are jobs of both warps, each
// with the kernel rows that land on its own side), its first job of layer L among its 18 jobs of a group.
__device__ __forceinline__ int tr_holo(int h, int L) { return max(tr_olo(L), 4 * h); }
__device__ __forceinline__ int tr_hohi(int h, int L) { return min(tr_ohi(L), 4 * h + 3); }
__device__ __forceinline__ int tr_tlo(int h, int L) { return L == 0 ? 4 * h : max(tr_mlo(L), tr_holo(h, L) - 1); }
__device__ __forceinline__ int tr_thi(int h, int L) { return L == 0 ? 4 * h + 3 : min(tr_mlo(L) + tr_mcnt(L) - 1, tr_hohi(h, L) + 1); }
__device__ __forceinline__ int tr_jloc(int L) { return L == 0 ? 0 : 5 * L - 1; }       // 0, 4, 9, 14
// TMEM column of the accumulator of output tile r of layer L.  (conv4's four accumulators at the two ends of its region,
// where conv2 of the next group writes last, so that its wait for conv4's items comes two jobs later: measured, no gain.)
__device__ __forceinline__ uint32_t tr_ocol(int L, int r) { return (uint32_t)((L & 1) * 256 + 32 * (7 - r)); }

// conv1's A operand for one row (tile r, lane): the 9 taps of player * grid around cell (r, c) as bf16 (+1 own, -1 opp,
// 0 empty / off-board), then zeros up to K = 16; c < 0 (pad column, unused lane, missing board) = a zero row.
__device__ __forceinline__ void tr_im2col_row(uint8_t* im2col, int row, int r, int c, uint64_t own, uint64_t opp) {
    uint32_t w[5] = {0u, 0u, 0u, 0u, 0u};                  // taps 0..8 as bf16 pairs
    if (c >= 0) {
#pragma unroll
        for (int dr = 0; dr < 3; ++dr) {
            const int ur = r + dr - 1;
            uint32_t o3 = 0u, p3 = 0u;                     // cells (ur, c-1 .. c+1) as bits 0..2
            if (ur >= 0 && ur < 8) {
                const uint32_t orow = (uint32_t)(own >> (ur * 8)) & 0xFFu, prow = (uint32_t)(opp >> (ur * 8)) & 0xFFu;
                o3 = ((orow << 1) >> c) & 7u;
                p3 = ((prow << 1) >> c) & 7u;
            }
#pragma unroll
            for (int dc = 0; dc < 3; ++dc) {
                const int t = dr * 3 + dc;
                const uint32_t v = ((o3 >> dc) & 1u) ? 0x3F80u : (((p3 >> dc) & 1u) ? 0xBF80u : 0u);   // +-1.0 bf16
                w[t >> 1] |= v << ((t & 1) * 16);
            }
        }
    }
    uint8_t* dst = im2col + (TR_MARGIN + row) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(dst + TR_PLANE) = make_uint4(w[4], 0u, 0u, 0u);
}

__global__ void __launch_bounds__(TR_THREADS, 1) k_net_trunk_rows(TrunkRowsParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_groups = (p.G + TR_NB - 1) / TR_NB;
    // debug_layer 200: cycle stamps instead of activations, int64 [n_ctas][64] (tools/trunk_rows_probe.py; [16 + jj] = what
    // job jj of conv2..4 of the CTA's second group waited for its input tile)
    // (timing only, wrong results: 201 = the same without the issuing warps' waits for the epilogue, 202 = also with
    // N = 32 MMAs only, 203 = with the waits but half the MMAs -- what the dependencies / the tensor pipe / the issue cost)
    const int hack = (p.debug && p.debug_layer >= 200) ? p.debug_layer - 200 : 0;
    long long* stamps = (p.debug && p.debug_layer >= 200) ? reinterpret_cast<long long*>(p.debug) + (size_t)blockIdx.x * 64 : nullptr;
    const long long t_entry = stamps ? clock64() : 0;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = smem_u32(smem + TR_SM_BARS);
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + TR_SM_TMEM_PTR);

    // ---- one-time setup: barriers, weights (bulk copies), the item / job tables, zeroed margins and unused lanes, TMEM
    if (tid < TRB_COUNT) {
        // arrivals are per WARP (lane 0 after __syncwarp): 32 lanes arriving on one barrier word serialise
        uint32_t count = 1u;                                       // job[][]: one tcgen05.commit
        if (tid >= TRB_READY && tid < TRB_IM2COL) count = 4u;        // ready[][]: the team that wrote the tile
        else if (tid == TRB_IM2COL) count = TR_PRODUCERS / 32;
        else if (tid == TRB_I2FREE) count = TR_MMA_WARPS;           // each issuing warp commits after its conv1 jobs
        else if (tid == TRB_O4DONE) count = 4u * 4u;                // conv4's four output tiles, 4 warps each
        mbar_init(bars + 8 * tid, count);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // the weights by two asynchronous bulk copies (they do not depend on the launch before this one): conv1's tile +
        // the four bias rows, which the first MMAs and items need, and conv2..conv4 in the kernel's own order
        // (AZ_TRUNK_ROWS_BLOB_OFFSET of the blob), which nothing needs before conv1 is through
        if (tid == TRB_W1 || tid == TRB_W) {
            const uint32_t bytes = tid == TRB_W1 ? 1024u + 512u : (uint32_t)AZ_TRUNK_ROWS_BLOB_BYTES;
            const uint8_t* src = p.wblob + (tid == TRB_W1 ? AZ_TRUNK_ROWS_BLOB_OFFSET - 1536 : AZ_TRUNK_ROWS_BLOB_OFFSET);
            const uint32_t dst = sbase + (tid == TRB_W1 ? TR_SM_W1 : TR_SM_W);
            asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}"
                         ::"r"(bars + 8 * tid), "r"(bytes) : "memory");
            asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                         ::"r"(dst), "l"(src), "r"(bytes), "r"(bars + 8 * tid) : "memory");
        }
    }
    if (tid >= 64 && tid < 64 + TR_ITEMS) {
        // item idx of a group: layer L, output tile r, the job whose completion completes the tile's accumulator (of the
        // warp that owns it, its last input tile that contributes: r + 1, or the warp's last of the layer), the
        // accumulator's TMEM column.  The item overwrites INPUT tile r in place: the tiles next to the border (3, 4)
        // are read by the other warp too, whose job on that tile is the item's second barrier.
        const int idx = tid - 64;
        const int L = idx < 8 ? 0 : (idx < 16 ? 1 : (idx < 22 ? 2 : 3));
        // (both the jobs of a warp and the items of a layer go from the border outwards -- tiles 3, 4, 2, 5, 1, 6, 0, 7 --
        // so that what the next layer needs first is complete first)
        const int pos = idx - (L == 0 ? 0 : (L == 1 ? 8 : (L == 2 ? 16 : 22)));
        const int r = (pos & 1) ? 4 + (pos >> 1) : 3 - (pos >> 1);
        const int h = r >> 2;
        int j;                                                     // the step of warp h that completes the accumulator
        if (L == 0) j = h ? r - 4 : 3 - r;
        else j = h ? min(r + 1, tr_thi(1, L)) - tr_tlo(1, L) : tr_thi(0, L) - max(r - 1, tr_tlo(0, L));
        const int job = h * TR_HJOBS + tr_jloc(L) + j;
        int job2 = job;
        if (L > 0 && (r == 3 || r == 4)) job2 = (1 - h) * TR_HJOBS + tr_jloc(L);      // tile 3 / 4 is the other warp's first job
        reinterpret_cast<uint32_t*>(smem + TR_SM_ITEMS)[idx] =
            (uint32_t)L | ((uint32_t)r << 4) | ((uint32_t)job << 8) | ((uint32_t)job2 << 14) | (tr_ocol(L, r) << 20);
    }
    if (tid >= 96 && tid < 96 + 2 * TR_CJOBS) {
        // job jj of issuing warp h in conv2..conv4 (its step j of layer L): everything about it that does not depend on the
        // group.  Warp 0 goes down from the border (tiles 4, 3, ..), warp 1 up (3, 4, ..).
        const int h = (tid - 96) / TR_CJOBS, jj = (tid - 96) % TR_CJOBS;
        const int L = jj < 5 ? 1 : (jj < 10 ? 2 : 3), j = jj - (5 * L - 5);
        const int tlo = tr_tlo(h, L), thi = tr_thi(h, L), olo = tr_holo(h, L), ohi = tr_hohi(h, L);
        const int r = h ? tlo + j : thi - j;
        // kernel rows whose output tile r - kh + 1 this warp produces: kh = klo .. khi, 32 accumulator columns each, output
        // tiles descending = columns ascending from that of tile r - klo + 1
        const int klo = max(0, r + 1 - ohi), khi = min(2, r + 1 - olo), nblk = khi - klo + 1;
        const uint32_t d_col = tr_ocol(L, r - klo + 1);
        // the first contribution to an output tile must not accumulate: all of the warp's first job of the layer; of every
        // other, the block of the output tile it reaches first -- tile r - 1 (kh = 2, the last block) going down, tile
        // r + 1 (kh = 0, the first block) going up -- by an N = 32 MMA of its own before the rest
        const bool first = j == 0, init = !first && (h ? klo == 0 : khi == 2);
        uint32_t z = d_col | ((uint32_t)nblk << 9);
        if (!init) z |= (0u << 11) | ((uint32_t)nblk << 13) | ((first ? 0u : 1u) << 15);
        else {
            const uint32_t ipos = h ? 0u : (uint32_t)(nblk - 1), rpos = h ? 1u : 0u;
            z |= (ipos << 11) | (1u << 13) | (0u << 15);
            if (nblk > 1) z |= (1u << 16) | (rpos << 17) | ((uint32_t)(nblk - 1) << 19);
        }
        if (L == 1 && j == 0) z |= 1u << 21;
        const uint32_t wl = sbase + TR_SM_W + (uint32_t)((L - 1) * 3) * TR_W_DC_BYTES + (uint32_t)(klo * 32 * 16);
        reinterpret_cast<uint4*>(smem + TR_SM_JOBS)[tid - 96] =
            make_uint4((uint32_t)(TR_MARGIN + r * 128 - 1),            // A: the tile's rows shifted by -1 column, in 16-byte rows
                       tr_desc_lo(wl, 96u * 16u), z,
                       (uint32_t)r | (((r >= 1 && r <= 6) ? 3u : 2u) << 3) | ((uint32_t)(L - 1) << 5));
    }
    // the rows nothing ever writes: the margins of the activation buffers (the row before tile 0 is lane 0's left
    // neighbour = zero padding) and the two unused lanes of every tile (lane 127 is the left neighbour of the next
    // tile's lane 0); the pad lanes are written with every conv1 output
    if (tid < 2 * 4 * 2 * TR_MARGIN) {
        const int bufi = tid >> 6, pl = (tid >> 4) & 3, side = (tid >> 3) & 1, row = tid & 7;
        *reinterpret_cast<uint4*>(smem + TR_SM_BUF(bufi) + pl * TR_PLANE + (side ? TR_MARGIN + TR_ROWS + row : row) * 16) =
            make_uint4(0u, 0u, 0u, 0u);
    } else if (tid < 2 * 4 * 2 * TR_MARGIN + 2 * 4 * TR_TILES * 2) {
        const int u = tid - 2 * 4 * 2 * TR_MARGIN;
        const int bufi = u >> 6, pl = (u >> 4) & 3, tile = (u >> 1) & 7, ln = TR_NB * TR_PITCH + (u & 1);
        *reinterpret_cast<uint4*>(smem + TR_SM_BUF(bufi) + pl * TR_PLANE + (TR_MARGIN + tile * 128 + ln) * 16) =
            make_uint4(0u, 0u, 0u, 0u);
    }
    if (warp == TR_MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr_smem;
    // launched with az_launch_pdl: everything above may run beside the tail of the tree step that selects this launch's
    // leaves; nothing below may
    az_pdl_wait();
    if (!p.late_trigger) az_pdl_trigger();
    if (stamps && tid == 0) stamps[0] = clock64() - t_entry;

    {   // conv1's A operand of the CTA's first group, by all threads: two rows each, both boards loaded before either row
        // is built (the loads are the first thing after the wait and come from the L2: their latency is most of this)
        static_assert(TR_ROWS == 2 * TR_THREADS, "two rows per thread");
        const int board0 = blockIdx.x * TR_NB, n_boards = min(TR_NB, p.G - board0);
        uint64_t o[2], q[2];
        int cc[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int l = (tid + j * TR_THREADS) & 127, bb = l / TR_PITCH;
            cc[j] = l - bb * TR_PITCH;
            const bool ok = bb < n_boards && l < TR_NB * TR_PITCH && cc[j] < 8;
            o[j] = ok ? p.own[board0 + bb] : 0ULL;
            q[j] = ok ? p.opp[board0 + bb] : 0ULL;
            if (!ok) cc[j] = -1;
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) tr_im2col_row(smem + TR_SM_IM2COL, tid + j * TR_THREADS, (tid + j * TR_THREADS) >> 7, cc[j], o[j], q[j]);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncthreads();
        if (tid < TR_PRODUCERS / 32) mbar_arrive(bars + 8 * TRB_IM2COL);
        if (stamps && tid == 0) stamps[1] = clock64() - t_entry;
    }

    if (warp >= TR_MMA_WARP + TR_MMA_WARPS) {
        // ================================================================== producers: conv1's A operand (9-tap im2col
        // of player * grid, K = 16) of every group of this CTA.  Thread = lanes ptid and ptid + 64 of all 8 tiles.
        const int ptid = tid - (TR_EPI_THREADS + 32 * TR_MMA_WARPS);
        uint8_t* im2col = smem + TR_SM_IM2COL;
        // (the first group's is built by all threads above)
        uint32_t k = 1;
        for (int grp = blockIdx.x + gridDim.x; grp < n_groups; grp += gridDim.x, ++k) {
            const int board0 = grp * TR_NB, n_boards = min(TR_NB, p.G - board0);
            uint64_t own[2], opp[2];
            int col[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int tl = ptid + 64 * j, b = tl / TR_PITCH;
                col[j] = tl - b * TR_PITCH;
                const bool ok = b < n_boards && tl < TR_NB * TR_PITCH && col[j] < 8;
                own[j] = ok ? p.own[board0 + b] : 0ULL;
                opp[j] = ok ? p.opp[board0 + b] : 0ULL;
                if (!ok) col[j] = -1;
            }
            tr_wait(bars + 8 * TRB_I2FREE, (k - 1u) & 1u);                 // conv1 of the previous group has read the buffer
#pragma unroll
            for (int j = 0; j < 2; ++j) {
#pragma unroll 2
                for (int r = 0; r < TR_TILES; ++r) {
                    tr_im2col_row(im2col, r * 128 + ptid + 64 * j, r, col[j], own[j], opp[j]);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + 8 * TRB_IM2COL);
        }
    } else if (warp >= TR_MMA_WARP) {
        // ================================================================== MMA issue: warp h issues the tile jobs that
        // accumulate into the output tiles 4 h .. 4 h + 3, in program order, by one elected thread of a converged warp
        // (a single lane in divergent code costs an election and seven register broadcasts per MMA: measured slower).
        // The issuing thread is the serial resource of this kernel (a tcgen05.mma costs it ~65 cycles; two of them
        // keep the tensor pipe busy), so what a job needs beyond the group's buffer is in the table built at set-up.
        const int h = warp - TR_MMA_WARP;
        const uint32_t jobs = bars + 8 * (TRB_JOB + h * TR_HJOBS);
        const uint32_t idesc32 = umma_idesc_bf16_m128(32u);
        constexpr uint32_t IDESC_BLK = (32u >> 3) << 17;          // one more kernel row = 32 more columns
        const uint4* table = reinterpret_cast<const uint4*>(smem + TR_SM_JOBS) + h * TR_CJOBS;
        const uint32_t w1_lo = tr_desc_lo(sbase + TR_SM_W1, 512u);
        uint32_t k = 0;                               // group counter of this CTA
        long long w_ready = 0, w_o4 = 0, t_issue = 0, w_im2col = 0;
        for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++k) {
            const uint32_t buf = k & 1u, use = k >> 1;
            const long long c_i = stamps ? clock64() : 0;
            tr_wait(bars + 8 * TRB_IM2COL, k & 1u);
            if (k == 0) tr_wait(bars + 8 * TRB_W1, 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const long long c_j = stamps ? clock64() : 0;
            if (stamps) w_im2col += c_j - c_i;
            // conv1: one K = 16 MMA per tile into its own accumulator (region 0: conv3's of the previous group have been
            // read -- its conv4 jobs waited for every conv3 output tile)
            if (elect_one()) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = h ? 4 + i : 3 - i;
                    tr_mma(tmem_base + tr_ocol(0, r), tr_desc_lo(sbase + TR_SM_IM2COL + (uint32_t)(TR_MARGIN + r * 128) * 16u, TR_PLANE), w1_lo,
                           idesc32, 0u);
                    umma_commit(jobs + 8 * i);
                }
                umma_commit(bars + 8 * TRB_I2FREE);    // conv1 has read the im2col buffer: the producers may build the next group's
            }
            __syncwarp();
            if (stamps) t_issue += clock64() - c_j;
            if (stamps && lane == 0 && k == 0 && h == 0) stamps[8] = clock64() - t_entry;
            if (k == 0) tr_wait(bars + 8 * TRB_W, 0u);
            const uint32_t abuf_lo = tr_desc_lo(sbase + TR_SM_BUF(buf), TR_PLANE);
            const uint32_t ready = bars + 8 * (TRB_READY + buf * 8);
#pragma unroll 1
            for (int jj = 0; jj < TR_CJOBS; ++jj) {
                const uint4 e = table[jj];
                const long long c0 = stamps ? clock64() : 0;
                {   // the input: layer L - 1's output tile r = phase (use * P + L - 1) of ready[buf][r], P = layers that write tile r
                    const uint32_t r = e.w & 7u, P = (e.w >> 3) & 3u, Lm1 = (e.w >> 5) & 3u;
                    if (hack == 0 || hack == 3) tr_wait(ready + 8 * r, (use * P + Lm1) & 1u);
                }
                const long long c1 = stamps ? clock64() : 0;
                // region 1 still holds conv4's accumulators of the previous group until its four epilogue items have read them
                if (((e.z >> 21) & 1u) && k > 0 && (hack == 0 || hack == 3)) tr_wait(bars + 8 * TRB_O4DONE, (k - 1u) & 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const long long c2 = stamps ? clock64() : 0;
                const uint32_t d_tmem = tmem_base + (e.z & 511u);
                const uint32_t idesc = idesc32 + (hack == 2 ? 0u : (((e.z >> 9) & 3u) - 1u) * IDESC_BLK);
                const uint32_t a_lo = abuf_lo + e.x, b_lo = e.y;
                const uint32_t ez = e.z;
                if (elect_one()) {
                {   // kernel column 0, channels 0..15: whole, or the block that must not accumulate and then the rest
                    const uint32_t pos = (ez >> 11) & 3u;
                    tr_mma(d_tmem + 32u * pos, a_lo, b_lo + 32u * pos, idesc32 + (((ez >> 13) & 3u) - 1u) * IDESC_BLK, (ez >> 15) & 1u);
                    if ((ez >> 16) & 1u) {
                        const uint32_t pos2 = (ez >> 17) & 3u;
                        tr_mma(d_tmem + 32u * pos2, a_lo, b_lo + 32u * pos2, idesc32 + (((ez >> 19) & 3u) - 1u) * IDESC_BLK, 1u);
                    }
                }
#pragma unroll
                for (int i = 1; i < 6; ++i) {
                    const int kw = i >> 1, kk = i & 1;
                    if (hack == 3 && kk) continue;               // timing only: half the MMAs
                    tr_mma(d_tmem, a_lo + (uint32_t)(kw + kk * (2 * TR_PLANE >> 4)),
                           b_lo + (uint32_t)(kw * (TR_W_DC_BYTES >> 4) + kk * (2 * 96 * 16 >> 4)), idesc, 1u);
                }
                umma_commit(jobs + 8 * (4 + jj));
                }
                __syncwarp();
                if (stamps) { const long long c3 = clock64(); w_ready += c1 - c0; w_o4 += c2 - c1; t_issue += c3 - c2; }
                if (stamps && lane == 0 && k == 1 && h == 0) stamps[16 + jj] = c1 - c0;
                if (stamps && lane == 0 && k == 1 && h == 0 && jj >= 5 && jj < 10) { stamps[32 + jj - 5] = c2 - t_entry; stamps[37 + jj - 5] = clock64() - t_entry; }
                if (stamps && lane == 0 && k == 0 && h == 0 && (jj == 4 || jj == 9)) stamps[jj == 4 ? 9 : 10] = clock64() - t_entry;   // group 0: conv2 / conv3 issued
            }
            if (stamps && lane == 0 && k == 0 && h == 0) stamps[3] = clock64() - t_entry;
        }
        if (stamps && lane == 0 && h == 0) { stamps[5] = w_ready; stamps[6] = w_o4; stamps[7] = t_issue; stamps[12] = w_im2col; }
    } else {
        // ================================================================== epilogue team `team`: items team, team + 3, ..
        const int team = warp >> 2, quarter = warp & 3;
        const int tl = quarter * 32 + lane;                     // TMEM lane = lane of the tile
        const int b = tl / TR_PITCH, c = tl - b * TR_PITCH;     // board of the group, column (8 = pad)
        const float* bias_all = reinterpret_cast<const float*>(smem + TR_SM_BIAS);
        const uint32_t t_quarter = tmem_base + ((uint32_t)(quarter * 32) << 16);
        tr_wait(bars + 8 * TRB_W1, 0u);                          // the biases
        uint32_t k = 0;
        long long e_wait = 0, e_work = 0;
        for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++k) {
            const uint32_t buf = k & 1u;
            const int board0 = grp * TR_NB, n_boards = min(TR_NB, p.G - board0);
            const bool used = tl < TR_NB * TR_PITCH;                // not one of the tile's two spare lanes
            const bool cell = b < n_boards && used && c < 8;        // a real cell of a real board
            // 26 items per group, taken round robin across the groups: item idx of group k by team (26 k + idx) % TR_TEAMS
#pragma unroll 1
            for (int idx = (int)((uint32_t)(team + (TR_TEAMS - TR_ITEMS % TR_TEAMS) * (int)(k % TR_TEAMS)) % TR_TEAMS); idx < TR_ITEMS; idx += TR_TEAMS) {
                const uint32_t it = reinterpret_cast<const uint32_t*>(smem + TR_SM_ITEMS)[idx];
                const int L = (int)(it & 3u), r = (int)((it >> 4) & 15u);
                const long long f0 = stamps ? clock64() : 0;
                tr_wait(bars + 8 * (TRB_JOB + ((it >> 8) & 63u)), k & 1u);
                if (((it >> 14) & 63u) != ((it >> 8) & 63u)) tr_wait(bars + 8 * (TRB_JOB + ((it >> 14) & 63u)), k & 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const long long f1 = stamps ? clock64() : 0;
                e_wait += f1 - f0;
                const float* bias = bias_all + L * 32;
                uint32_t acc[32], out[16];
                tmem_ld32(t_quarter + (it >> 20), acc);
#pragma unroll
                for (int co = 0; co < 32; co += 2) {
                    const float2 bb = *reinterpret_cast<const float2*>(bias + co);
                    out[co >> 1] = pack_relu_bf16x2(__uint_as_float(acc[co]) + bb.x, __uint_as_float(acc[co + 1]) + bb.y);
                }
                // ---- store: activations in place (tile r of the group's buffer) or, after conv4, the features
                if (L < 3) {
                    // conv2 pads with zeros: conv1's output must be zero in the pad column and the boards a partial group
                    // lacks.  (conv3 / conv4 do not pad: what they read there only reaches outputs that are not valid
                    // anyway, and it is finite.)  The spare lanes stay zero: lane 127 is what lane 0 of the NEXT tile
                    // reads as its left neighbour, possibly before or after this item runs.
                    if (L == 0 && !cell) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) out[i] = 0u;
                    }
                    if (used) {
                        uint8_t* d = smem + TR_SM_BUF(buf) + (size_t)(TR_MARGIN + r * 128 + tl) * 16;
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(d + ch * TR_PLANE) = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) tr_arrive(bars + 8 * (TRB_READY + buf * 8 + r));
                    if (stamps && lane == 0 && quarter == 0 && k == 1 && L == 1) { stamps[42 + idx - 8] = f1 - t_entry; stamps[50 + idx - 8] = clock64() - t_entry; }
                    if (p.debug && p.debug_layer == L + 1) {
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(p.debug + ((size_t)grp * TR_ROWS + r * 128 + tl) * 32 + ch * 8) =
                                used ? make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]) : make_uint4(0u, 0u, 0u, 0u);
                    }
                } else {
                    // the accumulator has been read (tmem_ld32 waits for the data): conv2 of the next group may overwrite it
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) tr_arrive(bars + 8 * TRB_O4DONE);
                    if (cell && c >= 2 && c < 6) {
                        // conv4: the outputs are the cells at an inset of 2; features[board][pos * 32 + co], pos = (r-2)*4 + (c-2)
                        const int pos = (r - 2) * 4 + (c - 2), g = board0 + b;
                        if (p.features_packed) {
                            // K = pos*32 + 8*ch: chunk pos / 2 (64 wide, 16 KB per row tile of 128 positions), plane (pos % 2)*4 + ch
                            uint8_t* f = reinterpret_cast<uint8_t*>(p.features) + ((size_t)(g >> 7) * 8 + (size_t)(pos >> 1)) * 16384 +
                                         (size_t)((pos & 1) * 4) * 2048 + (size_t)(g & 127) * 16;
#pragma unroll
                            for (int ch = 0; ch < 4; ++ch)
                                *reinterpret_cast<uint4*>(f + ch * 2048) = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                        } else {
                            uint4* f = reinterpret_cast<uint4*>(p.features + ((size_t)g * 16 + pos) * 32);
#pragma unroll
                            for (int ch = 0; ch < 4; ++ch)
                                f[ch] = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                        }
                    }
                }
                if (stamps) e_work += clock64() - f1;
            }
        }
        if (stamps && lane == 0 && quarter == 0 && team == 0) { stamps[13] = e_wait; stamps[14] = e_work; }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == TR_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    if (stamps && tid == 0) stamps[1 + 10] = clock64() - t_entry;      // [11] exit
}

// az_net_trunk takes this kernel for every 8 x 8 batch of AZ_TRUNK_ROWS_MIN_G positions and more unless
// AZ_B200_TRUNK_V1 is set in the environment (A/B switch) or bit 1 of its `features_packed` asks for the first kernel.
bool az_trunk_rows_enabled() {
    static const bool on = getenv("AZ_B200_TRUNK_V1") == nullptr;
    return on;
}

int az_trunk_rows_launch(const uint64_t* own, const uint64_t* opp, int64_t G, const void* wblob, void* features,
                     int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream) {
    static bool attr_set = false;
    if (!attr_set) {
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk_rows, cudaFuncAttributeMaxDynamicSharedMemorySize, TR_SM_TOTAL));
        attr_set = true;
    }
    TrunkRowsParams p;
    p.own = own; p.opp = opp; p.wblob = (const uint8_t*)wblob; p.features = (__nv_bfloat16*)features;
    p.debug = (__nv_bfloat16*)debug; p.G = (int)G; p.debug_layer = debug ? debug_layer : 0;
    p.features_packed = features_packed ? 1 : 0; p.late_trigger = late_trigger;
    const int n_groups = (int)((G + TR_NB - 1) / TR_NB);
    const int grid = n_groups < n_sms ? n_groups : n_sms;      // persistent: one CTA per SM
    cudaError_t e = az_launch_pdl(k_net_trunk_rows, dim3(grid), dim3(TR_THREADS), TR_SM_TOTAL, (cudaStream_t)stream, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_trunk_rows");
    AZ_CHECK_LAUNCH("k_net_trunk_rows");
    return AZ_OK;
}

