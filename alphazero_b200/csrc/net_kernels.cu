// Fused convolution trunk of the Othello policy-value network on the 5th-gen tensor cores (tcgen05 + TMEM),
// plus the fused policy/value head epilogue.  sm_100a only.
//
// Replaces, for the leaf batch of the MCTS step, conv1..conv4 + BatchNorm + ReLU of OthelloNet.forward
// (alphazero/games/othello.py:366-374).  Design:
//   * one persistent CTA per SM loops over groups of NB boards; all activations of a group stay in shared memory,
//     nothing but the 16-byte bitboards is read from HBM and only the 32*(n-4)^2 features are written;
//   * row space: a board is n rows of 8 cells (pitch 8, pad cells zero) preceded by one zero "ring" row of 8 cells
//     that it shares with the board before it: row = board*(n+1)*8 + (1+r)*8 + c.  32-lane warps and 8-row tensor
//     core matrices therefore never straddle a board row;
//   * a 3x3 convolution = 3 accumulating GEMM steps, one per kernel ROW dr: the A operand of `tcgen05.mma` is the
//     activation buffer with its descriptor start moved by dr*8 rows (K-major no-swizzle layout, 16 bytes per row per
//     8-channel chunk, so the shift is a whole number of 128-byte core matrices) and the B operand holds the THREE taps
//     of that kernel row side by side: N = 96 = (dc, c_out).  The accumulator row p then holds, per dc, the partial
//     sum that belongs to output p - dc; the epilogue adds the three blocks with two warp shuffles per channel
//     (lane +-1 = column +-1; the zero padding / row wrap at columns 0 and 7 is a predicate, not data).
//     Measured on B200, an M128 N32 K16 SS-mode MMA costs ~90 cycles of which ~50 are a per-instruction cost that
//     does not depend on N, so tripling N per instruction is what pays (profiles/r1_trunk_experiments.txt);
//   * B operands (BatchNorm folded in) sit in shared memory for the whole kernel;
//   * accumulators (128 rows x 96 fp32 columns) live in TMEM in four stages; two teams of four epilogue warps read
//     them with tcgen05.ld, combine the dc blocks, add the folded bias, apply ReLU, round to bf16 and write the next
//     layer's A operand;
//   * conv1 (1 input channel, K = 9 padded to 16) is one N = 32 MMA per tile on a 9-tap im2col built straight from
//     the bitboards (values -1/0/+1 are exact in bf16); its epilogue writes zeros outside the n x n cells, which is
//     the zero padding conv2 reads.
// Garbage computed on ring rows / pad cells / tail rows after conv1 is never read by a valid output: conv2 (pad 1)
// reads only conv1's zeros there, conv3 / conv4 (no padding) produce valid outputs only at an inset of 1 / 2 cells.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "net_heads.cuh"
#include "umma.cuh"
#include "trunk_rows.cuh"

#define NET_CH 32                 // channels of every conv layer
#define NET_TILE_M 128            // rows per MMA tile
#define NET_TILES 4               // tiles per group (at most)
#define NET_ROWS (NET_TILE_M * NET_TILES)   // 512 rows of activation per buffer
#define NET_DEBUG_ROWS 768        // row stride of the optional debug dump (kept from the 6-tile version of the ABI)
#define NET_SLOTS 2               // groups in flight per CTA, each with its own buffers, MMA warp and epilogue team
#define NET_PITCH 8               // cells per board row in the row space
#define NET_MARGIN 16             // zero rows before / after (|row shift| = 8)
#define NET_BUF_ROWS (NET_ROWS + 2 * NET_MARGIN)   // 544
#define NET_CHUNK_BYTES (NET_BUF_ROWS * 16)        // one 8-channel chunk plane: 8704 B  (= LBO of the A operand)
#define NET_BUF_BYTES (4 * NET_CHUNK_BYTES)        // 34816 B per activation buffer
#define NET_N3 96                 // N of the conv2..4 MMAs: 3 taps (dc) x 32 output channels
#define NET_DR_CHUNK_BYTES (NET_N3 * 16)           // one 8-input-channel chunk of a kernel row: 1536 B (= LBO of B)
#define NET_DR_BYTES (4 * NET_DR_CHUNK_BYTES)      // one kernel row: [ci chunk 4][(dc, co) 96][ci % 8] bf16 = 6144 B
#define NET_LAYER_BYTES (3 * NET_DR_BYTES)         // 18432
#define NET_CONV1_TILE_BYTES 1024 // conv1 as one K=16 MMA: [2 chunks][32 co][8] bf16 (taps 0..8, zero padded)
#define NET_WBLOB_MMA_BYTES (3 * NET_LAYER_BYTES + NET_CONV1_TILE_BYTES)  // conv2..conv4 tiles, then the conv1 tile
#define NET_WBLOB_F32 (4 * 32)                     // folded biases of conv1..conv4
#define NET_WBLOB_BYTES (NET_WBLOB_MMA_BYTES + NET_WBLOB_F32 * 4)
#define NET_WORKERS 256           // warps 0-7: epilogue, one team of 4 warps per slot (warp % 4 = TMEM lane quarter)
#define NET_MMA_WARPS NET_SLOTS   // warps 8-9: MMA issue, one warp per slot.  Issuing a tile costs the thread ~700 cycles
                                  // (60 per tcgen05.mma + barrier bookkeeping) against ~340 cycles of tensor-pipe time
                                  // and ~730 of epilogue per team, so two independent slot pipelines keep all three busy
#define NET_PRODUCERS 64          // warps 10-11: im2col of the slot's NEXT group while the current one runs conv2..conv4
#define NET_THREADS (NET_WORKERS + 32 * NET_MMA_WARPS + NET_PRODUCERS)
#define NET_STAGE_COLS 128        // TMEM columns per accumulator stage (96 used)
#define NET_STAGES 4              // accumulator stages: two per slot
#define NET_TMEM_COLS (NET_STAGES * NET_STAGE_COLS)   // 512: all of the SM's TMEM (one CTA per SM)

// shared memory map (bytes): per slot two activation buffers (A, B), per slot one im2col buffer (conv1's A operand,
// K = 16: two chunk planes), the weight blob, the barriers
#define SM_BUF(slot, which) (((slot) * 2 + (which)) * NET_BUF_BYTES)
#define NET_IM2COL_BYTES (2 * NET_CHUNK_BYTES)
#define SM_IM2COL(slot) (2 * NET_SLOTS * NET_BUF_BYTES + (slot) * NET_IM2COL_BYTES)
#define SM_W (SM_IM2COL(NET_SLOTS))
// mbarriers: full / empty per accumulator stage; ready[slot][tile] "rows of tile t of this layer are in shared
// memory"; im2col[slot] "conv1's A operand of the slot's next group is in place"; i2free[slot] "the slot's im2col
// buffer is no longer read" (conv1 MMAs done); weights "the bulk copy of the weight blob has landed"
#define SM_BARS (SM_W + NET_WBLOB_BYTES)
#define BAR_FULL 0
#define BAR_EMPTY NET_STAGES
#define BAR_READY (2 * NET_STAGES)
#define BAR_IM2COL (BAR_READY + NET_SLOTS * NET_TILES)
#define BAR_I2FREE (BAR_IM2COL + NET_SLOTS)
#define BAR_WEIGHTS (BAR_I2FREE + NET_SLOTS)
#define NET_N_BARS (BAR_WEIGHTS + 1)
#define SM_TMEM_PTR (SM_BARS + 24 * 8)   // room for NET_N_BARS = 21 barriers
#define SM_TOTAL (SM_TMEM_PTR + 16)
static_assert(NET_N_BARS <= 24, "barrier area");
static_assert(SM_TOTAL <= 232448, "shared memory per CTA on sm_100");

struct NetTrunkParams {
    const uint64_t* own;       // [G] side-relative bitboards of the leaves
    const uint64_t* opp;       // [G]
    const uint8_t* wblob;      // NET_WBLOB_BYTES: packed weights (see alphazero_b200/fused_net.py)
    __nv_bfloat16* features;   // [G][(n-4)^2 * 32]  position-major, channel-minor
    __nv_bfloat16* debug;      // optional [n_groups][NET_ROWS][32] dump of the activations after `debug_layer`
    int G;
    int rows, cols;            // image size: 8x8 / 6x6 (Othello) or 7x6 (Connect4Net's view of the 6x7 grid)
    int boards_per_group;      // NB: NB * (n+1)*8 + 8 <= n_tiles * 128
    int n_tiles;               // 128-row tiles per group: 2 or 4
    int debug_layer;           // 0 = off, 1..3 = dump the output of conv1..conv3
    int late_trigger;          // 1: no early programmatic launch of the kernel that follows (A/B measurements)
    int features_packed;       // 0: features[G][pos*32 + co]; 1: the packed A operand layout of az_net_fcstack
                               //    [G / 128][K / 64][8 K planes][128 rows][8] with K index = pos*32 + co
};

// Row of the group's row space -> (board in group, board row r, column c); r = -1 on a ring row.
// rpb = rows per board = (n + 1) * 8.
__device__ __forceinline__ void net_row_coords(int row, int rpb, int& b, int& r, int& c) {
    b = row / rpb;
    const int rem = row - b * rpb;
    r = (rem >> 3) - 1;
    c = rem & 7;
}

// 9-tap im2col of the network input (player*grid: +1 own, -1 opp, 0 empty / off-board) for one row:
// 16 bf16 = taps 0..8 then zeros, written as the two 16-byte K chunks of the conv1 A operand.
// (b, r, c) are the row's coordinates; own/opp the bitboards of its board (0 when the board does not exist).
__device__ __forceinline__ void net_im2col_row(uint8_t* im2col, int row, int n_rows, int n_cols, int r, int c,
                                               uint64_t own, uint64_t opp) {
    uint32_t w[5] = {0u, 0u, 0u, 0u, 0u};     // taps 0..8 as bf16 pairs (tap 8 in the low half of w[4])
    if (r >= 0 && c < n_cols) {
#pragma unroll
        for (int dr = 0; dr < 3; ++dr) {
            const int ur = r + dr - 1;
            // the three cells (ur, c-1 .. c+1) as bits 0..2; cells outside the image are 0 in both bitboards
            uint32_t o3 = 0u, p3 = 0u;
            if (ur >= 0 && ur < n_rows) {
                const uint32_t orow = (uint32_t)(own >> (ur * 8)) & 0xFFu, prow = (uint32_t)(opp >> (ur * 8)) & 0xFFu;
                o3 = ((orow << 1) >> c) & 7u;
                p3 = ((prow << 1) >> c) & 7u;
            }
#pragma unroll
            for (int dc = 0; dc < 3; ++dc) {
                const int t = dr * 3 + dc;
                const uint32_t v = ((o3 >> dc) & 1u) ? 0x3F80u : (((p3 >> dc) & 1u) ? 0xBF80u : 0u);   // +1.0 / -1.0 bf16
                w[t >> 1] |= v << ((t & 1) * 16);
            }
        }
    }
    uint8_t* dst = im2col + (NET_MARGIN + row) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(dst + NET_CHUNK_BYTES) = make_uint4(w[4], 0u, 0u, 0u);
}

// Warp roles (no block-wide barrier inside the group loop; everything is mbarrier-pipelined).  The CTA keeps
// NET_SLOTS = 2 groups in flight; slot s owns the CTA's groups k = s, s + 2, ... with its own activation buffers,
// im2col buffer, two TMEM stages and barriers, so that the two pipelines never wait for each other:
//   warps 0-3 / 4-7   epilogue team of slot 0 / 1 (warp % 4 = TMEM lane quarter);
//   warp 8 / 9        MMA issue of slot 0 / 1 (one elected lane): tile t of layer L+1 only waits for tiles t-1 .. t+1 of
//                     layer L (row shifts are +-8 rows), so layers and groups of a slot follow each other without a drain;
//   warps 10-11       producers: im2col of the next group of either slot while that slot runs conv2..conv4.
// Measured per tile (B200): 6 MMAs occupy the tensor pipe ~340 cycles (operand fetch, 128 B per clock), cost the issuing
// thread ~700 cycles, and the epilogue team ~730 cycles (TMEM read of 96 columns + 64 shuffles per row) -- one slot
// alone leaves each of the three idle most of the time (profiles/r1_trunk_experiments.txt).
template <int IMG_ROWS, int IMG_COLS>
__global__ void __launch_bounds__(NET_THREADS, 1) k_net_trunk(NetTrunkParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const bool timing = p.debug && p.debug_layer == 100;   // cycle accounting into debug[blockIdx.x][0..15] (int64)
    const long long t_entry = timing ? clock64() : 0;
    unsigned long long g_entry = 0ULL;
    if (timing) asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_entry));
    constexpr int RPB = (IMG_ROWS + 1) * NET_PITCH;    // rows per board: ring row + image rows
    const int NB = p.boards_per_group;
    const int n_groups = (p.G + NB - 1) / NB;
    const int n_tiles = p.n_tiles, n_rows = n_tiles * NET_TILE_M;
    const uint32_t bars = smem_u32(smem + SM_BARS);
    const uint32_t bar_full0 = bars + 8 * BAR_FULL, bar_empty0 = bars + 8 * BAR_EMPTY, bar_ready0 = bars + 8 * BAR_READY;
    const uint32_t bar_im2col0 = bars + 8 * BAR_IM2COL, bar_i2free0 = bars + 8 * BAR_I2FREE, bar_weights = bars + 8 * BAR_WEIGHTS;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + SM_TMEM_PTR);
    constexpr int MMA_WARP = NET_WORKERS / 32;

    // ---- one-time setup: barriers, async bulk copy of the weights, zeroed activation buffers, TMEM
    if (tid == 0) {
        // (arrivals are per warp -- lane 0 after __syncwarp: 32 lanes arriving on one barrier word serialise)
        for (int st = 0; st < NET_STAGES; ++st) { mbar_init(bar_full0 + 8 * st, 1); mbar_init(bar_empty0 + 8 * st, 4); }
        for (int t = 0; t < NET_SLOTS * NET_TILES; ++t) mbar_init(bar_ready0 + 8 * t, 4);
        for (int sl = 0; sl < NET_SLOTS; ++sl) { mbar_init(bar_im2col0 + 8 * sl, NET_PRODUCERS / 32); mbar_init(bar_i2free0 + 8 * sl, 1); }
        mbar_init(bar_weights, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}"
                     ::"r"(bar_weights), "r"((uint32_t)NET_WBLOB_BYTES) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(smem + SM_W)), "l"(p.wblob), "r"((uint32_t)NET_WBLOB_BYTES), "r"(bar_weights) : "memory");
    }
    for (int i = tid; i < 2 * NET_SLOTS * NET_BUF_BYTES / 16; i += NET_THREADS)
        reinterpret_cast<uint4*>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
    if (warp == MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(NET_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr_smem;
    // launched with az_launch_pdl: everything above (barriers, TMEM, the weight copy in flight) may run beside the tail
    // of the tree step that selects this launch's leaves; nothing below may
    az_pdl_wait();
    // the fc stack that follows (az_net_fcstack, also launched with az_launch_pdl) may start its prologue on every SM a
    // CTA of this grid has left; it waits for the whole grid before it reads the features
    if (!p.late_trigger) az_pdl_trigger();
    const long long t_setup = timing ? clock64() : 0;

    if (warp >= MMA_WARP + NET_MMA_WARPS) {
        // ================================================================== producers: im2col of every group, in the
        // order the CTA starts them (k = 0, 1, 2, ...: slot k & 1, the slot's j-th group with j = k >> 1)
        const int ptid = tid - (NET_WORKERS + 32 * NET_MMA_WARPS);
        // (the first group of each slot is converted by the slot's epilogue team, which has nothing else to do yet)
        uint32_t k = NET_SLOTS;
        for (int grp = blockIdx.x + NET_SLOTS * gridDim.x; grp < n_groups; grp += gridDim.x, ++k) {
            const uint32_t slot = k & 1u, j = k >> 1;
            const int board0 = grp * NB;
            const int n_boards = min(NB, p.G - board0);
            // bitboards of the rows this thread converts (global loads issued before waiting for the buffer)
            constexpr int RPT = NET_ROWS / NET_PRODUCERS;
            uint64_t own[RPT], opp[RPT];
#pragma unroll
            for (int q = 0; q < RPT; ++q) {
                const int b = (ptid + q * NET_PRODUCERS) / RPB;
                const bool ok = b < n_boards && ptid + q * NET_PRODUCERS < n_rows;
                own[q] = ok ? p.own[board0 + b] : 0ULL;
                opp[q] = ok ? p.opp[board0 + b] : 0ULL;
            }
            mbar_wait(bar_i2free0 + 8 * slot, (j - 1u) & 1u);   // conv1 of the slot's previous group has read it
            uint8_t* im2col = smem + SM_IM2COL(slot);
#pragma unroll
            for (int q = 0; q < RPT; ++q) {
                const int row = ptid + q * NET_PRODUCERS;
                if (row < n_rows) {
                    int b, r, c;
                    net_row_coords(row, RPB, b, r, c);
                    net_im2col_row(im2col, row, IMG_ROWS, IMG_COLS, r, c, own[q], opp[q]);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bar_im2col0 + 8 * slot);
        }
    } else if (warp >= MMA_WARP) {
        // ================================================================== MMA issue of slot `slot`
        const uint32_t slot = (uint32_t)(warp - MMA_WARP);
        const uint32_t idesc32 = umma_idesc_bf16_m128(32u), idesc96 = umma_idesc_bf16_m128(NET_N3);
        const uint32_t bar_ready = bar_ready0 + 8 * NET_TILES * slot;
        mbar_wait(bar_weights, 0u);
        uint32_t j = 0, tile_it = 0;
        long long tm_ready = 0, tm_empty = 0, tm_issue = 0, tm_im2col = 0;
        const long long tm_start = timing ? clock64() : 0;
        for (int grp = blockIdx.x + (int)slot * gridDim.x; grp < n_groups; grp += NET_SLOTS * gridDim.x, ++j) {
            long long c0 = timing ? clock64() : 0;
            mbar_wait(bar_im2col0 + 8 * slot, j & 1u);
            if (timing) tm_im2col += clock64() - c0;
            // ---- layer 0: conv1 (im2col buffer -> A), 1: conv2 (A -> B), 2: conv3 (B -> A), 3: conv4 (A -> features)
            for (int layer = 0; layer < 4; ++layer) {
                const uint32_t src = smem_u32(smem + (layer == 0 ? SM_IM2COL(slot) : SM_BUF(slot, (layer & 1) ? 0 : 1)));
                const uint64_t a_layer = umma_desc(src + NET_MARGIN * 16, NET_CHUNK_BYTES, 128);
                const uint64_t b_layer = layer == 0
                    ? umma_desc(smem_u32(smem + SM_W) + 3 * NET_LAYER_BYTES, 512, 128)
                    : umma_desc(smem_u32(smem + SM_W) + (layer - 1) * NET_LAYER_BYTES, NET_DR_CHUNK_BYTES, 128);
                // ready[] barriers complete one phase per layer 0..2 of every group of the slot (conv4 writes no activations)
                const uint32_t ready_parity = (j * 3u + (uint32_t)(layer - 1)) & 1u;
                for (int t = 0; t < n_tiles; ++t, ++tile_it) {
                    long long c1 = timing ? clock64() : 0;
                    if (layer > 0) {
                        // input rows of tile t: tiles t-1 .. t+1 of the previous layer; t-1 and t were awaited for tile t-1
                        if (t == 0) mbar_wait(bar_ready, ready_parity);
                        if (t + 1 < n_tiles) mbar_wait(bar_ready + 8 * (t + 1), ready_parity);
                    }
                    long long c2 = timing ? clock64() : 0;
                    const uint32_t stage = slot * 2u + (tile_it & 1u), phase = (tile_it >> 1) & 1u;
                    mbar_wait(bar_empty0 + 8 * stage, phase ^ 1u);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    long long c3 = timing ? clock64() : 0;
                    const uint32_t d_tmem = tmem_base + stage * NET_STAGE_COLS;
                    const uint64_t a_tile = a_layer + (uint64_t)(t * (NET_TILE_M * 16 / 16));
                    if (elect_one()) {
                        if (layer == 0) {
                            umma_bf16(d_tmem, a_tile, b_layer, idesc32, 0u);
                        } else {
#pragma unroll
                            for (int dr = 0; dr < 3; ++dr) {
#pragma unroll
                                for (int kk = 0; kk < 2; ++kk) {
                                    // A: rows shifted by (dr - 1) board rows, input channels 16*kk .. 16*kk + 15
                                    const int64_t a_off = (dr - 1) * NET_PITCH + kk * (2 * NET_CHUNK_BYTES / 16);
                                    const uint64_t b_off = (uint64_t)((dr * NET_DR_BYTES + kk * 2 * NET_DR_CHUNK_BYTES) / 16);
                                    umma_bf16(d_tmem, (uint64_t)((int64_t)a_tile + a_off), b_layer + b_off, idesc96,
                                              (dr | kk) ? 1u : 0u);
                                }
                            }
                        }
                        umma_commit(bar_full0 + 8 * stage);
                        // conv1 has read the im2col buffer: the producers may build the slot's next group
                        if (layer == 0 && t == n_tiles - 1) umma_commit(bar_i2free0 + 8 * slot);
                    }
                    __syncwarp();
                    if (timing) {
                        const long long c4 = clock64();
                        tm_ready += c2 - c1; tm_empty += c3 - c2; tm_issue += c4 - c3;
                    }
                }
            }
        }
        if (timing && lane == 0 && slot == 0) {
            long long* o = reinterpret_cast<long long*>(p.debug) + (size_t)blockIdx.x * 16;
            o[0] = tm_ready; o[1] = tm_empty; o[2] = tm_issue; o[3] = clock64() - tm_start; o[4] = tm_im2col;
            o[8] = t_setup - t_entry; o[9] = tm_start - t_entry; o[10] = (long long)g_entry;
        }
    } else {
        // ================================================================== epilogue team of slot `slot`
        const uint32_t slot = (uint32_t)(warp >> 2);
        const int quarter = warp & 3;          // TMEM lane quarter this warp may read
        constexpr int M_R = IMG_ROWS - 4, M_C = IMG_COLS - 4;   // conv4 output size
        const float* bias_all = reinterpret_cast<const float*>(smem + SM_W + NET_WBLOB_MMA_BYTES);   // b1..b4
        const uint32_t bar_ready = bar_ready0 + 8 * NET_TILES * slot;
        {   // im2col of the slot's first group by this team (128 threads, 4 rows each): halves the start-up latency
            const int grp0 = blockIdx.x + (int)slot * gridDim.x;
            if (grp0 < n_groups) {
                const int tt = tid - (int)slot * 128, board0 = grp0 * NB, n_boards = min(NB, p.G - board0);
                uint8_t* im2col = smem + SM_IM2COL(slot);
#pragma unroll
                for (int q = 0; q < NET_ROWS / 128; ++q) {
                    const int row = tt + q * 128;
                    if (row < n_rows) {
                        int b, r, c;
                        net_row_coords(row, RPB, b, r, c);
                        const bool ok = b < n_boards;
                        net_im2col_row(im2col, row, IMG_ROWS, IMG_COLS, r, c, ok ? p.own[board0 + b] : 0ULL,
                                       ok ? p.opp[board0 + b] : 0ULL);
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                asm volatile("bar.sync %0, 128;" ::"r"(1 + (int)slot) : "memory");     // the team's named barrier
                if (tt < NET_PRODUCERS / 32) mbar_arrive(bar_im2col0 + 8 * slot);
            }
        }
        mbar_wait(bar_weights, 0u);
        uint32_t it = 0;
        long long te_full = 0, te_work = 0;
        for (int grp = blockIdx.x + (int)slot * gridDim.x; grp < n_groups; grp += NET_SLOTS * gridDim.x) {
            const int board0 = grp * NB;
            const int n_boards = min(NB, p.G - board0);
            for (int layer = 0; layer < 4; ++layer) {
                uint8_t* dst = smem + SM_BUF(slot, (layer & 1) ? 1 : 0);
                float bias[32];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 v = reinterpret_cast<const float4*>(bias_all + layer * 32)[i];
                    bias[4 * i] = v.x; bias[4 * i + 1] = v.y; bias[4 * i + 2] = v.z; bias[4 * i + 3] = v.w;
                }
                for (int t = 0; t < n_tiles; ++t, ++it) {
                    const uint32_t stage = slot * 2u + (it & 1u), phase = (it >> 1) & 1u;
                    const int row = t * NET_TILE_M + quarter * 32 + lane;     // TMEM lane = row within the tile
                    int b, r, c;
                    net_row_coords(row, RPB, b, r, c);
                    const bool on_board = b < n_boards && r >= 0 && c < IMG_COLS;
                    const long long e0 = timing ? clock64() : 0;
                    mbar_wait(bar_full0 + 8 * stage, phase);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const long long e1 = timing ? clock64() : 0;
                    te_full += e1 - e0;
                    const uint32_t t_lane = tmem_base + ((uint32_t)(quarter * 32) << 16) + stage * NET_STAGE_COLS;
                    uint32_t out[16];
                    if (layer == 0) {
                        uint32_t acc[32];
                        tmem_ld32(t_lane, acc);
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_empty0 + 8 * stage);
#pragma unroll
                        for (int co = 0; co < 32; co += 2) {
                            float x0, x1;
                            f32x2_unpack(f32x2_add(f32x2_pack(__uint_as_float(acc[co]), __uint_as_float(acc[co + 1])),
                                                   f32x2_pack(bias[co], bias[co + 1])), x0, x1);
                            out[co >> 1] = pack_relu_bf16x2(x0, x1);
                        }
                        if (!on_board) {
#pragma unroll
                            for (int i = 0; i < 16; ++i) out[i] = 0u;         // zero ring / pad cells = conv2's padding
                        }
                    } else {
                        // accumulator row p, block dc holds the contribution to output p - dc:
                        //   out[p] = blk(-1)[p - 1] + blk(0)[p] + blk(+1)[p + 1]; column 0 / 7 have no left / right
                        //   neighbour (zero padding for conv2, never-valid outputs for conv3 / conv4).
                        // two halves of 16 channels: the TMEM read of the second overlaps the arithmetic of the first
                        uint32_t lo0[16], mid0[16], hi0[16], lo1[16], mid1[16], hi1[16];
                        tmem_ld48_issue(t_lane, lo0, mid0, hi0);
                        tmem_ld48_wait(lo0, mid0, hi0);
                        tmem_ld48_issue(t_lane + 16, lo1, mid1, hi1);
                        // 0/1 factors instead of selects: one FFMA per neighbour term (all operands are finite)
                        const float hl = c > 0 ? 1.f : 0.f, hr = c < 7 ? 1.f : 0.f;
                        const uint64_t has_left = f32x2_pack(hl, hl), has_right = f32x2_pack(hr, hr);
#pragma unroll
                        for (int co = 0; co < 16; co += 2) {
                            const float up0 = __shfl_up_sync(0xffffffffu, __uint_as_float(lo0[co]), 1);
                            const float up1 = __shfl_up_sync(0xffffffffu, __uint_as_float(lo0[co + 1]), 1);
                            const float dn0 = __shfl_down_sync(0xffffffffu, __uint_as_float(hi0[co]), 1);
                            const float dn1 = __shfl_down_sync(0xffffffffu, __uint_as_float(hi0[co + 1]), 1);
                            uint64_t x = f32x2_add(f32x2_pack(__uint_as_float(mid0[co]), __uint_as_float(mid0[co + 1])),
                                                   f32x2_pack(bias[co], bias[co + 1]));
                            x = f32x2_fma(f32x2_pack(up0, up1), has_left, x);
                            x = f32x2_fma(f32x2_pack(dn0, dn1), has_right, x);
                            float x0, x1;
                            f32x2_unpack(x, x0, x1);
                            out[co >> 1] = pack_relu_bf16x2(x0, x1);
                        }
                        tmem_ld48_wait(lo1, mid1, hi1);
                        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_empty0 + 8 * stage);
#pragma unroll
                        for (int co = 0; co < 16; co += 2) {
                            const float up0 = __shfl_up_sync(0xffffffffu, __uint_as_float(lo1[co]), 1);
                            const float up1 = __shfl_up_sync(0xffffffffu, __uint_as_float(lo1[co + 1]), 1);
                            const float dn0 = __shfl_down_sync(0xffffffffu, __uint_as_float(hi1[co]), 1);
                            const float dn1 = __shfl_down_sync(0xffffffffu, __uint_as_float(hi1[co + 1]), 1);
                            uint64_t x = f32x2_add(f32x2_pack(__uint_as_float(mid1[co]), __uint_as_float(mid1[co + 1])),
                                                   f32x2_pack(bias[16 + co], bias[17 + co]));
                            x = f32x2_fma(f32x2_pack(up0, up1), has_left, x);
                            x = f32x2_fma(f32x2_pack(dn0, dn1), has_right, x);
                            float x0, x1;
                            f32x2_unpack(x, x0, x1);
                            out[8 + (co >> 1)] = pack_relu_bf16x2(x0, x1);
                        }
                    }
                    if (layer < 3) {
                        uint8_t* d = dst + (NET_MARGIN + row) * 16;
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(d + ch * NET_CHUNK_BYTES) =
                                make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        __syncwarp();
                        if (lane == 0) mbar_arrive(bar_ready + 8 * t);
                        if (p.debug && p.debug_layer == layer + 1)
                            for (int ch = 0; ch < 4; ++ch)
                                *reinterpret_cast<uint4*>(p.debug + ((size_t)grp * NET_DEBUG_ROWS + row) * 32 + ch * 8) =
                                    make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    } else if (on_board && r >= 2 && r < 2 + M_R && c >= 2 && c < 2 + M_C) {
                        // conv4: the outputs are the cells at an inset of 2; features[board][pos*32 + co]
                        const int pos = (r - 2) * M_C + (c - 2);
                        if (p.features_packed) {
                            // K = pos*32 + 8*ch: chunk pos / 2 (64 wide, 16 KB per row tile), plane (pos % 2)*4 + ch
                            static_assert((M_R * M_C) % 2 == 0, "64-wide K chunks");
                            const int g = board0 + b;
                            uint8_t* f = reinterpret_cast<uint8_t*>(p.features) +
                                         ((size_t)(g >> 7) * (M_R * M_C / 2) + (size_t)(pos >> 1)) * 16384 +
                                         (size_t)((pos & 1) * 4) * 2048 + (size_t)(g & 127) * 16;
#pragma unroll
                            for (int ch = 0; ch < 4; ++ch)
                                *reinterpret_cast<uint4*>(f + ch * 2048) =
                                    make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                        } else {
                            uint4* f = reinterpret_cast<uint4*>(p.features + ((size_t)(board0 + b) * (M_R * M_C) + pos) * 32);
#pragma unroll
                            for (int ch = 0; ch < 4; ++ch)
                                f[ch] = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                        }
                    }
                    if (timing) te_work += clock64() - e1;
                }
            }
        }
        if (timing && lane == 0 && quarter == 0) {
            long long* o = reinterpret_cast<long long*>(p.debug) + (size_t)blockIdx.x * 16;
            o[5 + slot] = te_full;
            if (slot == 0) o[7] = te_work;
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(NET_TMEM_COLS));
    if (timing && tid == 0) {
        unsigned long long g_end;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(g_end));
        long long* o = reinterpret_cast<long long*>(p.debug) + (size_t)blockIdx.x * 16;
        o[11] = (long long)g_end; o[12] = clock64() - t_entry;
    }
}

// ------------------------------------------------------------------------------------------------ heads epilogue

// One warp per position: probs = softmax(logits[0 .. A)) in float32, value = tanh(logits[A]).
// Fuses log_softmax + exp (PolicyValueNetwork.predict, base.py:350-355) and tanh (othello.py:382) and writes straight
// into the tree's probs / value buffers.
__global__ void __launch_bounds__(128) k_net_heads(const __nv_bfloat16* __restrict__ logits, int ld, int A, int64_t G,
                                                   float* __restrict__ probs, float* __restrict__ value) {
    az_pdl_trigger();   // the tree step that follows (az_launch_pdl) waits for this grid's completion before reading
    const int64_t g = (int64_t)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= G) return;
    const __nv_bfloat16* row = logits + g * ld;
    float x[3], pr[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) x[i] = (lane + 32 * i) < A ? __bfloat162float(row[lane + 32 * i]) : 0.f;
    az_heads_softmax(x, lane, A, pr);
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        const int a = lane + 32 * i;
        if (a < A) probs[g * A + a] = pr[i];
    }
    if (lane == 0) value[g] = az_heads_value(row[A]);
}

// ------------------------------------------------------------------------------------------------ heads GEMM

// logits = h2 x Wh^T + bh, probs = softmax(policy logits), value = tanh(value logit) in ONE kernel: the two output
// layers fc_probs / fc_value (othello.py:361-362, 379-382; connect4.py:389-390, 408-411) as a tcgen05 GEMM whose
// epilogue is PolicyValueNetwork.predict's exp(log_softmax) / tanh (base.py:350-355).  Replaces a library GEMM + the
// k_net_heads launch; the logits stay fp32 (they never exist in memory).
//
// One CTA per 128 positions.  A = h2 tile, K-major no-swizzle [K/8][128][8] bf16, filled by 8 warps (lane = 8 rows x
// 4 K-chunks: 64-byte row segments from global, 128-byte contiguous shared-memory stores) in four K quarters, each
// with its mbarrier, so that the MMAs of a quarter start while the next one is still loading; B = the packed head
// weights [K/8][NH][8] by one bulk async copy.  K/16 MMAs of shape 128 x NH x 16 into NH TMEM columns.
#define HG_ROWS 128
#define HG_LOAD_WARPS 8
#define HG_THREADS (32 * (HG_LOAD_WARPS + 1))
#define HG_MAX_K 512
#define HG_MAX_NH 96
#define HG_QUARTERS 4

struct HeadGemmParams {
    const __nv_bfloat16* h2;   // [G][K]
    const uint8_t* wblob;      // [K/8][NH][8] bf16, then float bias[NH]
    float* probs;              // [G][A]
    float* value;              // [G]
    int64_t G;
    int K, NH, A;
};

template <int NH>
__global__ void __launch_bounds__(HG_THREADS, 1) k_net_headgemm(HeadGemmParams p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int K = p.K, n_kc = K >> 3;                       // 8-element K chunks
    const uint32_t a_plane = HG_ROWS * 16, b_plane = NH * 16;
    uint8_t* a_smem = smem;                                 // n_kc planes of 2 KB
    // the A region doubles as the probs staging area of the epilogue ([128][A] floats): sized for the larger of the two
    const uint32_t a_bytes = max((uint32_t)n_kc * a_plane, (uint32_t)(HG_ROWS * NH * 4));
    uint8_t* b_smem = smem + a_bytes;                       // n_kc planes of NH*16 B, then float bias[NH]
    const uint32_t w_bytes = (uint32_t)n_kc * b_plane + NH * 4u;
    uint8_t* tail = b_smem + ((w_bytes + 127u) & ~127u);
    const uint32_t bars = smem_u32(tail);                   // quarter[4], weights, full
    const uint32_t bar_w = bars + 8 * HG_QUARTERS, bar_full = bar_w + 8;
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tail + 64);
    const int64_t row0 = (int64_t)blockIdx.x * HG_ROWS;

    if (tid == 0) {
        for (int q = 0; q < HG_QUARTERS; ++q) mbar_init(bars + 8 * q, HG_LOAD_WARPS * 32);
        mbar_init(bar_w, 1);
        mbar_init(bar_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}"
                     ::"r"(bar_w), "r"(w_bytes) : "memory");
        asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                     ::"r"(smem_u32(b_smem)), "l"(p.wblob), "r"(w_bytes), "r"(bar_w) : "memory");
    }
    if (warp == HG_LOAD_WARPS) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(128));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_base = *tmem_ptr_smem;

    if (warp < HG_LOAD_WARPS) {
        // ---- A tile: blocks of 8 rows x 4 K-chunks, K-chunk-block major so that the quarters complete in order
        const int r8 = lane & 7, k4 = lane >> 3;
        const int kc_blocks = n_kc >> 2, per_q = (kc_blocks + HG_QUARTERS - 1) / HG_QUARTERS;
        for (int q = 0; q < HG_QUARTERS; ++q) {
            const int kb0 = q * per_q, kb1 = min(kc_blocks, kb0 + per_q);
            const int n_blocks = (kb1 - kb0) * (HG_ROWS / 8);
            // 8 independent 16-byte loads in flight per lane before the first store (latency-bound otherwise)
            for (int b0 = warp; b0 < n_blocks; b0 += 8 * HG_LOAD_WARPS) {
                uint4 v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int b = b0 + u * HG_LOAD_WARPS;
                    const int kc = (kb0 + b / (HG_ROWS / 8)) * 4 + k4, row = (b % (HG_ROWS / 8)) * 8 + r8;
                    v[u] = make_uint4(0u, 0u, 0u, 0u);
                    if (b < n_blocks && row0 + row < p.G)
                        v[u] = *reinterpret_cast<const uint4*>(p.h2 + (row0 + row) * K + kc * 8);
                }
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int b = b0 + u * HG_LOAD_WARPS;
                    const int kc = (kb0 + b / (HG_ROWS / 8)) * 4 + k4, row = (b % (HG_ROWS / 8)) * 8 + r8;
                    if (b < n_blocks) *reinterpret_cast<uint4*>(a_smem + (size_t)kc * a_plane + row * 16) = v[u];
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            mbar_arrive(bars + 8 * q);
        }
    } else {
        // ---- MMA issue
        const uint32_t idesc = umma_idesc_bf16_m128(NH);
        const int kc_blocks = n_kc >> 2, per_q = (kc_blocks + HG_QUARTERS - 1) / HG_QUARTERS;
        mbar_wait(bar_w, 0u);
        for (int q = 0; q < HG_QUARTERS; ++q) {
            mbar_wait(bars + 8 * q, 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int ks0 = q * per_q * 2, ks1 = min(n_kc >> 1, (q + 1) * per_q * 2);   // K = 16 steps of this quarter
            if (elect_one()) {
                for (int ks = ks0; ks < ks1; ++ks) {
                    const uint64_t ad = umma_desc(smem_u32(a_smem) + (uint32_t)ks * 2u * a_plane, a_plane, 128);
                    const uint64_t bd = umma_desc(smem_u32(b_smem) + (uint32_t)ks * 2u * b_plane, b_plane, 128);
                    umma_bf16(tmem_base, ad, bd, idesc, ks > 0 ? 1u : 0u);
                }
                if (q == HG_QUARTERS - 1) umma_commit(bar_full);
            }
            __syncwarp();
        }
    }

    // ---- epilogue: warps 0-3, thread = row (TMEM lane); logits + bias -> softmax / tanh -> staged -> coalesced stores
    if (warp < 4) {
        const float* bias = reinterpret_cast<const float*>(b_smem + (size_t)n_kc * b_plane);
        mbar_wait(bar_full, 0u);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        uint32_t acc[NH];
#pragma unroll
        for (int c = 0; c < NH / 16; ++c) {
            uint32_t (&blk)[16] = *reinterpret_cast<uint32_t (*)[16]>(&acc[c * 16]);
            tmem_ld16_issue(tmem_base + ((uint32_t)(warp * 32) << 16) + c * 16, blk);
        }
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        const int A = p.A;
        float mx = -INFINITY;
#pragma unroll
        for (int a = 0; a < NH; ++a) {
            const float x = __uint_as_float(acc[a]) + bias[a];
            acc[a] = __float_as_uint(x);
            if (a < A) mx = fmaxf(mx, x);
        }
        float sum = 0.f;
#pragma unroll
        for (int a = 0; a < NH; ++a)
            if (a < A) {
                const float e = __expf(__uint_as_float(acc[a]) - mx);
                acc[a] = __float_as_uint(e);
                sum += e;
            }
        const float inv = 1.f / sum;
        // the A operand is dead once the accumulator is complete: stage probs [128][A] there for coalesced stores
        float* stage = reinterpret_cast<float*>(a_smem);
        const int row = warp * 32 + lane;
        float v = 0.f;
#pragma unroll
        for (int a = 0; a < NH; ++a) {
            if (a < A) stage[row * A + a] = __uint_as_float(acc[a]) * inv;
            else if (a == A) v = __uint_as_float(acc[a]);
        }
        if (row0 + row < p.G) p.value[row0 + row] = tanhf(v);
        asm volatile("bar.sync 1, 128;" ::: "memory");
        const int64_t n_valid = min((int64_t)HG_ROWS, p.G - row0) * A;
        float* out = p.probs + row0 * A;
        for (int i = tid; i < n_valid; i += 128) out[i] = stage[i];
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == HG_LOAD_WARPS)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128));
}

// ------------------------------------------------------------------------------------------------ C ABI

// k_net_trunk's blob, then conv2..conv4 once more in the order k_net_trunk_rows stages them by one bulk copy
// ([layer][kernel column][ci chunk][(kernel row, co)][8]: trunk_rows.cuh)
static_assert(NET_WBLOB_BYTES == AZ_TRUNK_ROWS_BLOB_OFFSET, "trunk_rows.cuh knows where its section of the blob starts");
extern "C" int az_net_wblob_bytes(void) { return NET_WBLOB_BYTES + AZ_TRUNK_ROWS_BLOB_BYTES; }

// Boards per group / tiles per group for a launch over G positions on n_sms persistent CTAs: a group is 2, 4 or 6
// tiles of 128 rows (rows per board = (n + 1) * 8, plus one closing ring row); the shape that needs the fewest
// tile-steps on the busiest CTA wins (8x8 on 148 SMs: 4096 positions -> 3 rounds of 10 boards / 6 tiles,
// 2048 positions -> 2 rounds of 7 boards / 4 tiles).
static void net_group_shape(int rows, int64_t G, int n_sms, int* boards_per_group, int* n_tiles) {
    const int rpb = (rows + 1) * NET_PITCH;
    int64_t best_cost = -1;
    for (int tiles = 2; tiles <= NET_TILES; tiles += 2) {
        const int nb = (tiles * NET_TILE_M - NET_PITCH) / rpb;
        const int64_t groups = (G + nb - 1) / nb;
        // a CTA runs its groups two at a time (NET_SLOTS pipelines); + 1 tile-step per group for the ramp
        const int64_t per_cta = (groups + n_sms - 1) / n_sms;
        const int64_t cost = ((per_cta + NET_SLOTS - 1) / NET_SLOTS) * (tiles + 1);
        if (best_cost < 0 || cost <= best_cost) { best_cost = cost; *boards_per_group = nb; *n_tiles = tiles; }
    }
}

static bool net_valid_image(int rows, int cols) {
    return (rows == 8 && cols == 8) || (rows == 6 && cols == 6) || (rows == 7 && cols == 6);
}

// Boards per group az_net_trunk uses for (rows, G) on the current device (the layout of its debug dump depends on it).
extern "C" int az_net_trunk_boards_per_group(int rows, int64_t G) {
    int dev = 0, n_sms = 0, nb = 0, tiles = 0;
    if (rows < 6 || rows > 8 || G <= 0) return AZ_ERR_ARG;
    if (cudaGetDevice(&dev) != cudaSuccess ||
        cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess)
        return AZ_ERR_CUDA;
    if (rows == 8 && G >= AZ_TRUNK_ROWS_MIN_G && az_trunk_rows_enabled()) return AZ_TRUNK_ROWS_BOARDS_PER_GROUP;    // k_net_trunk_rows / rows2
    if (rows == 7 && az_trunk_rows_enabled() && az_trunk_rows2_enabled() && (G + 17) / 18 > n_sms) return 18;       // k_net_trunk_rows2<7, 6>
    net_group_shape(rows, G, n_sms, &nb, &tiles);
    return nb;
}

extern "C" int az_net_trunk(const uint64_t* own, const uint64_t* opp, int rows, int cols, int64_t G, const void* wblob,
                            void* features, int features_packed, void* debug, int debug_layer, void* stream) {
    if (!net_valid_image(rows, cols) || G < 0 || (G > 0 && (!own || !opp || !wblob || !features))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    static int n_sms = 0;
    static bool attr_set = false;
    if (!attr_set) {
        int dev = 0;
        AZ_CUDA_TRY(cudaGetDevice(&dev));
        AZ_CUDA_TRY(cudaDeviceGetAttribute(&n_sms, cudaDevAttrMultiProcessorCount, dev));
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL));
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk<6, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL));
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk<7, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, SM_TOTAL));
        attr_set = true;
    }
    static const int late_trigger = getenv("AZ_B200_TRUNK_LATE_TRIGGER") ? atoi(getenv("AZ_B200_TRUNK_LATE_TRIGGER")) : 0;
    // 8 x 8 boards in large batches: k_net_trunk_rows (trunk_rows_kernels.cu) unless the caller (bit 1) or the
    // environment asks for this file's kernel; the cycle-accounting mode (debug_layer 100) belongs to this file's kernel
    if (rows == 8 && cols == 8 && G >= AZ_TRUNK_ROWS_MIN_G && !(features_packed & 2) && az_trunk_rows_enabled() &&
        !(debug && debug_layer == 100))
        return (az_trunk_rows2_enabled() && !(features_packed & 4) && (G + AZ_TRUNK_ROWS_BOARDS_PER_GROUP - 1) / AZ_TRUNK_ROWS_BOARDS_PER_GROUP > n_sms &&
                !(debug && debug_layer >= 200))
                   ? az_trunk_rows2_launch(own, opp, 8, 8, G, wblob, features, features_packed & 1, debug, debug_layer, n_sms, late_trigger, stream)
                   : az_trunk_rows_launch(own, opp, G, wblob, features, features_packed & 1, debug, debug_layer, n_sms, late_trigger, stream);
    // Connect4Net's 7 x 6 image: the same two-CTAs-per-SM kernel once there are more groups of 18 boards than SMs
    if (rows == 7 && cols == 6 && !(features_packed & 6) && az_trunk_rows_enabled() && az_trunk_rows2_enabled() &&
        (G + az_trunk_rows2_boards_per_group(7) - 1) / az_trunk_rows2_boards_per_group(7) > n_sms && !(debug && debug_layer >= 100))
        return az_trunk_rows2_launch(own, opp, 7, 6, G, wblob, features, features_packed & 1, debug, debug_layer, n_sms, late_trigger, stream);
    NetTrunkParams p;
    p.own = own; p.opp = opp; p.wblob = (const uint8_t*)wblob;
    p.features = (__nv_bfloat16*)features; p.debug = (__nv_bfloat16*)debug;
    p.G = (int)G; p.rows = rows; p.cols = cols;
    net_group_shape(rows, G, n_sms, &p.boards_per_group, &p.n_tiles);
    p.debug_layer = debug ? debug_layer : 0;
    p.features_packed = features_packed & 1;
    p.late_trigger = late_trigger;
    const int n_groups = (int)((G + p.boards_per_group - 1) / p.boards_per_group);
    const int grid = n_groups < n_sms ? n_groups : n_sms;   // persistent: one CTA per SM
    cudaError_t e;
    if (rows == 8) e = az_launch_pdl(k_net_trunk<8, 8>, dim3(grid), dim3(NET_THREADS), SM_TOTAL, (cudaStream_t)stream, p);
    else if (rows == 6) e = az_launch_pdl(k_net_trunk<6, 6>, dim3(grid), dim3(NET_THREADS), SM_TOTAL, (cudaStream_t)stream, p);
    else e = az_launch_pdl(k_net_trunk<7, 6>, dim3(grid), dim3(NET_THREADS), SM_TOTAL, (cudaStream_t)stream, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_trunk");
    AZ_CHECK_LAUNCH("k_net_trunk");
    return AZ_OK;
}

// Connect4Net VIEWS the row-major (height, width) grid as a (width, height) image (connect4.py:399): image cell
// (i, j) is flat index i*height + j of the grid, i.e. grid cell (flat / width, flat % width).  One thread per board
// moves the canvas bits (grid r*8 + c) to image-canvas bits (i*8 + j) for az_net_trunk.
__global__ void __launch_bounds__(256) k_connect4_image_bits(const uint64_t* __restrict__ own,
                                                             const uint64_t* __restrict__ opp, int width, int height,
                                                             int64_t G, uint64_t* __restrict__ own_img,
                                                             uint64_t* __restrict__ opp_img) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t o = own[g], p = opp[g];
        uint64_t oi = 0ULL, pi = 0ULL;
        for (int flat = 0; flat < width * height; ++flat) {
            const int src = (flat / width) * 8 + (flat % width), dst = (flat / height) * 8 + (flat % height);
            oi |= ((o >> src) & 1ULL) << dst;
            pi |= ((p >> src) & 1ULL) << dst;
        }
        own_img[g] = oi;
        opp_img[g] = pi;
    }
}

extern "C" int az_connect4_image_bits(const uint64_t* own, const uint64_t* opp, int width, int height, int64_t G,
                                      uint64_t* own_img, uint64_t* opp_img, void* stream) {
    if (width < 4 || width > 8 || height < 4 || height > 8 || G < 0 || (G > 0 && (!own || !opp || !own_img || !opp_img)))
        return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_connect4_image_bits<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, width, height, G, own_img, opp_img);
    AZ_CHECK_LAUNCH("k_connect4_image_bits");
    return AZ_OK;
}

extern "C" int az_net_heads(const void* logits, int ld, int action_size, int64_t G, float* probs, float* value,
                            void* stream) {
    const int A = action_size;
    if (A < 1 || A > 96 || G < 0 || ld < A + 1 || (G > 0 && (!logits || !probs || !value))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_net_heads<<<(unsigned)((G + 3) / 4), 128, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)logits, ld, A, G, probs,
                                                                       value);
    AZ_CHECK_LAUNCH("k_net_heads");
    return AZ_OK;
}

static size_t hg_smem_bytes(int K, int NH) {
    const size_t w = (size_t)(K / 8) * NH * 16 + NH * 4;
    const size_t a = (size_t)(K / 8) * HG_ROWS * 16, stage = (size_t)HG_ROWS * NH * 4;
    return (a > stage ? a : stage) + ((w + 127) & ~(size_t)127) + 128;
}

extern "C" int64_t az_net_headgemm_blob_bytes(int K, int NH) {
    if (K < 16 || K > HG_MAX_K || (K % 32) || NH < 16 || NH > HG_MAX_NH || (NH % 16)) return AZ_ERR_ARG;
    return (int64_t)(K / 8) * NH * 16 + NH * 4;
}

extern "C" int az_net_headgemm(const void* h2, int K, const void* wblob, int NH, int action_size, int64_t G,
                               float* probs, float* value, void* stream) {
    if (az_net_headgemm_blob_bytes(K, NH) < 0 || action_size < 1 || action_size + 1 > NH || G < 0 ||
        (G > 0 && (!h2 || !wblob || !probs || !value)))
        return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    HeadGemmParams p;
    p.h2 = (const __nv_bfloat16*)h2; p.wblob = (const uint8_t*)wblob; p.probs = probs; p.value = value;
    p.G = G; p.K = K; p.NH = NH; p.A = action_size;
    const size_t bytes = hg_smem_bytes(K, NH);
    const unsigned grid = (unsigned)((G + HG_ROWS - 1) / HG_ROWS);
    static size_t attr_bytes[HG_MAX_NH / 16 + 1] = {0};     // opt-in shared memory already granted per instantiation
#define HG_CASE(N)                                                                                              \
    case N:                                                                                                     \
        if (attr_bytes[N / 16] < bytes) {                                                                       \
            AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_headgemm<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes)); \
            attr_bytes[N / 16] = bytes;                                                                         \
        }                                                                                                       \
        k_net_headgemm<N><<<grid, HG_THREADS, bytes, (cudaStream_t)stream>>>(p);                                \
        break;
    switch (NH) {
        HG_CASE(16) HG_CASE(32) HG_CASE(48) HG_CASE(64) HG_CASE(80) HG_CASE(96)
        default: return AZ_ERR_ARG;
    }
    AZ_CHECK_LAUNCH("k_net_headgemm");
    return AZ_OK;
}
