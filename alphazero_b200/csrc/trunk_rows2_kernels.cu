// Convolution trunk of OthelloNet(n = 8) for batches of 1024 positions and more, TWO CTAs PER SM.
// (alphazero/games/othello.py:366-374: conv1..conv4 + BatchNorm + ReLU, pads 1, 1, 0, 0; same weight blob, same outputs
// as k_net_trunk<8, 8> in net_kernels.cu and k_net_trunk_rows in trunk_rows_kernels.cu.)
//
// The formulation is k_net_trunk_rows' (read its header first): a tile = one board ROW of 14 boards (lane = board * 9 +
// column), kernel column = a +-16-byte shift of the A operand, kernel row = the N dimension, ONE TMEM ACCUMULATOR PER
// OUTPUT TILE (O[r] at column 32 * (7 - r)) into which the tensor core scatters the kernel-row blocks of the input tiles,
// activations converted in place.  What limits that kernel is not a unit of the SM but the round trips of ONE group's
// dependency chain (issue -> pipe -> epilogue -> next layer; profiles/r2_trunk_experiments.txt, item 7): tensor pipe
// 29 % busy, issue slots 35 %.  So this kernel makes a group small enough for TWO CTAs to share an SM, each with its
// own group, and lets the hardware interleave them:
//   * TMEM: one region of 256 columns instead of two -- a layer's accumulators REPLACE the previous layer's in place.
//     The job that first touches output tile x of layer L + 1 (its N = 32 initialising MMA) therefore also waits for
//     layer L's item of tile x to have read its accumulator -- which is the "tile ready" barrier of the NEXT job's
//     input: the same number of polls, one tile earlier.
//   * shared memory: ONE activation buffer (66.5 KB; a CTA runs its groups one after the other); conv1's im2col operand
//     lives in planes 2 - 3 of that buffer (conv1's MMAs do not shift rows, and the item that converts tile r in place
//     waits for the MMA that read tile r's im2col rows); the weights of ONE layer at a time in two 18 KB buffers (conv2
//     and conv3 copied at set-up, conv4 into conv2's buffer once conv2's MMAs have completed): 103.4 KB per CTA.
//   * threads: two issuing warps that own the output tiles 0..3 / 4..7 and walk from the border outwards, as in
//     k_net_trunk_rows, + two epilogue teams (one per issuing warp's half of the tiles) = 320, 96 registers; the im2col
//     operand is built by all of them at the start of a group.  (One issuing warp per CTA -- no tile a job of two warps,
//     154 instead of 200 MMAs per group -- was measured too: 17.3 against 16.7 us, the single thread's chain is longer.)
// A CTA with more than one group runs them back to back with a block barrier in between (the other CTA of the SM fills
// the bubble).  az_net_trunk takes this kernel when there are more groups than SMs, k_net_trunk_rows below that.
// Templated on the image: T rows x C columns, a tile = image row r of floor(128 / (C + 1)) boards -- 8 x 8 (Othello, 14
// boards per group) and 7 x 6 (Connect4Net's (width, height) view of the board, connect4.py:397-403: 18 boards per
// group).  The shared-memory and TMEM layouts are those of 8 tiles for both; the 7 x 6 image leaves the last tile unused.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "trunk_rows.cuh"

#define R2_TILES 8                // tiles per layer in the layout = the largest number of image rows
#define R2_ROWS (R2_TILES * 128)
#define R2_MARGIN 8               // zero rows before / after the buffer (the +-1 row shifts of the first / last tile)
#define R2_PLANE ((R2_ROWS + 2 * R2_MARGIN) * 16)      // one 8-channel chunk plane: 16 640 B (= LBO of the A operand)
#define R2_BUF (4 * R2_PLANE)     // activation buffer: 66 560 B
#define R2_TEAMS 2
#define R2_EPI_THREADS (128 * R2_TEAMS)
#define R2_MMA_WARP (R2_EPI_THREADS / 32)
#define R2_THREADS (R2_EPI_THREADS + 64)
#define R2_HJOBS 18               // MMA tile jobs per group and issuing warp, at most: 4 + 5 + 5 + 4
#define R2_CJOBS 14               // of which conv2..conv4
#define R2_ITEMS 26               // epilogue items per group = output tiles that are used, at most: 8 + 8 + 6 + 4
#define R2_LAYER_BYTES 18432      // one layer's weights: [kw 3][ci chunk 4][(kh, co) 96][8] bf16
#define R2_W_DC_BYTES (4 * 96 * 16)                    // one kernel column of them: 6 144 B
// shared memory map
#define R2_SM_BUF 0
#define R2_SM_WB(i) (R2_BUF + (i) * R2_LAYER_BYTES)    // buffer 0: conv2, later conv4; buffer 1: conv3
#define R2_SM_W1 (R2_SM_WB(2))                         // conv1: [k chunk 2][co 32][8] bf16
#define R2_SM_BIAS (R2_SM_W1 + 1024)                   // float [4][32]
#define R2_SM_BARS (R2_SM_BIAS + 512)
#define R2B_JOB 0                                      // [2][18] the MMAs of tile job j of issuing warp h have completed
#define R2B_READY 36                                   // [8] output tile r of the layer is in the buffer and its accumulator read
#define R2B_W1 44                                      // conv1's weights and the biases have landed
#define R2B_WB 45                                      // [2] a layer's weights have landed in buffer i
#define R2B_COUNT 47
#define R2_SM_TMEM_PTR (R2_SM_BARS + 8 * 48)
#define R2_SM_TOTAL (R2_SM_TMEM_PTR + 16)
static_assert(2 * (R2_SM_TOTAL + 1024) <= 233472, "two CTAs per SM");
#define R2_WAIT_LIMIT (4000000000LL)

struct TrunkRows2Params {
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
__device__ __forceinline__ void r2_wait(uint32_t bar, uint32_t parity) {
    long long t0 = 0;
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        const long long t = clock64();
        if (t0 == 0) t0 = t;
        else if (t - t0 > R2_WAIT_LIMIT) __trap();
    }
}
__device__ __forceinline__ void r2_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void r2_bulk_load(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
    asm volatile("{\n\t.reg .b64 st;\n\tmbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1;\n\t}" ::"r"(bar), "r"(bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(dst), "l"(src), "r"(bytes), "r"(bar) : "memory");
}

// tcgen05.mma with K-major no-swizzle descriptors given by their low words (trunk_rows_kernels.cu: tr_mma)
__device__ __forceinline__ void r2_mma(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
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
__device__ __forceinline__ uint32_t r2_desc_lo(uint32_t smem_addr, uint32_t lbo) {
    return ((smem_addr >> 4) & 0x3FFFu) | ((lbo >> 4) << 16);
}

// Per layer L = 0..3 (conv1..conv4) of a T-row image: first MMA tile, number of MMA tiles; first / last output tile that
// is used; per issuing warp h (output tiles 0..3 / 4..T-1): the output tiles it produces, the input tiles that contribute
// to them = its jobs of the layer, their number, its first job of the layer among its jobs of a group.
__host__ __device__ constexpr int r2_min(int a, int b) { return a < b ? a : b; }
__host__ __device__ constexpr int r2_max(int a, int b) { return a > b ? a : b; }
__host__ __device__ constexpr int r2_mlo(int L) { return L == 3 ? 1 : 0; }
__host__ __device__ constexpr int r2_mcnt(int L, int T) { return L == 3 ? T - 2 : T; }
__host__ __device__ constexpr int r2_olo(int L) { return L < 2 ? 0 : L - 1; }                  // conv3: 1, conv4: 2
__host__ __device__ constexpr int r2_ohi(int L, int T) { return L < 2 ? T - 1 : T - L; }       // conv3: T - 2, conv4: T - 3
__host__ __device__ constexpr int r2_holo(int h, int L) { return r2_max(r2_olo(L), 4 * h); }
__host__ __device__ constexpr int r2_hohi(int h, int L, int T) { return r2_min(r2_ohi(L, T), 4 * h + 3); }
__host__ __device__ constexpr int r2_tlo(int h, int L) { return L == 0 ? 4 * h : r2_max(r2_mlo(L), r2_holo(h, L) - 1); }
__host__ __device__ constexpr int r2_thi(int h, int L, int T) {
    return L == 0 ? r2_min(4 * h + 3, T - 1) : r2_min(r2_mlo(L) + r2_mcnt(L, T) - 1, r2_hohi(h, L, T) + 1);
}
__host__ __device__ constexpr int r2_cnt(int h, int L, int T) { return r2_thi(h, L, T) - r2_tlo(h, L) + 1; }
__host__ __device__ constexpr int r2_jloc(int h, int L, int T) {
    int j = 0;
    for (int l = 0; l < L; ++l) j += r2_cnt(h, l, T);
    return j;
}
// layers that write tile r into the buffer per group (conv1, conv2 all of them, conv3 tiles 1..T-2): its "ready" phases
__host__ __device__ constexpr uint32_t r2_writes(int r, int T) { return (r >= 1 && r <= T - 2) ? 3u : 2u; }

// What does not depend on the launch, computed by the compiler and read from constant memory (every access is uniform
// across its warp): the items of the epilogue teams and the jobs of the issuing warps.
struct R2Tables {
    uint32_t items[R2_ITEMS + 2];            // layer | tile << 4 | job barrier << 8 | second job barrier << 14 | TMEM column << 20; 0xFFFFFFFF ends the list
    uint32_t jobs[2][R2_CJOBS][4];           // per issuing warp and job of conv2..conv4, see r2_make_tables
    int cnt[2][4];                           // jobs of issuing warp h in layer L
};
constexpr R2Tables r2_make_tables(int T) {
    R2Tables t{};
    for (int h = 0; h < 2; ++h)
        for (int L = 0; L < 4; ++L) t.cnt[h][L] = r2_cnt(h, L, T);
    // Items: layer L, output tile r, the job whose completion completes the tile's accumulator (of the warp that owns it,
    // its last input tile that contributes), the accumulator's TMEM column.  The item overwrites INPUT tile r in place:
    // the tiles next to the border (3, 4) are read by the other warp too, whose job on that tile is the item's second
    // barrier.  Jobs and items go from the border outwards -- tiles 3, 4, 2, 5, 1, 6, 0, 7 as far as the layer produces
    // them --, so that what the next layer needs first is complete first; tiles 0..3 are team 0's, the others team 1's.
    int n = 0;
    for (int L = 0; L < 4; ++L)
        for (int pos = 0; pos < 8; ++pos) {
            const int r = (pos & 1) ? 4 + (pos >> 1) : 3 - (pos >> 1);
            if (r < r2_olo(L) || r > r2_ohi(L, T)) continue;
            const int h = r >> 2;
            int j = 0;                                                  // the step of warp h that completes the accumulator
            if (L == 0) j = h ? r - 4 : 3 - r;
            else j = h ? r2_min(r + 1, r2_thi(1, L, T)) - r2_tlo(1, L) : r2_thi(0, L, T) - r2_max(r - 1, r2_tlo(0, L));
            const int job = h * R2_HJOBS + r2_jloc(h, L, T) + j;
            int job2 = job;
            if (L > 0 && (r == 3 || r == 4)) job2 = (1 - h) * R2_HJOBS + r2_jloc(1 - h, L, T);    // tile 3 / 4 is the other warp's first job
            t.items[n++] = (uint32_t)L | ((uint32_t)r << 4) | ((uint32_t)job << 8) | ((uint32_t)job2 << 14) | ((uint32_t)(32 * (7 - r)) << 20);
        }
    for (; n < R2_ITEMS + 2; ++n) t.items[n] = 0xFFFFFFFFu;
    // Jobs: job jj of issuing warp h in conv2..conv4 (its step j of layer L).  Warp 0 goes down from the border (tiles 4, 3,
    // ..), warp 1 up (3, 4, ..).
    for (int h = 0; h < 2; ++h) {
        const int c1 = t.cnt[h][1], c2 = t.cnt[h][2], c3 = t.cnt[h][3];
        for (int jj = 0; jj < c1 + c2 + c3; ++jj) {
            const int L = jj < c1 ? 1 : (jj < c1 + c2 ? 2 : 3), j = jj - (L == 1 ? 0 : (L == 2 ? c1 : c1 + c2));
            const int tlo = r2_tlo(h, L), thi = r2_thi(h, L, T), olo = r2_holo(h, L), ohi = r2_hohi(h, L, T);
            const int r = h ? tlo + j : thi - j;
            // kernel rows whose output tile r - kh + 1 this warp produces: kh = klo .. khi, 32 accumulator columns each,
            // output tiles descending = columns ascending from that of tile r - klo + 1
            const int klo = r2_max(0, r + 1 - ohi), khi = r2_min(2, r + 1 - olo), nblk = khi - klo + 1;
            const uint32_t d_col = (uint32_t)(32 * (6 - r + klo));
            // the first contribution to an output tile must not accumulate: all of the warp's first job of the layer; of
            // every other, the block of the output tile it reaches first -- tile r - 1 (kh = 2, the last block) going
            // down, tile r + 1 (kh = 0, the first block) going up -- by an N = 32 MMA of its own before the rest.  That
            // tile's accumulator is the previous layer's until its item has read it: the job waits for tile x to be
            // "ready" too.
            const bool first = j == 0, fresh = h ? klo == 0 : khi == 2, init = !first && fresh;
            uint32_t z = d_col | ((uint32_t)nblk << 9);
            if (!init) z |= (0u << 11) | ((uint32_t)nblk << 13) | ((first ? 0u : 1u) << 15);
            else {
                const uint32_t ipos = h ? 0u : (uint32_t)(nblk - 1), rpos = h ? 1u : 0u;
                z |= (ipos << 11) | (1u << 13) | (0u << 15);
                if (nblk > 1) z |= (1u << 16) | (rpos << 17) | ((uint32_t)(nblk - 1) << 19);
            }
            const int x = fresh ? (h ? r + 1 : r - 1) : 15;
            t.jobs[h][jj][0] = (uint32_t)(R2_MARGIN + r * 128 - 1);    // A: the tile's rows shifted by -1 column, in 16-byte rows
            t.jobs[h][jj][1] = (uint32_t)((R2_SM_WB(L == 2 ? 1 : 0) + klo * 32 * 16) >> 4);      // B: the layer's buffer, kernel rows from klo
            t.jobs[h][jj][2] = z;
            t.jobs[h][jj][3] = (uint32_t)r | ((uint32_t)(L - 1) << 5) | ((uint32_t)x << 7);
        }
    }
    return t;
}
// The schedule the tables encode, checked by the compiler: per layer and output tile, the owning warp's jobs in order --
// (1) every (input tile, kernel row) that contributes appears exactly once, on the right accumulator columns, (2) the
// first MMA that touches the accumulator does not accumulate and every later one does, (3) a job that initialises an
// accumulator waits for the tile whose previous-layer accumulator it replaces (its own input tile or tile x), (4) the
// tile's item waits for the last of those jobs, and -- tiles 3 / 4 -- for the other warp's job that reads the tile.
constexpr bool r2_check_tables(int T) {
    const R2Tables t = r2_make_tables(T);
    int n_items = 0;
    for (int L = 1; L < 4; ++L) {
        const int mlo = r2_mlo(L), mhi = mlo + r2_mcnt(L, T) - 1;
        for (int o = r2_olo(L); o <= r2_ohi(L, T); ++o) {
            const int h = o >> 2, off = (L > 1 ? t.cnt[h][1] : 0) + (L > 2 ? t.cnt[h][2] : 0);
            int seen = 0, last = -1;
            bool inited = false;
            for (int j = 0; j < t.cnt[h][L]; ++j) {
                const uint32_t* e = t.jobs[h][off + j];
                const int r = (int)(e[3] & 7u), x = (int)((e[3] >> 7) & 15u);
                if ((int)((e[3] >> 5) & 3u) != L - 1 || r < mlo || r > mhi) return false;
                const int klo = ((int)(e[1] << 4) - R2_SM_WB(L == 2 ? 1 : 0)) / 512, nblk = (int)((e[2] >> 9) & 3u);
                const int p1 = (int)((e[2] >> 11) & 3u), n1 = (int)((e[2] >> 13) & 3u), acc1 = (int)((e[2] >> 15) & 1u);
                const bool has2 = (e[2] >> 16) & 1u;
                const int p2 = (int)((e[2] >> 17) & 3u), n2 = (int)((e[2] >> 19) & 3u);
                if (has2 ? !((p1 == 0 && p2 == n1 && n1 + n2 == nblk) || (p2 == 0 && p1 == n2 && n1 + n2 == nblk)) : !(p1 == 0 && n1 == nblk))
                    return false;                                       // the first MMA (or its two parts) covers the job's blocks once
                for (int blk = 0; blk < nblk; ++blk) {
                    const int kh = klo + blk, tile = 7 - ((int)(e[2] & 511u) + 32 * blk) / 32;
                    if (kh > 2 || tile != r - kh + 1 || (tile >> 2) != h) return false;     // right columns, the warp's own half
                    if (tile != o) continue;
                    const bool in1 = blk >= p1 && blk < p1 + n1;
                    const int acc = in1 ? acc1 : 1;
                    if (inited ? acc != 1 : acc != 0) return false;
                    if (!inited && tile != r && tile != x) return false;                   // waits for what it replaces
                    inited = true;
                    if (seen & (1 << kh)) return false;
                    seen |= 1 << kh;
                    last = j;
                }
            }
            int want = 0;
            for (int kh = 0; kh < 3; ++kh)
                if (o + kh - 1 >= mlo && o + kh - 1 <= mhi) want |= 1 << kh;
            if (seen != want) return false;
            // the tile's item
            bool found = false;
            for (int i = 0; i < R2_ITEMS; ++i) {
                const uint32_t it = t.items[i];
                if (it == 0xFFFFFFFFu || (int)(it & 3u) != L || (int)((it >> 4) & 15u) != o) continue;
                if (found) return false;
                found = true;
                if ((int)((it >> 8) & 63u) != h * R2_HJOBS + r2_jloc(h, L, T) + last) return false;
                if ((int)(it >> 20) != 32 * (7 - o)) return false;
                const int job2 = (int)((it >> 14) & 63u);
                if (o == 3 || o == 4) {
                    const int g = 1 - h, off2 = (L > 1 ? t.cnt[g][1] : 0) + (L > 2 ? t.cnt[g][2] : 0);
                    if (job2 != g * R2_HJOBS + r2_jloc(g, L, T) || (int)(t.jobs[g][off2][3] & 7u) != o) return false;
                } else if (job2 != (int)((it >> 8) & 63u)) return false;
            }
            if (!found) return false;
            ++n_items;
        }
    }
    for (int i = 0; i < R2_ITEMS + 2; ++i)
        if (t.items[i] != 0xFFFFFFFFu && (t.items[i] & 3u) == 0u) ++n_items;    // conv1: one item per tile
    int listed = 0;
    for (int i = 0; i < R2_ITEMS + 2; ++i) listed += t.items[i] != 0xFFFFFFFFu;
    return n_items == listed && listed == T + T + (T - 2) + (T - 4) && t.cnt[0][0] + t.cnt[1][0] == T &&
           t.cnt[0][0] + t.cnt[0][1] + t.cnt[0][2] + t.cnt[0][3] <= R2_HJOBS && t.cnt[1][0] + t.cnt[1][1] + t.cnt[1][2] + t.cnt[1][3] <= R2_HJOBS;
}
static_assert(r2_check_tables(8), "schedule of the 8 x 8 image");
static_assert(r2_check_tables(7), "schedule of the 7 x 6 image");
__constant__ R2Tables c_r2_tables_8 = r2_make_tables(8);
__constant__ R2Tables c_r2_tables_7 = r2_make_tables(7);

// conv1's A operand for one row (tile r, lane): the 9 taps of player * grid around cell (r, c) as bf16 (+1 own, -1 opp,
// 0 empty / off-board), then zeros up to K = 16; c < 0 (pad column, unused lane, missing board) = a zero row.
__device__ __forceinline__ void r2_im2col_row(uint8_t* im2col, int row, int r, int c, uint64_t own, uint64_t opp) {
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
    uint8_t* dst = im2col + (R2_MARGIN + row) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(dst + R2_PLANE) = make_uint4(w[4], 0u, 0u, 0u);
}

template <int T, int C>
__global__ void __launch_bounds__(R2_THREADS, 2) k_net_trunk_rows2(TrunkRows2Params p) {
    constexpr int PITCH = C + 1, NB = 128 / PITCH, NPOS = (T - 4) * (C - 4);     // lanes per board, boards per group, outputs per board
    static_assert(T >= 7 && T <= R2_TILES && C <= 8 && NPOS % 2 == 0, "image shapes this kernel is laid out for");
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const R2Tables& tb = T == 8 ? c_r2_tables_8 : c_r2_tables_7;
    const int n_groups = (p.G + NB - 1) / NB;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = smem_u32(smem + R2_SM_BARS);
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + R2_SM_TMEM_PTR);
    const uint8_t* w_rows = p.wblob + AZ_TRUNK_ROWS_BLOB_OFFSET;       // conv2 | conv3 | conv4 in this kernel's order

    // ---- one-time setup: barriers, weights (bulk copies), the item / job tables, zeroed margins and unused lanes, TMEM
    if (tid < R2B_COUNT) {
        // arrivals are per WARP (lane 0 after __syncwarp): 32 lanes arriving on one barrier word serialise
        mbar_init(bars + 8 * tid, (tid >= R2B_READY && tid < R2B_READY + 8) ? 4u : 1u);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        // the weights do not depend on the launch before this one: conv1's tile + the four bias rows, conv2 and conv3
        if (tid == R2B_W1) r2_bulk_load(sbase + R2_SM_W1, w_rows - 1536, 1024u + 512u, bars + 8 * tid);
        if (tid == R2B_WB) r2_bulk_load(sbase + R2_SM_WB(0), w_rows, R2_LAYER_BYTES, bars + 8 * tid);
        if (tid == R2B_WB + 1) r2_bulk_load(sbase + R2_SM_WB(1), w_rows + R2_LAYER_BYTES, R2_LAYER_BYTES, bars + 8 * tid);
    }
    // the rows nothing ever writes: the margins of the buffer (the row before tile 0 is lane 0's left neighbour = zero
    // padding) and the two unused lanes of every tile (lane 127 is the left neighbour of the next tile's lane 0)
    if (tid >= 128 && tid < 192) {
        const int u = tid - 128, pl = (u >> 4) & 3, side = (u >> 3) & 1, row = u & 7;
        *reinterpret_cast<uint4*>(smem + R2_SM_BUF + pl * R2_PLANE + (side ? R2_MARGIN + R2_ROWS + row : row) * 16) = make_uint4(0u, 0u, 0u, 0u);
    } else if (tid >= 192 && tid < 256) {
        const int u = tid - 192, pl = (u >> 4) & 3, tile = (u >> 1) & 7, ln = NB * PITCH + (u & 1);
        static_assert(NB * PITCH == 126, "two spare lanes per tile");
        *reinterpret_cast<uint4*>(smem + R2_SM_BUF + pl * R2_PLANE + (R2_MARGIN + tile * 128 + ln) * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
    if (warp == R2_MMA_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(256));
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

    const uint32_t idesc32 = umma_idesc_bf16_m128(32u);
    constexpr uint32_t IDESC_BLK = (32u >> 3) << 17;              // one more kernel row = 32 more columns
    uint32_t k = 0;                                               // group counter of this CTA
    for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++k) {
        const int board0 = grp * NB, n_boards = min(NB, p.G - board0);
        // (conv4 of the previous group has completed -- block barrier at the end of the loop: its weights make room for conv2's)
        if (k > 0 && tid == 0) r2_bulk_load(sbase + R2_SM_WB(0), w_rows, R2_LAYER_BYTES, bars + 8 * R2B_WB);
        {   // conv1's A operand (9-tap im2col of player * grid, K = 16) in planes 2 - 3 of the buffer, by all threads: up to
            // four rows each, the boards loaded before the first row is built
            uint8_t* im2col = smem + R2_SM_BUF + 2 * R2_PLANE;
            uint64_t o[4], q[4];
            int cc[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int row = tid + j * R2_THREADS, l = row & 127, bb = l / PITCH;
                cc[j] = l - bb * PITCH;
                const bool ok = row < T * 128 && bb < n_boards && l < NB * PITCH && cc[j] < C;
                o[j] = ok ? p.own[board0 + bb] : 0ULL;
                q[j] = ok ? p.opp[board0 + bb] : 0ULL;
                if (!ok) cc[j] = -1;
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const int row = tid + j * R2_THREADS;
                if (row < T * 128) r2_im2col_row(im2col, row, row >> 7, cc[j], o[j], q[j]);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
        }

        if (warp >= R2_MMA_WARP) {
            // ============================================================== MMA issue: warp h issues the tile jobs that
            // accumulate into the output tiles 4 h .. 4 h + 3, in program order, by one elected thread of a converged warp
            const int h = warp - R2_MMA_WARP;
            const uint32_t jobs = bars + 8 * (R2B_JOB + h * R2_HJOBS);
            const int c0 = tb.cnt[h][0], c1 = tb.cnt[h][1], c2 = tb.cnt[h][2], n_jobs = c1 + c2 + tb.cnt[h][3];
            const uint32_t wbase_lo = r2_desc_lo(sbase, 96u * 16u);
            if (k == 0) r2_wait(bars + 8 * R2B_W1, 0u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // conv1: one K = 16 MMA per tile into the tile's accumulator
            if (elect_one()) {
                const uint32_t w1_lo = r2_desc_lo(sbase + R2_SM_W1, 512u);
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = h ? 4 + i : 3 - i;
                    if (r >= T) break;
                    r2_mma(tmem_base + (uint32_t)(32 * (7 - r)),
                           r2_desc_lo(sbase + R2_SM_BUF + 2 * R2_PLANE + (uint32_t)(R2_MARGIN + r * 128) * 16u, R2_PLANE), w1_lo, idesc32, 0u);
                    umma_commit(jobs + 8 * i);
                }
            }
            __syncwarp();
            const uint32_t abuf_lo = r2_desc_lo(sbase + R2_SM_BUF, R2_PLANE);
            const uint32_t ready = bars + 8 * R2B_READY;
#pragma unroll 1
            for (int jj = 0; jj < n_jobs; ++jj) {
                const uint4 e = make_uint4(tb.jobs[h][jj][0], wbase_lo + tb.jobs[h][jj][1], tb.jobs[h][jj][2], tb.jobs[h][jj][3]);
                // the layer's weights: conv2 = load 2 k of buffer 0, conv4 = load 2 k + 1 of it; conv3 stays in buffer 1
                if (jj == 0) r2_wait(bars + 8 * R2B_WB, 0u);
                else if (jj == c1) r2_wait(bars + 8 * (R2B_WB + 1), 0u);
                else if (jj == c1 + c2) r2_wait(bars + 8 * R2B_WB, 1u);
                {   // the input: layer L - 1's output tile r = write (k * P + L - 1) of tile r into the buffer (P = layers that
                    // write tile r); and tile x, whose accumulator this job initialises for layer L
                    const uint32_t r = e.w & 7u, Lm1 = (e.w >> 5) & 3u, x = (e.w >> 7) & 15u;
                    r2_wait(ready + 8 * r, (k * r2_writes((int)r, T) + Lm1) & 1u);
                    if (x != 15u) r2_wait(ready + 8 * x, (k * r2_writes((int)x, T) + Lm1) & 1u);
                }
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t d_tmem = tmem_base + (e.z & 511u);
                const uint32_t idesc = idesc32 + (((e.z >> 9) & 3u) - 1u) * IDESC_BLK;
                const uint32_t a_lo = abuf_lo + e.x, b_lo = e.y;
                const bool reload = h == 0 && jj == c1 + 2;       // warp 0's third job of conv3
                if (reload) {
                    // conv2's MMAs of both warps have long completed: conv4's weights may replace conv2's
                    constexpr int last_a = r2_jloc(0, 2, T) - 1, last_b = r2_jloc(1, 2, T) - 1;      // the warps' last jobs of conv2
                    r2_wait(bars + 8 * (R2B_JOB + last_a), k & 1u);
                    r2_wait(bars + 8 * (R2B_JOB + R2_HJOBS + last_b), k & 1u);
                }
                if (elect_one()) {
                    if (reload) r2_bulk_load(sbase + R2_SM_WB(0), w_rows + 2 * R2_LAYER_BYTES, R2_LAYER_BYTES, bars + 8 * R2B_WB);
                    {   // kernel column 0, channels 0..15: whole, or the block that must not accumulate and then the rest
                        const uint32_t pos = (e.z >> 11) & 3u;
                        r2_mma(d_tmem + 32u * pos, a_lo, b_lo + 32u * pos, idesc32 + (((e.z >> 13) & 3u) - 1u) * IDESC_BLK, (e.z >> 15) & 1u);
                        if ((e.z >> 16) & 1u) {
                            const uint32_t pos2 = (e.z >> 17) & 3u;
                            r2_mma(d_tmem + 32u * pos2, a_lo, b_lo + 32u * pos2, idesc32 + (((e.z >> 19) & 3u) - 1u) * IDESC_BLK, 1u);
                        }
                    }
#pragma unroll
                    for (int i = 1; i < 6; ++i) {
                        const int kw = i >> 1, kk = i & 1;
                        r2_mma(d_tmem, a_lo + (uint32_t)(kw + kk * (2 * R2_PLANE >> 4)),
                               b_lo + (uint32_t)(kw * (R2_W_DC_BYTES >> 4) + kk * (2 * 96 * 16 >> 4)), idesc, 1u);
                    }
                    umma_commit(jobs + 8 * (c0 + jj));
                }
                __syncwarp();
            }
        } else {
            // ============================================================== epilogue team `team`: the items of its issuing
            // warp's half of the tiles
            const int team = warp >> 2, quarter = warp & 3;
            const int tl = quarter * 32 + lane;                     // TMEM lane = lane of the tile
            const int b = tl / PITCH, c = tl - b * PITCH;           // board of the group, column (C = pad)
            const float* bias_all = reinterpret_cast<const float*>(smem + R2_SM_BIAS);
            const uint32_t t_quarter = tmem_base + ((uint32_t)(quarter * 32) << 16);
            const bool used = tl < NB * PITCH;                      // not one of the tile's two spare lanes
            const bool cell = b < n_boards && used && c < C;        // a real cell of a real board
            if (k == 0) r2_wait(bars + 8 * R2B_W1, 0u);             // the biases
#pragma unroll 1
            for (int idx = 0; idx < R2_ITEMS; ++idx) {
                const uint32_t it = tb.items[idx];
                if (it == 0xFFFFFFFFu) break;
                const int L = (int)(it & 3u), r = (int)((it >> 4) & 15u);
                if ((r >> 2) != team) continue;                     // the other team's half of the tiles
                // (one polling warp per team, the other three behind a named barrier: measured, no difference)
                r2_wait(bars + 8 * (R2B_JOB + ((it >> 8) & 63u)), k & 1u);
                if (((it >> 14) & 63u) != ((it >> 8) & 63u)) r2_wait(bars + 8 * (R2B_JOB + ((it >> 14) & 63u)), k & 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const float* bias = bias_all + L * 32;
                uint32_t acc[32], out[16];
                tmem_ld32(t_quarter + (it >> 20), acc);
#pragma unroll
                for (int co = 0; co < 32; co += 2) {
                    const float2 bb = *reinterpret_cast<const float2*>(bias + co);
                    out[co >> 1] = pack_relu_bf16x2(__uint_as_float(acc[co]) + bb.x, __uint_as_float(acc[co + 1]) + bb.y);
                }
                // ---- store: activations in place (tile r of the buffer) or, after conv4, the features
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
                        uint8_t* d = smem + R2_SM_BUF + (size_t)(R2_MARGIN + r * 128 + tl) * 16;
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(d + ch * R2_PLANE) = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    }
                    // "ready" = the tile is in the buffer AND its accumulator has been read (tmem_ld32 waits for the data)
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) r2_arrive(bars + 8 * (R2B_READY + r));
                    if (p.debug && p.debug_layer == L + 1) {
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(p.debug + ((size_t)grp * R2_ROWS + r * 128 + tl) * 32 + ch * 8) =
                                used ? make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]) : make_uint4(0u, 0u, 0u, 0u);
                    }
                } else if (cell && c >= 2 && c < C - 2) {
                    // conv4: the outputs are the cells at an inset of 2; features[board][pos * 32 + co], pos = (r-2)*(C-4) + (c-2)
                    const int pos = (r - 2) * (C - 4) + (c - 2), g = board0 + b;
                    if (p.features_packed) {
                        // K = pos*32 + 8*ch: chunk pos / 2 (64 wide, 16 KB per row tile of 128 positions), plane (pos % 2)*4 + ch
                        uint8_t* f = reinterpret_cast<uint8_t*>(p.features) + ((size_t)(g >> 7) * (NPOS / 2) + (size_t)(pos >> 1)) * 16384 +
                                     (size_t)((pos & 1) * 4) * 2048 + (size_t)(g & 127) * 16;
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(f + ch * 2048) = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    } else {
                        uint4* f = reinterpret_cast<uint4*>(p.features + ((size_t)g * NPOS + pos) * 32);
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            f[ch] = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    }
                }
            }
        }
        // the group is through when conv4's items are (they waited for the last job of either warp): the next group may
        // overwrite the buffer, the accumulators and conv4's weights
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }

    if (warp == R2_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256));
}

// az_net_trunk takes this kernel instead of k_net_trunk_rows when the batch has more groups of 14 boards than the GPU
// has SMs (two CTAs per SM then; with fewer groups a CTA would be alone on its SM, where k_net_trunk_rows' three
// teams and two accumulator regions are faster: 10.6 against 11.8 us at 1024 positions).  AZ_B200_TRUNK_ROWS1=1 in the
// environment keeps k_net_trunk_rows for every size (A/B switch).
bool az_trunk_rows2_enabled() {
    static const bool on = getenv("AZ_B200_TRUNK_ROWS1") == nullptr;
    return on;
}

int az_trunk_rows2_boards_per_group(int rows) { return rows == 8 ? 14 : 18; }

int az_trunk_rows2_launch(const uint64_t* own, const uint64_t* opp, int rows, int cols, int64_t G, const void* wblob, void* features,
                          int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream) {
    if (!((rows == 8 && cols == 8) || (rows == 7 && cols == 6))) return AZ_ERR_ARG;
    static bool attr_set = false;
    if (!attr_set) {
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk_rows2<8, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, R2_SM_TOTAL));
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk_rows2<8, 8>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk_rows2<7, 6>, cudaFuncAttributeMaxDynamicSharedMemorySize, R2_SM_TOTAL));
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk_rows2<7, 6>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
        attr_set = true;
    }
    TrunkRows2Params p;
    p.own = own; p.opp = opp; p.wblob = (const uint8_t*)wblob; p.features = (__nv_bfloat16*)features;
    p.debug = (__nv_bfloat16*)debug; p.G = (int)G; p.debug_layer = debug ? debug_layer : 0;
    p.features_packed = features_packed ? 1 : 0; p.late_trigger = late_trigger;
    const int nb = az_trunk_rows2_boards_per_group(rows);
    const int n_groups = (int)((G + nb - 1) / nb);
    const int grid = n_groups < 2 * n_sms ? n_groups : 2 * n_sms;      // persistent: two CTAs per SM
    cudaError_t e = rows == 8 ? az_launch_pdl(k_net_trunk_rows2<8, 8>, dim3(grid), dim3(R2_THREADS), R2_SM_TOTAL, (cudaStream_t)stream, p)
                              : az_launch_pdl(k_net_trunk_rows2<7, 6>, dim3(grid), dim3(R2_THREADS), R2_SM_TOTAL, (cudaStream_t)stream, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_trunk_rows2");
    AZ_CHECK_LAUNCH("k_net_trunk_rows2");
    return AZ_OK;
}
