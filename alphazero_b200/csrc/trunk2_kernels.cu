// Convolution trunk of OthelloNet(n = 8), second kernel: no cross-lane traffic in the epilogue.
// (alphazero/games/othello.py:366-374: conv1..conv4 + BatchNorm + ReLU, pads 1, 1, 0, 0; same weight blob, same outputs
// as k_net_trunk<8, 8> in net_kernels.cu.)
//
// Why.  In the first kernel a 3x3 convolution is 3 row-shifted A windows (the kernel ROW as a shift of the operand by 8
// rows) times N = 96 = (kernel COLUMN, c_out), and the epilogue moves the column terms to their output with 64 warp
// shuffles per row: measured, those shuffles are 315 of a tile's 765 epilogue cycles and the epilogue is what the
// kernel waits for (profiles/r2_trunk_experiments.txt).  Here the roles are swapped and the board is laid out so that
// BOTH shifts are free:
//   * a tile (128 TMEM lanes) is ONE BOARD ROW of 14 boards: lane = board * 9 + column (8 cells + 1 zero pad column,
//     lanes 126 / 127 unused and zero); a group of 14 boards is 8 tiles per layer;
//   * the kernel COLUMN is a shift of the A operand by +-1 row = 16 bytes -- a K-major no-swizzle operand may start at
//     any 16-byte row (tools/mma2_probe.py, profiles/r2_mma2_probe.txt) -- and the pad column / unused lanes supply
//     the zero padding between boards;
//   * the kernel ROW is N: N = 96 = (kernel row kh, c_out).  Block kh of the accumulator of INPUT tile r' belongs to
//     OUTPUT tile r' - kh + 1 at the SAME lane: out[r] = D[r-1].block0 + D[r].block1 + D[r+1].block2.  The epilogue of
//     an output tile reads one 32-column block from each of three accumulators and adds them: no shuffles, and the
//     missing accumulators at the top / bottom edge are conv2's zero padding.
// Same six N = 96 MMAs per tile as before (the tensor-pipe cost is unchanged), the same TMEM read volume, a third of
// the epilogue's instructions gone; conv3 / conv4 only convert the output rows that can be valid (6 and 4 of 8 tiles).
//
// Pipeline.  One MMA warp issues tile jobs in program order -- per group conv1 x 8, conv2 x 8, conv3 x 8, conv4 x 6 -- into a
// ring of 5 accumulator stages (96 TMEM columns each).  Layer L + 1's MMA on tile r only needs layer L's OUTPUT tile r
// (the +-1 row bleed at a tile's ends lands in unused lanes), so layers overlap tile by tile and the activations are
// converted IN PLACE: output tile r overwrites input tile r once the accumulators r-1 .. r+1 are complete, which
// implies that the MMAs that read tile r have finished.  Three epilogue teams take the output tiles round robin; an
// accumulator stage is released by its three consumer items (one arrival per warp; "virtual" items just outside a layer's tile range
// supply the arrivals the edges lack, conv1 items arrive three times).  Two activation buffers alternate between the
// CTA's groups; two producer warps build conv1's im2col operand of the next group while this one runs.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "trunk2.cuh"

#define T2_NB 14                  // boards per group
#define T2_PITCH 9                // lanes per board in a tile
#define T2_TILES 8                // tiles per layer = board rows
#define T2_ROWS (T2_TILES * 128)
#define T2_MARGIN 8               // zero rows before / after a buffer (the +-1 row shifts of the first / last tile)
#define T2_PLANE ((T2_ROWS + 2 * T2_MARGIN) * 16)      // one 8-channel chunk plane: 16 640 B (= LBO of the A operand)
#define T2_BUF (4 * T2_PLANE)     // activation buffer: 66 560 B
#define T2_IM2COL (2 * T2_PLANE)  // conv1's A operand (K = 16: 9 taps + zeros): 33 280 B
#define T2_STAGES 5
#define T2_STAGE_COLS 96
#define T2_TEAMS 3
#define T2_EPI_THREADS (128 * T2_TEAMS)
#define T2_MMA_WARP (T2_EPI_THREADS / 32)
#define T2_MMA_WARPS 2            // tile jobs alternate between two issuing warps: a job costs its thread ~400 cycles (six MMAs,
                                  // three barrier waits, a commit) against 337 cycles of tensor-pipe time
#define T2_PRODUCERS 64
#define T2_THREADS (T2_EPI_THREADS + 32 * T2_MMA_WARPS + T2_PRODUCERS)
#define T2_JOBS 30                // MMA tile jobs per group: 8 + 8 + 8 + 6
#define T2_ITEMS 36               // epilogue items per group: 8 + 10 + 10 + 8 (a multiple of T2_TEAMS)
// shared memory map
#define T2_SM_BUF(i) ((i) * T2_BUF)
#define T2_SM_IM2COL (2 * T2_BUF)
#define T2_SM_W (T2_SM_IM2COL + T2_IM2COL)             // conv2..4 reordered: [layer 3][kw 3][ci chunk 4][(kh, co) 96][8] bf16
#define T2_W_DC_BYTES (4 * 96 * 16)                    // one (layer, kernel column): 6 144 B
#define T2_SM_W1 (T2_SM_W + 9 * T2_W_DC_BYTES)         // conv1: [k chunk 2][co 32][8] bf16
#define T2_SM_BIAS (T2_SM_W1 + 1024)                   // float [4][32]
#define T2_SM_BARS (T2_SM_BIAS + 512)
#define TB_FULL 0                                      // [5]  accumulator stage complete (tcgen05.commit)
#define TB_EMPTY 5                                     // [5]  stage read by its three consumers (3 x 128 arrivals)
#define TB_READY 10                                    // [2][8] output tile r of the group in buffer b is in shared memory
#define TB_IM2COL 26                                   // conv1's A operand of the next group is in place
#define TB_I2FREE 27                                   // conv1's MMAs of the current group have read it
#define TB_COUNT 28
#define T2_SM_ITEMS (T2_SM_BARS + 8 * 32)                // uint2 [36]: what each epilogue item reads / waits for (static per group)
#define T2_SM_TMEM_PTR (T2_SM_ITEMS + 8 * 40)
#define T2_SM_TOTAL (T2_SM_TMEM_PTR + 16)
static_assert(T2_SM_TOTAL <= 232448, "shared memory per CTA on sm_100");
static_assert(T2_ITEMS % T2_TEAMS == 0, "the team of an item does not depend on the group");
static_assert(T2_JOBS % (2 * T2_STAGES) == 0, "stage and barrier parity of a job do not depend on the group");
#define T2_WAIT_LIMIT (4000000000LL)

struct Trunk2Params {
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
__device__ __forceinline__ void t2_wait(uint32_t bar, uint32_t parity) {
    long long t0 = 0;
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        const long long t = clock64();
        if (t0 == 0) t0 = t;
        else if (t - t0 > T2_WAIT_LIMIT) __trap();
    }
}
// arrivals that do not wait for the barrier's state word to come back
__device__ __forceinline__ void t2_arrive_n(uint32_t bar, uint32_t n) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(n) : "memory");
}
__device__ __forceinline__ void t2_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// Per layer L = 0..3 (conv1..conv4): first MMA tile, number of MMA tiles, job offset inside a group; first epilogue
// item tile (virtual edge items included), number of items, item offset; first / last REAL output tile.
__device__ __forceinline__ int t2_mlo(int L) { return L == 3 ? 1 : 0; }
__device__ __forceinline__ int t2_mcnt(int L) { return L == 3 ? 6 : 8; }
__device__ __forceinline__ int t2_joff(int L) { return L * 8; }
__device__ __forceinline__ int t2_ilo(int L) { return L == 0 ? 0 : (L == 3 ? 0 : -1); }
__device__ __forceinline__ int t2_ioff(int L) { return L == 0 ? 0 : (L == 1 ? 8 : (L == 2 ? 18 : 28)); }
__device__ __forceinline__ int t2_olo(int L) { return L < 2 ? 0 : L - 1; }          // conv3: 1, conv4: 2
__device__ __forceinline__ int t2_ohi(int L) { return L < 2 ? 7 : 8 - L; }          // conv3: 6, conv4: 5

// conv1's A operand for one row (tile r, lane): the 9 taps of player * grid around cell (r, c) as bf16 (+1 own, -1 opp,
// 0 empty / off-board), then zeros up to K = 16; c < 0 (pad column, unused lane, missing board) = a zero row.
__device__ __forceinline__ void t2_im2col_row(uint8_t* im2col, int row, int r, int c, uint64_t own, uint64_t opp) {
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
    uint8_t* dst = im2col + (T2_MARGIN + row) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(dst + T2_PLANE) = make_uint4(w[4], 0u, 0u, 0u);
}

__global__ void __launch_bounds__(T2_THREADS, 1) k_net_trunk2(Trunk2Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_groups = (p.G + T2_NB - 1) / T2_NB;
    // debug_layer 200: cycle stamps instead of activations, int64 [n_ctas][16] (tools/trunk2_probe.py)
    long long* stamps = (p.debug && p.debug_layer == 200) ? reinterpret_cast<long long*>(p.debug) + (size_t)blockIdx.x * 16 : nullptr;
    const long long t_entry = stamps ? clock64() : 0;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = smem_u32(smem + T2_SM_BARS);
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + T2_SM_TMEM_PTR);

    // ---- one-time setup: barriers, zeroed buffers (margins, pad lanes), weights (reordered from the blob), TMEM
    if (tid < TB_COUNT) {
        // arrivals are per WARP (lane 0 after __syncwarp): 32 lanes arriving on one barrier word serialise
        uint32_t count = 1u;                                       // full[]: one tcgen05.commit
        if (tid >= TB_EMPTY && tid < TB_READY) count = 3u * 4u;    // empty[]: three consumer items of 4 warps
        else if (tid >= TB_READY && tid < TB_IM2COL) count = 4u;   // ready[][]: the team that wrote the tile
        else if (tid == TB_IM2COL) count = T2_PRODUCERS / 32;
        else if (tid == TB_I2FREE) count = T2_MMA_WARPS;           // each issuing warp commits after its last conv1 job
        mbar_init(bars + 8 * tid, count);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid >= 64 && tid < 64 + T2_ITEMS) {
        // item idx of a group: layer L, output tile r, which accumulators D[r-1], D[r], D[r+1] exist (bits 0..2), whether the
        // item produces an output (bit 3); their stages and the parity of their `full` phase.  T2_JOBS is a multiple of
        // 2 * T2_STAGES, so none of this depends on the group.
        const int idx = tid - 64;
        const int L = idx < 8 ? 0 : (idx < 18 ? 1 : (idx < 28 ? 2 : 3));
        const int r = t2_ilo(L) + idx - t2_ioff(L);
        const int mlo = t2_mlo(L), mhi = mlo + t2_mcnt(L) - 1;
        const int j1 = t2_joff(L) + r - mlo;                    // job of D[r] within the group (if it exists)
        uint32_t fl = 0u, st = 0u;
        for (int i = 0; i < 3; ++i) {
            const int rr = r - 1 + i, j = j1 - 1 + i;
            const bool ex = L == 0 ? i == 1 : (rr >= mlo && rr <= mhi);
            if (ex) { fl |= 1u << i; st |= (uint32_t)(j % T2_STAGES) << (4 * i); st |= (uint32_t)((j / T2_STAGES) & 1) << (12 + i); }
        }
        if (r >= t2_olo(L) && r <= t2_ohi(L)) fl |= 8u;
        reinterpret_cast<uint2*>(smem + T2_SM_ITEMS)[idx] = make_uint2((uint32_t)L | ((uint32_t)(r + 1) << 4) | (fl << 8), st);
    }
    // the only rows nothing ever writes: the margins of the activation buffers (the row before tile 0 is lane 0's left
    // neighbour = zero padding); pad lanes and unused lanes are written with every conv1 output
    if (tid < 2 * 4 * 2 * T2_MARGIN) {
        const int bufi = tid >> 6, pl = (tid >> 4) & 3, side = (tid >> 3) & 1, row = tid & 7;
        *reinterpret_cast<uint4*>(smem + T2_SM_BUF(bufi) + pl * T2_PLANE + (side ? T2_MARGIN + T2_ROWS + row : row) * 16) =
            make_uint4(0u, 0u, 0u, 0u);
    }
    {   // blob: [layer][kh][ci chunk][(kw, co)][8] -> shared memory [layer][kw][ci chunk][(kh, co)][8], 16-byte units
        const uint4* src = reinterpret_cast<const uint4*>(p.wblob);
        uint4* dst = reinterpret_cast<uint4*>(smem + T2_SM_W);
        constexpr int UNITS = 3 * 3 * 4 * 96, PER = (UNITS + T2_THREADS - 1) / T2_THREADS;
        uint4 v[PER];                                              // all loads in flight before the first store
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int u = tid + i * T2_THREADS;
            v[i] = u < UNITS ? src[u] : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int u = tid + i * T2_THREADS;
            if (u < UNITS) {
                const int co = u & 31, kw = (u >> 5) % 3, cc = (u / 96) & 3, kh = (u / 384) % 3, li = u / 1152;
                dst[(((li * 3 + kw) * 4 + cc) * 3 + kh) * 32 + co] = v[i];
            }
        }
        const uint4* tail = src + 3 * 3 * 4 * 96;                  // conv1 tile (1024 B), then the four bias rows (512 B)
        for (int u = tid; u < (1024 + 512) / 16; u += T2_THREADS) reinterpret_cast<uint4*>(smem + T2_SM_W1)[u] = tail[u];
    }
    if (warp == T2_MMA_WARP) {
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

    if (warp >= T2_MMA_WARP + T2_MMA_WARPS) {
        // ================================================================== producers: conv1's A operand (9-tap im2col
        // of player * grid, K = 16) of every group of this CTA.  Thread = lanes ptid and ptid + 64 of all 8 tiles.
        const int ptid = tid - (T2_EPI_THREADS + 32 * T2_MMA_WARPS);
        uint8_t* im2col = smem + T2_SM_IM2COL;
        // (the first group's is built by the epilogue teams, which have nothing else to do yet)
        uint32_t k = 1;
        for (int grp = blockIdx.x + gridDim.x; grp < n_groups; grp += gridDim.x, ++k) {
            const int board0 = grp * T2_NB, n_boards = min(T2_NB, p.G - board0);
            uint64_t own[2], opp[2];
            int col[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int tl = ptid + 64 * j, b = tl / T2_PITCH;
                col[j] = tl - b * T2_PITCH;
                const bool ok = b < n_boards && tl < T2_NB * T2_PITCH && col[j] < 8;
                own[j] = ok ? p.own[board0 + b] : 0ULL;
                opp[j] = ok ? p.opp[board0 + b] : 0ULL;
                if (!ok) col[j] = -1;
            }
            t2_wait(bars + 8 * TB_I2FREE, (k - 1u) & 1u);                 // conv1 of the previous group has read the buffer
#pragma unroll
            for (int j = 0; j < 2; ++j) {
#pragma unroll 2
                for (int r = 0; r < T2_TILES; ++r) {
                    t2_im2col_row(im2col, r * 128 + ptid + 64 * j, r, col[j], own[j], opp[j]);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + 8 * TB_IM2COL);
        }
    } else if (warp >= T2_MMA_WARP) {
        // ================================================================== MMA issue: tile jobs in program order, job q by warp q % 2
        const uint32_t mw = (uint32_t)(warp - T2_MMA_WARP);
        const uint32_t idesc32 = umma_idesc_bf16_m128(32u), idesc96 = umma_idesc_bf16_m128(96u);
        uint32_t q = 0, k = 0;                        // job counter (stage = q % 5), group counter of this CTA
        long long w_ready = 0, w_empty = 0, t_issue = 0, w_im2col = 0;
        for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++k) {
            const uint32_t buf = k & 1u, use = k >> 1;
            const long long c_i = stamps ? clock64() : 0;
            t2_wait(bars + 8 * TB_IM2COL, k & 1u);
            if (stamps) w_im2col += clock64() - c_i;
#pragma unroll 1
            for (int L = 0; L < 4; ++L) {
                const int lo = t2_mlo(L), cnt = t2_mcnt(L);
#pragma unroll 1
                for (int r = lo; r < lo + cnt; ++r, ++q) {
                    if ((q & 1u) != mw) continue;
                    const long long c0 = stamps ? clock64() : 0;
                    if (L > 0) {   // layer L - 1's output tile r: phase (use * P + L - 1) of ready[buf][r], P = layers that write tile r
                        const uint32_t P = (r >= 1 && r <= 6) ? 3u : 2u;
                        t2_wait(bars + 8 * (TB_READY + buf * 8 + r), (use * P + (uint32_t)(L - 1)) & 1u);
                    }
                    const uint32_t s = q % T2_STAGES, n = q / T2_STAGES;
                    const long long c1 = stamps ? clock64() : 0;
                    t2_wait(bars + 8 * (TB_EMPTY + s), (n & 1u) ^ 1u);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const long long c2 = stamps ? clock64() : 0;
                    if (elect_one()) {
                        const uint32_t d_tmem = tmem_base + s * T2_STAGE_COLS;
                        if (L == 0) {
                            umma_bf16(d_tmem, umma_desc(sbase + T2_SM_IM2COL + (uint32_t)(T2_MARGIN + r * 128) * 16u, T2_PLANE, 128u),
                                      umma_desc(sbase + T2_SM_W1, 512u, 128u), idesc32, 0u);
                        } else {
                            const uint32_t a0 = sbase + T2_SM_BUF(buf) + (uint32_t)(T2_MARGIN + r * 128) * 16u;
                            const uint32_t b0 = sbase + T2_SM_W + (uint32_t)((L - 1) * 3) * T2_W_DC_BYTES;
#pragma unroll
                            for (int kw = 0; kw < 3; ++kw) {
#pragma unroll
                                for (int kk = 0; kk < 2; ++kk) {
                                    // A: the tile's rows shifted by kw - 1 columns (= rows of 16 bytes), channels 16 kk ..
                                    umma_bf16(d_tmem,
                                              umma_desc(a0 + (uint32_t)((kw - 1) * 16 + kk * 2 * T2_PLANE), T2_PLANE, 128u),
                                              umma_desc(b0 + (uint32_t)(kw * T2_W_DC_BYTES + kk * 2 * 96 * 16), 96u * 16u, 128u),
                                              idesc96, (kw | kk) ? 1u : 0u);
                                }
                            }
                        }
                        umma_commit(bars + 8 * (TB_FULL + s));
                        // conv1 has read the im2col buffer: the producers may build the next group's
                        if (L == 0 && r >= T2_TILES - 2) umma_commit(bars + 8 * TB_I2FREE);     // this warp's last conv1 job
                    }
                    __syncwarp();
                    if (stamps) { const long long c3 = clock64(); w_ready += c1 - c0; w_empty += c2 - c1; t_issue += c3 - c2; }
                }
                if (stamps && lane == 0 && k == 0 && mw == 0) stamps[8 + L] = clock64() - t_entry;      // group 0: layer L issued
            }
            if (stamps && lane == 0 && k == 0 && mw == 0) stamps[3] = clock64() - t_entry;
        }
        if (stamps && lane == 0 && mw == 0) { stamps[5] = w_ready; stamps[6] = w_empty; stamps[7] = t_issue; stamps[12] = w_im2col; }
    } else {
        // ================================================================== epilogue team `team`: items team, team + 3, ..
        const int team = warp >> 2, quarter = warp & 3;
        const int tl = quarter * 32 + lane;                     // TMEM lane = lane of the tile
        const int b = tl / T2_PITCH, c = tl - b * T2_PITCH;     // board of the group, column (8 = pad)
        const float* bias_all = reinterpret_cast<const float*>(smem + T2_SM_BIAS);
        const uint32_t t_quarter = tmem_base + ((uint32_t)(quarter * 32) << 16);
        {   // conv1's A operand of the CTA's first group, by all epilogue threads (three rows each)
            const int board0 = blockIdx.x * T2_NB, n_boards = min(T2_NB, p.G - board0);
            uint8_t* im2col = smem + T2_SM_IM2COL;
            for (int row = tid; row < T2_ROWS; row += T2_EPI_THREADS) {
                const int r = row >> 7, l = row & 127, bb = l / T2_PITCH, cc = l - bb * T2_PITCH;
                const bool ok = bb < n_boards && l < T2_NB * T2_PITCH && cc < 8;
                t2_im2col_row(im2col, row, r, ok ? cc : -1, ok ? p.own[board0 + bb] : 0ULL, ok ? p.opp[board0 + bb] : 0ULL);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("bar.sync 1, %0;" ::"r"(T2_EPI_THREADS) : "memory");
            if (tid < T2_PRODUCERS / 32) mbar_arrive(bars + 8 * TB_IM2COL);
            if (stamps && tid == 0) stamps[1] = clock64() - t_entry;
        }
        uint32_t k = 0;
        long long e_wait = 0, e_work = 0;
        for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++k) {
            const uint32_t buf = k & 1u;
            const int board0 = grp * T2_NB, n_boards = min(T2_NB, p.G - board0);
            const bool cell = b < n_boards && tl < T2_NB * T2_PITCH && c < 8;      // a real cell of a real board
#pragma unroll 1
            for (int idx = team; idx < T2_ITEMS; idx += T2_TEAMS) {
                const uint2 it = reinterpret_cast<const uint2*>(smem + T2_SM_ITEMS)[idx];
                const int L = (int)(it.x & 3u), r = (int)((it.x >> 4) & 15u) - 1;
                const bool e0 = (it.x >> 8) & 1u, e1 = (it.x >> 9) & 1u, e2 = (it.x >> 10) & 1u, real = (it.x >> 11) & 1u;
                const uint32_t s0 = it.y & 15u, s1 = (it.y >> 4) & 15u, s2 = (it.y >> 8) & 15u;      // stages of D[r-1], D[r], D[r+1]
                const long long f0 = stamps ? clock64() : 0;
                // (two issuing warps: completion is not in job order, wait for every accumulator the item touches)
                if (e0) t2_wait(bars + 8 * (TB_FULL + s0), (it.y >> 12) & 1u);
                if (e1) t2_wait(bars + 8 * (TB_FULL + s1), (it.y >> 13) & 1u);
                if (e2) t2_wait(bars + 8 * (TB_FULL + s2), (it.y >> 14) & 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const long long f1 = stamps ? clock64() : 0;
                e_wait += f1 - f0;
                const float* bias = bias_all + L * 32;
                uint32_t out[16];
                if (L == 0) {
                    uint32_t acc[32];
                    tmem_ld32(t_quarter + s1 * T2_STAGE_COLS, acc);
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) t2_arrive_n(bars + 8 * (TB_EMPTY + s1), 3u);      // conv1's accumulator has one consumer
#pragma unroll
                    for (int co = 0; co < 32; co += 2) {
                        const float2 bb = *reinterpret_cast<const float2*>(bias + co);
                        out[co >> 1] = pack_relu_bf16x2(__uint_as_float(acc[co]) + bb.x, __uint_as_float(acc[co + 1]) + bb.y);
                    }
                } else {
                    // out[r] = D[r-1].block0 + D[r].block1 + D[r+1].block2 at the same lane (a missing accumulator at the
                    // top / bottom edge = a row of zero padding); two halves of 16 channels, the TMEM read of the second
                    // under the arithmetic of the first
                    uint32_t lo[2][16], mid[2][16], hi[2][16];
                    if (real) {
                        const uint32_t t0 = t_quarter + s0 * T2_STAGE_COLS;
                        const uint32_t t1 = t_quarter + s1 * T2_STAGE_COLS + 32u;
                        const uint32_t t2 = t_quarter + s2 * T2_STAGE_COLS + 64u;
                        if (!(e0 && e2)) {       // conv2's first / last row: the missing accumulator is a row of zero padding
#pragma unroll
                            for (int h = 0; h < 2; ++h) {
#pragma unroll
                                for (int i = 0; i < 16; ++i) { lo[h][i] = 0u; hi[h][i] = 0u; }
                            }
                        }
                        if (e0) tmem_ld16_issue(t0, lo[0]);
                        tmem_ld16_issue(t1, mid[0]);
                        if (e2) tmem_ld16_issue(t2, hi[0]);
                        tmem_ld48_wait(lo[0], mid[0], hi[0]);
                        if (e0) tmem_ld16_issue(t0 + 16u, lo[1]);
                        tmem_ld16_issue(t1 + 16u, mid[1]);
                        if (e2) tmem_ld16_issue(t2 + 16u, hi[1]);
#pragma unroll
                        for (int co = 0; co < 16; co += 2) {
                            const float2 bb = *reinterpret_cast<const float2*>(bias + co);
                            uint64_t x = f32x2_add(f32x2_pack(__uint_as_float(mid[0][co]), __uint_as_float(mid[0][co + 1])), f32x2_pack(bb.x, bb.y));
                            x = f32x2_add(x, f32x2_pack(__uint_as_float(lo[0][co]), __uint_as_float(lo[0][co + 1])));
                            x = f32x2_add(x, f32x2_pack(__uint_as_float(hi[0][co]), __uint_as_float(hi[0][co + 1])));
                            float x0, x1;
                            f32x2_unpack(x, x0, x1);
                            out[co >> 1] = pack_relu_bf16x2(x0, x1);
                        }
                        tmem_ld48_wait(lo[1], mid[1], hi[1]);
                    }
                    // every accumulator this item touches (read or not) gets this item's arrival: three per accumulator
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) {
                        if (e0) t2_arrive(bars + 8 * (TB_EMPTY + s0));
                        if (e1) t2_arrive(bars + 8 * (TB_EMPTY + s1));
                        if (e2) t2_arrive(bars + 8 * (TB_EMPTY + s2));
                    }
                    if (!real) continue;               // a virtual item (outside the layer's valid output rows)
#pragma unroll
                    for (int co = 0; co < 16; co += 2) {
                        const float2 bb = *reinterpret_cast<const float2*>(bias + 16 + co);
                        uint64_t x = f32x2_add(f32x2_pack(__uint_as_float(mid[1][co]), __uint_as_float(mid[1][co + 1])), f32x2_pack(bb.x, bb.y));
                        x = f32x2_add(x, f32x2_pack(__uint_as_float(lo[1][co]), __uint_as_float(lo[1][co + 1])));
                        x = f32x2_add(x, f32x2_pack(__uint_as_float(hi[1][co]), __uint_as_float(hi[1][co + 1])));
                        float x0, x1;
                        f32x2_unpack(x, x0, x1);
                        out[8 + (co >> 1)] = pack_relu_bf16x2(x0, x1);
                    }
                }
                // ---- store: activations in place (tile r of the group's buffer) or, after conv4, the features
                if (L < 3) {
                    // conv2 pads with zeros: conv1's output must be zero in the pad column, the unused lanes and the boards a
                    // partial group lacks.  (conv3 / conv4 do not pad: what they read there only reaches outputs that are
                    // not valid anyway, and it is finite.)
                    if (L == 0 && !cell) {
#pragma unroll
                        for (int i = 0; i < 16; ++i) out[i] = 0u;
                    }
                    uint8_t* d = smem + T2_SM_BUF(buf) + (size_t)(T2_MARGIN + r * 128 + tl) * 16;
#pragma unroll
                    for (int ch = 0; ch < 4; ++ch)
                        *reinterpret_cast<uint4*>(d + ch * T2_PLANE) = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) t2_arrive(bars + 8 * (TB_READY + buf * 8 + r));
                    if (p.debug && p.debug_layer == L + 1) {
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(p.debug + ((size_t)grp * T2_ROWS + r * 128 + tl) * 32 + ch * 8) =
                                make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    }
                } else if (cell && c >= 2 && c < 6) {
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
                if (stamps) e_work += clock64() - f1;
            }
        }
        if (stamps && lane == 0 && quarter == 0 && team == 0) { stamps[13] = e_wait; stamps[14] = e_work; }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == T2_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    if (stamps && tid == 0) stamps[1 + 10] = clock64() - t_entry;      // [11] exit
}

// Opt-in (measured slower than the first kernel so far, profiles/r2_trunk_experiments.txt): AZ_B200_TRUNK_V2=1 selects
// this kernel for every 8 x 8 batch of AZ_TRUNK2_MIN_G positions and more; bit 1 of az_net_trunk's `features_packed`
// selects it for one call.
bool az_trunk2_enabled() {
    static const bool on = getenv("AZ_B200_TRUNK_V2") != nullptr;
    return on;
}

int az_trunk2_launch(const uint64_t* own, const uint64_t* opp, int64_t G, const void* wblob, void* features,
                     int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream) {
    static bool attr_set = false;
    if (!attr_set) {
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk2, cudaFuncAttributeMaxDynamicSharedMemorySize, T2_SM_TOTAL));
        attr_set = true;
    }
    Trunk2Params p;
    p.own = own; p.opp = opp; p.wblob = (const uint8_t*)wblob; p.features = (__nv_bfloat16*)features;
    p.debug = (__nv_bfloat16*)debug; p.G = (int)G; p.debug_layer = debug ? debug_layer : 0;
    p.features_packed = features_packed ? 1 : 0; p.late_trigger = late_trigger;
    const int n_groups = (int)((G + T2_NB - 1) / T2_NB);
    const int grid = n_groups < n_sms ? n_groups : n_sms;      // persistent: one CTA per SM
    cudaError_t e = az_launch_pdl(k_net_trunk2, dim3(grid), dim3(T2_THREADS), T2_SM_TOTAL, (cudaStream_t)stream, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_trunk2");
    AZ_CHECK_LAUNCH("k_net_trunk2");
    return AZ_OK;
}

