// Convolution trunk of OthelloNet(n = 8), third kernel: no cross-lane traffic in the epilogue AND one accumulator per
// output tile.  (alphazero/games/othello.py:366-374: conv1..conv4 + BatchNorm + ReLU, pads 1, 1, 0, 0; same weight blob,
// same outputs as k_net_trunk<8, 8> in net_kernels.cu.)
//
// Layout as in trunk2_kernels.cu: a tile (128 TMEM lanes) is ONE BOARD ROW of 14 boards, lane = board * 9 + column (8
// cells + 1 zero pad column); the kernel COLUMN is a +-16-byte shift of the A operand; the kernel ROW is the N dimension,
// N = 96 = (kernel row kh, c_out), so block kh of what the MMAs on INPUT tile r' produce belongs to OUTPUT tile
// r' - kh + 1 at the same lane.  What the second kernel did with that -- one accumulator per input tile, an epilogue
// item reading a block from each of three accumulators -- made a stage live for three items and serialised everything
// on a ring of five (profiles/r2_trunk_experiments.txt).  Here the ACCUMULATORS ARE PER OUTPUT TILE and the tensor
// core does the scatter: the outputs of a layer sit side by side in TMEM in DESCENDING tile order, 32 columns each,
//     O[r] at column 32 * (7 - r) of the layer's region,
// so the three blocks of input tile r' -- for O[r'+1], O[r'], O[r'-1] -- are 96 CONSECUTIVE columns starting at
// 32 * (6 - r'), and ONE N = 96 MMA with that D address accumulates all three where they belong (N = 64 at the top /
// bottom tile, whose third block would be conv2's zero padding).  Successive tile jobs overlap in two of their three
// blocks; MMAs of one thread execute in order, which is all the accumulation needs.  The first contribution to an
// output tile must not accumulate: the first MMA of a job is split into an N = 32 MMA that initialises the block of
// O[r'+1] and an N = 64 MMA that accumulates into the other two (7 MMAs per job, 6 for the first / last tile).
// Two regions of 8 x 32 columns alternate between the layers = all 512 TMEM columns; a region is reused two layers
// later, when the MMAs of the layer in between have already waited for every tile the old one produced.
// The epilogue item of an output tile is then what conv1's always was: one 32-column TMEM read, bias, ReLU, bf16, four
// 16-byte stores IN PLACE, one barrier arrival -- a third of the TMEM traffic of the other two kernels, no adds, no
// shuffles, no accumulator hand-back; and the MMA warp can run a whole layer ahead of it.
// One MMA warp (the order of the jobs is what makes the accumulation deterministic), three epilogue teams taking the
// output tiles round robin, two producer warps for conv1's im2col operand of the next group.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "umma.cuh"
#include "trunk3.cuh"

#define T3_NB 14                  // boards per group
#define T3_PITCH 9                // lanes per board in a tile
#define T3_TILES 8                // tiles per layer = board rows
#define T3_ROWS (T3_TILES * 128)
#define T3_MARGIN 8               // zero rows before / after a buffer (the +-1 row shifts of the first / last tile)
#define T3_PLANE ((T3_ROWS + 2 * T3_MARGIN) * 16)      // one 8-channel chunk plane: 16 640 B (= LBO of the A operand)
#define T3_BUF (4 * T3_PLANE)     // activation buffer: 66 560 B
#define T3_IM2COL (2 * T3_PLANE)  // conv1's A operand (K = 16: 9 taps + zeros): 33 280 B
#define T3_TEAMS 3
#define T3_EPI_THREADS (128 * T3_TEAMS)
#define T3_MMA_WARP (T3_EPI_THREADS / 32)
#define T3_MMA_WARPS 2            // issuing warp h accumulates the output tiles 4 h .. 4 h + 3 of every layer
#define T3_PRODUCERS 64
#define T3_THREADS (T3_EPI_THREADS + 32 * T3_MMA_WARPS + T3_PRODUCERS)
#define T3_HJOBS 18               // MMA tile jobs per group and issuing warp: 4 + 5 + 5 + 4
#define T3_ITEMS 26               // epilogue items per group = output tiles that are used: 8 + 8 + 6 + 4
// shared memory map
#define T3_SM_BUF(i) ((i) * T3_BUF)
#define T3_SM_IM2COL (2 * T3_BUF)
#define T3_SM_W (T3_SM_IM2COL + T3_IM2COL)             // conv2..4 reordered: [layer 3][kw 3][ci chunk 4][(kh, co) 96][8] bf16
#define T3_W_DC_BYTES (4 * 96 * 16)                    // one (layer, kernel column): 6 144 B
#define T3_SM_W1 (T3_SM_W + 9 * T3_W_DC_BYTES)         // conv1: [k chunk 2][co 32][8] bf16
#define T3_SM_BIAS (T3_SM_W1 + 1024)                   // float [4][32]
#define T3_SM_BARS (T3_SM_BIAS + 512)
#define TC_JOB 0                                       // [2][18] the MMAs of tile job j of issuing warp h have completed
#define TC_READY 36                                    // [2][8] output tile r of the group in buffer b is in shared memory
#define TC_IM2COL 52                                   // conv1's A operand of the next group is in place
#define TC_I2FREE 53                                   // conv1's MMAs of the current group have read it
#define TC_O4DONE 54                                   // conv4's accumulators of the group have been read (16 warps)
#define TC_COUNT 55
#define T3_SM_ITEMS (T3_SM_BARS + 8 * 56)              // uint32 [26]: layer, tile, last job, TMEM column of each epilogue item
#define T3_CJOBS 14                                    // jobs of conv2..conv4 per group and issuing warp
#define T3_SM_JOBS (T3_SM_ITEMS + 4 * 28)              // uint4 [2][14]: what the issuing thread needs to know about a job
#define T3_SM_TMEM_PTR (T3_SM_JOBS + 16 * 2 * T3_CJOBS)
#define T3_SM_TOTAL (T3_SM_TMEM_PTR + 16)
static_assert(T3_SM_TOTAL <= 232448, "shared memory per CTA on sm_100");
#define T3_WAIT_LIMIT (4000000000LL)

struct Trunk3Params {
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
__device__ __forceinline__ void t3_wait(uint32_t bar, uint32_t parity) {
    long long t0 = 0;
    for (;;) {
        uint32_t done;
        asm volatile(
            "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
            : "=r"(done) : "r"(bar), "r"(parity) : "memory");
        if (done) return;
        const long long t = clock64();
        if (t0 == 0) t0 = t;
        else if (t - t0 > T3_WAIT_LIMIT) __trap();
    }
}
// arrivals that do not wait for the barrier's state word to come back
__device__ __forceinline__ void t3_arrive(uint32_t bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}

// tcgen05.mma with K-major no-swizzle descriptors given by their low words (start address >> 4 | LBO >> 4 << 16; the
// high word -- SBO = 128 bytes, descriptor version 1 -- is the same for every operand of this kernel): moving an
// operand by d bytes adds d >> 4 to the low word (shared memory addresses stay below 2^18), which is all the issuing
// thread computes per instruction
__device__ __forceinline__ void t3_mma(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
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
__device__ __forceinline__ uint32_t t3_desc_lo(uint32_t smem_addr, uint32_t lbo) {
    return ((smem_addr >> 4) & 0x3FFFu) | ((lbo >> 4) << 16);
}

// Per layer L = 0..3 (conv1..conv4): first MMA tile, number of MMA tiles, job offset inside a group; first / last
// output tile that is used.
__device__ __forceinline__ int t3_mlo(int L) { return L == 3 ? 1 : 0; }
__device__ __forceinline__ int t3_mcnt(int L) { return L == 3 ? 6 : 8; }
__device__ __forceinline__ int t3_olo(int L) { return L < 2 ? 0 : L - 1; }          // conv3: 1, conv4: 2
__device__ __forceinline__ int t3_ohi(int L) { return L < 2 ? 7 : 8 - L; }          // conv3: 6, conv4: 5
// Issuing warp h owns the accumulators of the output tiles 4 h .. 4 h + 3: the output tiles it produces in layer L, the
// input tiles that contribute to them (its tile jobs: the two tiles next to the border are jobs of both warps, each
// with the kernel rows that land on its own side), its first job of layer L among its 18 jobs of a group.
__device__ __forceinline__ int t3_holo(int h, int L) { return max(t3_olo(L), 4 * h); }
__device__ __forceinline__ int t3_hohi(int h, int L) { return min(t3_ohi(L), 4 * h + 3); }
__device__ __forceinline__ int t3_tlo(int h, int L) { return L == 0 ? 4 * h : max(t3_mlo(L), t3_holo(h, L) - 1); }
__device__ __forceinline__ int t3_thi(int h, int L) { return L == 0 ? 4 * h + 3 : min(t3_mlo(L) + t3_mcnt(L) - 1, t3_hohi(h, L) + 1); }
__device__ __forceinline__ int t3_jloc(int L) { return L == 0 ? 0 : 5 * L - 1; }       // 0, 4, 9, 14
// TMEM column of the accumulator of output tile r of layer L
__device__ __forceinline__ uint32_t t3_ocol(int L, int r) { return (uint32_t)((L & 1) * 256 + 32 * (7 - r)); }

// conv1's A operand for one row (tile r, lane): the 9 taps of player * grid around cell (r, c) as bf16 (+1 own, -1 opp,
// 0 empty / off-board), then zeros up to K = 16; c < 0 (pad column, unused lane, missing board) = a zero row.
__device__ __forceinline__ void t3_im2col_row(uint8_t* im2col, int row, int r, int c, uint64_t own, uint64_t opp) {
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
    uint8_t* dst = im2col + (T3_MARGIN + row) * 16;
    *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
    *reinterpret_cast<uint4*>(dst + T3_PLANE) = make_uint4(w[4], 0u, 0u, 0u);
}

__global__ void __launch_bounds__(T3_THREADS, 1) k_net_trunk3(Trunk3Params p) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_groups = (p.G + T3_NB - 1) / T3_NB;
    // debug_layer 200: cycle stamps instead of activations, int64 [n_ctas][16] (tools/trunk2_probe.py)
    // (201: the same without the MMA warps' waits for the epilogue, 202: also with N = 32 MMAs only -- wrong results, the
    // cost of the MMA stream alone)
    const int hack = p.debug_layer >= 200 ? p.debug_layer - 200 : 0;
    long long* stamps = (p.debug && p.debug_layer >= 200) ? reinterpret_cast<long long*>(p.debug) + (size_t)blockIdx.x * 16 : nullptr;
    const long long t_entry = stamps ? clock64() : 0;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bars = smem_u32(smem + T3_SM_BARS);
    uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(smem + T3_SM_TMEM_PTR);

    // ---- one-time setup: barriers, zeroed buffers (margins, unused lanes), weights (reordered from the blob), TMEM
    if (tid < TC_COUNT) {
        // arrivals are per WARP (lane 0 after __syncwarp): 32 lanes arriving on one barrier word serialise
        uint32_t count = 1u;                                       // job[][]: one tcgen05.commit
        if (tid >= TC_READY && tid < TC_IM2COL) count = 4u;        // ready[][]: the team that wrote the tile
        else if (tid == TC_IM2COL) count = T3_PRODUCERS / 32;
        else if (tid == TC_I2FREE) count = T3_MMA_WARPS;           // each issuing warp commits after its conv1 jobs
        else if (tid == TC_O4DONE) count = 4u * 4u;                // conv4's four output tiles, 4 warps each
        mbar_init(bars + 8 * tid, count);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (tid >= 64 && tid < 64 + T3_ITEMS) {
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
        else j = h ? min(r + 1, t3_thi(1, L)) - t3_tlo(1, L) : t3_thi(0, L) - max(r - 1, t3_tlo(0, L));
        const int job = h * T3_HJOBS + t3_jloc(L) + j;
        int job2 = job;
        if (L > 0 && (r == 3 || r == 4)) job2 = (1 - h) * T3_HJOBS + t3_jloc(L);      // tile 3 / 4 is the other warp's first job
        reinterpret_cast<uint32_t*>(smem + T3_SM_ITEMS)[idx] =
            (uint32_t)L | ((uint32_t)r << 4) | ((uint32_t)job << 8) | ((uint32_t)job2 << 14) | (t3_ocol(L, r) << 20);
    }
    if (tid >= 96 && tid < 96 + 2 * T3_CJOBS) {
        // job jj of issuing warp h in conv2..conv4 (its step j of layer L): everything about it that does not depend on the
        // group.  Warp 0 goes down from the border (tiles 4, 3, ..), warp 1 up (3, 4, ..).
        const int h = (tid - 96) / T3_CJOBS, jj = (tid - 96) % T3_CJOBS;
        const int L = jj < 5 ? 1 : (jj < 10 ? 2 : 3), j = jj - (5 * L - 5);
        const int tlo = t3_tlo(h, L), thi = t3_thi(h, L), olo = t3_holo(h, L), ohi = t3_hohi(h, L);
        const int r = h ? tlo + j : thi - j;
        // kernel rows whose output tile r - kh + 1 this warp produces: kh = klo .. khi, 32 accumulator columns each, output
        // tiles descending = columns ascending from that of tile r - klo + 1
        const int klo = max(0, r + 1 - ohi), khi = min(2, r + 1 - olo), nblk = khi - klo + 1;
        const uint32_t d_col = (uint32_t)((L & 1) * 256 + 32 * (6 - r + klo));
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
        const uint32_t wl = sbase + T3_SM_W + (uint32_t)((L - 1) * 3) * T3_W_DC_BYTES + (uint32_t)(klo * 32 * 16);
        reinterpret_cast<uint4*>(smem + T3_SM_JOBS)[tid - 96] =
            make_uint4((uint32_t)(T3_MARGIN + r * 128 - 1),            // A: the tile's rows shifted by -1 column, in 16-byte rows
                       t3_desc_lo(wl, 96u * 16u), z,
                       (uint32_t)r | (((r >= 1 && r <= 6) ? 3u : 2u) << 3) | ((uint32_t)(L - 1) << 5));
    }
    // the rows nothing ever writes: the margins of the activation buffers (the row before tile 0 is lane 0's left
    // neighbour = zero padding) and the two unused lanes of every tile (lane 127 is the left neighbour of the next
    // tile's lane 0); the pad lanes are written with every conv1 output
    if (tid < 2 * 4 * 2 * T3_MARGIN) {
        const int bufi = tid >> 6, pl = (tid >> 4) & 3, side = (tid >> 3) & 1, row = tid & 7;
        *reinterpret_cast<uint4*>(smem + T3_SM_BUF(bufi) + pl * T3_PLANE + (side ? T3_MARGIN + T3_ROWS + row : row) * 16) =
            make_uint4(0u, 0u, 0u, 0u);
    } else if (tid < 2 * 4 * 2 * T3_MARGIN + 2 * 4 * T3_TILES * 2) {
        const int u = tid - 2 * 4 * 2 * T3_MARGIN;
        const int bufi = u >> 6, pl = (u >> 4) & 3, tile = (u >> 1) & 7, ln = T3_NB * T3_PITCH + (u & 1);
        *reinterpret_cast<uint4*>(smem + T3_SM_BUF(bufi) + pl * T3_PLANE + (T3_MARGIN + tile * 128 + ln) * 16) =
            make_uint4(0u, 0u, 0u, 0u);
    }
    {   // blob: [layer][kh][ci chunk][(kw, co)][8] -> shared memory [layer][kw][ci chunk][(kh, co)][8], 16-byte units
        const uint4* src = reinterpret_cast<const uint4*>(p.wblob);
        uint4* dst = reinterpret_cast<uint4*>(smem + T3_SM_W);
        constexpr int UNITS = 3 * 3 * 4 * 96, PER = (UNITS + T3_THREADS - 1) / T3_THREADS;
        uint4 v[PER];                                              // all loads in flight before the first store
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int u = tid + i * T3_THREADS;
            v[i] = u < UNITS ? src[u] : make_uint4(0u, 0u, 0u, 0u);
        }
#pragma unroll
        for (int i = 0; i < PER; ++i) {
            const int u = tid + i * T3_THREADS;
            if (u < UNITS) {
                const int co = u & 31, kw = (u >> 5) % 3, cc = (u / 96) & 3, kh = (u / 384) % 3, li = u / 1152;
                dst[(((li * 3 + kw) * 4 + cc) * 3 + kh) * 32 + co] = v[i];
            }
        }
        const uint4* tail = src + 3 * 3 * 4 * 96;                  // conv1 tile (1024 B), then the four bias rows (512 B)
        for (int u = tid; u < (1024 + 512) / 16; u += T3_THREADS) reinterpret_cast<uint4*>(smem + T3_SM_W1)[u] = tail[u];
    }
    if (warp == T3_MMA_WARP) {
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

    if (warp >= T3_MMA_WARP + T3_MMA_WARPS) {
        // ================================================================== producers: conv1's A operand (9-tap im2col
        // of player * grid, K = 16) of every group of this CTA.  Thread = lanes ptid and ptid + 64 of all 8 tiles.
        const int ptid = tid - (T3_EPI_THREADS + 32 * T3_MMA_WARPS);
        uint8_t* im2col = smem + T3_SM_IM2COL;
        // (the first group's is built by the epilogue teams, which have nothing else to do yet)
        uint32_t k = 1;
        for (int grp = blockIdx.x + gridDim.x; grp < n_groups; grp += gridDim.x, ++k) {
            const int board0 = grp * T3_NB, n_boards = min(T3_NB, p.G - board0);
            uint64_t own[2], opp[2];
            int col[2];
#pragma unroll
            for (int j = 0; j < 2; ++j) {
                const int tl = ptid + 64 * j, b = tl / T3_PITCH;
                col[j] = tl - b * T3_PITCH;
                const bool ok = b < n_boards && tl < T3_NB * T3_PITCH && col[j] < 8;
                own[j] = ok ? p.own[board0 + b] : 0ULL;
                opp[j] = ok ? p.opp[board0 + b] : 0ULL;
                if (!ok) col[j] = -1;
            }
            t3_wait(bars + 8 * TC_I2FREE, (k - 1u) & 1u);                 // conv1 of the previous group has read the buffer
#pragma unroll
            for (int j = 0; j < 2; ++j) {
#pragma unroll 2
                for (int r = 0; r < T3_TILES; ++r) {
                    t3_im2col_row(im2col, r * 128 + ptid + 64 * j, r, col[j], own[j], opp[j]);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncwarp();
            if (lane == 0) mbar_arrive(bars + 8 * TC_IM2COL);
        }
    } else if (warp >= T3_MMA_WARP) {
        // ================================================================== MMA issue: warp h issues the tile jobs that
        // accumulate into the output tiles 4 h .. 4 h + 3, in program order, by one elected thread of a converged warp
        // (a single lane in divergent code costs an election and seven register broadcasts per MMA: measured slower).
        // The issuing thread is the serial resource of this kernel (a tcgen05.mma costs it ~65 cycles; two of them
        // keep the tensor pipe busy), so what a job needs beyond the group's buffer is in the table built at set-up.
        const int h = warp - T3_MMA_WARP;
        const uint32_t jobs = bars + 8 * (TC_JOB + h * T3_HJOBS);
        const uint32_t idesc32 = umma_idesc_bf16_m128(32u);
        constexpr uint32_t IDESC_BLK = (32u >> 3) << 17;          // one more kernel row = 32 more columns
        const uint4* table = reinterpret_cast<const uint4*>(smem + T3_SM_JOBS) + h * T3_CJOBS;
        const uint32_t w1_lo = t3_desc_lo(sbase + T3_SM_W1, 512u);
        uint32_t k = 0;                               // group counter of this CTA
        long long w_ready = 0, w_o4 = 0, t_issue = 0, w_im2col = 0;
        for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++k) {
            const uint32_t buf = k & 1u, use = k >> 1;
            const long long c_i = stamps ? clock64() : 0;
            t3_wait(bars + 8 * TC_IM2COL, k & 1u);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const long long c_j = stamps ? clock64() : 0;
            if (stamps) w_im2col += c_j - c_i;
            // conv1: one K = 16 MMA per tile into its own accumulator (region 0: conv3's of the previous group have been
            // read -- its conv4 jobs waited for every conv3 output tile)
            if (elect_one()) {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int r = h ? 4 + i : 3 - i;
                    t3_mma(tmem_base + t3_ocol(0, r), t3_desc_lo(sbase + T3_SM_IM2COL + (uint32_t)(T3_MARGIN + r * 128) * 16u, T3_PLANE), w1_lo,
                           idesc32, 0u);
                    umma_commit(jobs + 8 * i);
                }
                umma_commit(bars + 8 * TC_I2FREE);    // conv1 has read the im2col buffer: the producers may build the next group's
            }
            __syncwarp();
            if (stamps) t_issue += clock64() - c_j;
            if (stamps && lane == 0 && k == 0 && h == 0) stamps[8] = clock64() - t_entry;
            const uint32_t abuf_lo = t3_desc_lo(sbase + T3_SM_BUF(buf), T3_PLANE);
            const uint32_t ready = bars + 8 * (TC_READY + buf * 8);
#pragma unroll 1
            for (int jj = 0; jj < T3_CJOBS; ++jj) {
                const uint4 e = table[jj];
                const long long c0 = stamps ? clock64() : 0;
                {   // the input: layer L - 1's output tile r = phase (use * P + L - 1) of ready[buf][r], P = layers that write tile r
                    const uint32_t r = e.w & 7u, P = (e.w >> 3) & 3u, Lm1 = (e.w >> 5) & 3u;
                    if (hack == 0) t3_wait(ready + 8 * r, (use * P + Lm1) & 1u);
                }
                const long long c1 = stamps ? clock64() : 0;
                // region 1 still holds conv4's accumulators of the previous group until its four epilogue items have read them
                if (((e.z >> 21) & 1u) && k > 0 && hack == 0) t3_wait(bars + 8 * TC_O4DONE, (k - 1u) & 1u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const long long c2 = stamps ? clock64() : 0;
                const uint32_t d_tmem = tmem_base + (e.z & 511u);
                const uint32_t idesc = idesc32 + (hack == 2 ? 0u : (((e.z >> 9) & 3u) - 1u) * IDESC_BLK);
                const uint32_t a_lo = abuf_lo + e.x, b_lo = e.y;
                if (elect_one()) {
                {   // kernel column 0, channels 0..15: whole, or the block that must not accumulate and then the rest
                    const uint32_t pos = (e.z >> 11) & 3u;
                    t3_mma(d_tmem + 32u * pos, a_lo, b_lo + 32u * pos, idesc32 + (((e.z >> 13) & 3u) - 1u) * IDESC_BLK, (e.z >> 15) & 1u);
                    if ((e.z >> 16) & 1u) {
                        const uint32_t pos2 = (e.z >> 17) & 3u;
                        t3_mma(d_tmem + 32u * pos2, a_lo, b_lo + 32u * pos2, idesc32 + (((e.z >> 19) & 3u) - 1u) * IDESC_BLK, 1u);
                    }
                }
#pragma unroll
                for (int i = 1; i < 6; ++i) {
                    const int kw = i >> 1, kk = i & 1;
                    t3_mma(d_tmem, a_lo + (uint32_t)(kw + kk * (2 * T3_PLANE >> 4)),
                           b_lo + (uint32_t)(kw * (T3_W_DC_BYTES >> 4) + kk * (2 * 96 * 16 >> 4)), idesc, 1u);
                }
                umma_commit(jobs + 8 * (4 + jj));
                }
                __syncwarp();
                if (stamps) { const long long c3 = clock64(); w_ready += c1 - c0; w_o4 += c2 - c1; t_issue += c3 - c2; }
                if (stamps && lane == 0 && k == 0 && h == 0 && (jj == 4 || jj == 9)) stamps[jj == 4 ? 9 : 10] = clock64() - t_entry;   // group 0: conv2 / conv3 issued
            }
            if (stamps && lane == 0 && k == 0 && h == 0) stamps[3] = clock64() - t_entry;
        }
        if (stamps && lane == 0 && h == 0) { stamps[5] = w_ready; stamps[6] = w_o4; stamps[7] = t_issue; stamps[12] = w_im2col; }
    } else {
        // ================================================================== epilogue team `team`: items team, team + 3, ..
        const int team = warp >> 2, quarter = warp & 3;
        const int tl = quarter * 32 + lane;                     // TMEM lane = lane of the tile
        const int b = tl / T3_PITCH, c = tl - b * T3_PITCH;     // board of the group, column (8 = pad)
        const float* bias_all = reinterpret_cast<const float*>(smem + T3_SM_BIAS);
        const uint32_t t_quarter = tmem_base + ((uint32_t)(quarter * 32) << 16);
        {   // conv1's A operand of the CTA's first group, by all epilogue threads (three rows each)
            const int board0 = blockIdx.x * T3_NB, n_boards = min(T3_NB, p.G - board0);
            uint8_t* im2col = smem + T3_SM_IM2COL;
            for (int row = tid; row < T3_ROWS; row += T3_EPI_THREADS) {
                const int r = row >> 7, l = row & 127, bb = l / T3_PITCH, cc = l - bb * T3_PITCH;
                const bool ok = bb < n_boards && l < T3_NB * T3_PITCH && cc < 8;
                t3_im2col_row(im2col, row, r, ok ? cc : -1, ok ? p.own[board0 + bb] : 0ULL, ok ? p.opp[board0 + bb] : 0ULL);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            asm volatile("bar.sync 1, %0;" ::"r"(T3_EPI_THREADS) : "memory");
            if (tid < T3_PRODUCERS / 32) mbar_arrive(bars + 8 * TC_IM2COL);
            if (stamps && tid == 0) stamps[1] = clock64() - t_entry;
        }
        uint32_t k = 0;
        long long e_wait = 0, e_work = 0;
        for (int grp = blockIdx.x; grp < n_groups; grp += gridDim.x, ++k) {
            const uint32_t buf = k & 1u;
            const int board0 = grp * T3_NB, n_boards = min(T3_NB, p.G - board0);
            const bool used = tl < T3_NB * T3_PITCH;                // not one of the tile's two spare lanes
            const bool cell = b < n_boards && used && c < 8;        // a real cell of a real board
            // 26 items per group, three teams: the team of an item rotates with the group (26 = 2 mod 3)
#pragma unroll 1
            for (int idx = (team + (int)(k % 3u)) % 3; idx < T3_ITEMS; idx += T3_TEAMS) {
                const uint32_t it = reinterpret_cast<const uint32_t*>(smem + T3_SM_ITEMS)[idx];
                const int L = (int)(it & 3u), r = (int)((it >> 4) & 15u);
                const long long f0 = stamps ? clock64() : 0;
                t3_wait(bars + 8 * (TC_JOB + ((it >> 8) & 63u)), k & 1u);
                if (((it >> 14) & 63u) != ((it >> 8) & 63u)) t3_wait(bars + 8 * (TC_JOB + ((it >> 14) & 63u)), k & 1u);
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
                        uint8_t* d = smem + T3_SM_BUF(buf) + (size_t)(T3_MARGIN + r * 128 + tl) * 16;
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(d + ch * T3_PLANE) = make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]);
                    }
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) t3_arrive(bars + 8 * (TC_READY + buf * 8 + r));
                    if (p.debug && p.debug_layer == L + 1) {
#pragma unroll
                        for (int ch = 0; ch < 4; ++ch)
                            *reinterpret_cast<uint4*>(p.debug + ((size_t)grp * T3_ROWS + r * 128 + tl) * 32 + ch * 8) =
                                used ? make_uint4(out[ch * 4], out[ch * 4 + 1], out[ch * 4 + 2], out[ch * 4 + 3]) : make_uint4(0u, 0u, 0u, 0u);
                    }
                } else {
                    // the accumulator has been read (tmem_ld32 waits for the data): conv2 of the next group may overwrite it
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    __syncwarp();
                    if (lane == 0) t3_arrive(bars + 8 * TC_O4DONE);
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
    if (warp == T3_MMA_WARP)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(512));
    if (stamps && tid == 0) stamps[1 + 10] = clock64() - t_entry;      // [11] exit
}

// AZ_B200_TRUNK_V3=1 selects this kernel for every 8 x 8 batch of AZ_TRUNK3_MIN_G positions and more; bit 2 of
// az_net_trunk's `features_packed` selects it for one call.
bool az_trunk3_enabled() {
    static const bool on = getenv("AZ_B200_TRUNK_V3") != nullptr;
    return on;
}

int az_trunk3_launch(const uint64_t* own, const uint64_t* opp, int64_t G, const void* wblob, void* features,
                     int features_packed, void* debug, int debug_layer, int n_sms, int late_trigger, void* stream) {
    static bool attr_set = false;
    if (!attr_set) {
        AZ_CUDA_TRY(cudaFuncSetAttribute(k_net_trunk3, cudaFuncAttributeMaxDynamicSharedMemorySize, T3_SM_TOTAL));
        attr_set = true;
    }
    Trunk3Params p;
    p.own = own; p.opp = opp; p.wblob = (const uint8_t*)wblob; p.features = (__nv_bfloat16*)features;
    p.debug = (__nv_bfloat16*)debug; p.G = (int)G; p.debug_layer = debug ? debug_layer : 0;
    p.features_packed = features_packed ? 1 : 0; p.late_trigger = late_trigger;
    const int n_groups = (int)((G + T3_NB - 1) / T3_NB);
    const int grid = n_groups < n_sms ? n_groups : n_sms;      // persistent: one CTA per SM
    cudaError_t e = az_launch_pdl(k_net_trunk3, dim3(grid), dim3(T3_THREADS), T3_SM_TOTAL, (cudaStream_t)stream, p);
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_net_trunk3");
    AZ_CHECK_LAUNCH("k_net_trunk3");
    return AZ_OK;
}

