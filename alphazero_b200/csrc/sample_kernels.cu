// Training-sample kernels + their C-ABI entry points (include/az_b200.h): the symmetry augmentation of the self-play
// memory (alphazero/trainer.py:80-118, 275-284 with the games' reflect/rotate_neural_output, othello.py:414-450,
// connect4.py:437-445) and the mini-batch assembly of AlphaZeroTrainer._batch_generator (trainer.py:288-322).
// sm_100a only.
//
// Both are HBM-bound byte movers: a sample is 16 B of bitboards + action_size floats of pi + 10 B of indices; the
// augmentation reads each sample once and writes 1 + 7 (Othello) or 1 + 1 (Connect4) samples, the batch assembly
// reads one sample and writes rows*cols + action_size + 1 floats.  One warp owns one sample: lane j moves cells j
// and j + 32 of the board (the transformed bitboard is assembled by two ballots) and entries j, j + 32, j + 64 of pi,
// so every global access of a warp is one contiguous segment.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/az_b200.h"
#include "common.cuh"

#define AZ_SAMPLE_WARPS 8            // warps per CTA
#define AZ_SCAN_BLOCK 1024           // samples per block of the eligibility scan

// Source cell of output cell (i, j) under transformation `code` (0 identity, 1 reflect_h, 2/4/6 rot90/180/270,
// 3/5/7 the same rotations applied to the reflected board): out[i][j] = in[si][sj].
//   reflect_h: np.flip(axis=1)            -> in[i][cols-1-j]                       (trainer.py:91)
//   rot90 k  : np.rot90(m, k), square n   -> k=1 m[j][n-1-i], k=2 m[n-1-i][n-1-j], k=3 m[n-1-j][i]   (trainer.py:111)
__device__ __forceinline__ void az_sym_source(int code, int i, int j, int rows, int cols, int& si, int& sj) {
    const int n = cols;
    switch (code) {
        case 0: si = i; sj = j; break;
        case 1: si = i; sj = cols - 1 - j; break;
        case 2: si = j; sj = n - 1 - i; break;
        case 3: si = j; sj = i; break;                          // reflect, then rot90: m[j][n-1-i] with m = refl
        case 4: si = n - 1 - i; sj = n - 1 - j; break;
        case 5: si = n - 1 - i; sj = j; break;
        case 6: si = n - 1 - j; sj = i; break;
        default: si = n - 1 - j; sj = n - 1 - i; break;         // 7
    }
    (void)rows;
}

// ------------------------------------------------------------------------------------------------ eligibility scan

// Pass 1: local exclusive rank of every sample with move_idx >= min_move_idx inside its block + the block's count.
__global__ void __launch_bounds__(AZ_SCAN_BLOCK) k_augment_rank(const int32_t* __restrict__ move_idx, int64_t S,
                                                               int min_move_idx, int32_t* __restrict__ rank,
                                                               int32_t* __restrict__ block_sums) {
    __shared__ int warp_tot[32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int64_t s = (int64_t)blockIdx.x * AZ_SCAN_BLOCK + threadIdx.x;
    const bool el = s < S && move_idx[s] >= min_move_idx;
    const unsigned b = __ballot_sync(0xffffffffu, el);
    const int in_warp = __popc(b & ((1u << lane) - 1u));
    if (lane == 0) warp_tot[w] = __popc(b);
    __syncthreads();
    if (w == 0) {
        int v = warp_tot[lane], x = v;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        warp_tot[lane] = x - v;                                  // exclusive prefix of the warp totals
        if (lane == 31) block_sums[blockIdx.x] = x;
    }
    __syncthreads();
    if (s < S) rank[s] = el ? warp_tot[w] + in_warp : -1;
}

// Pass 2: exclusive prefix of the block counts in place (one CTA), grand total in block_sums[n_blocks] and *total.
__global__ void __launch_bounds__(1024) k_augment_block_offsets(int32_t* __restrict__ block_sums, int n_blocks,
                                                                int64_t* __restrict__ total) {
    __shared__ int warp_tot[32];
    __shared__ int carry;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    if (threadIdx.x == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < n_blocks; base += 1024) {
        const int i = base + threadIdx.x;
        const int v = i < n_blocks ? block_sums[i] : 0;
        int x = v;
        #pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, d);
            if (lane >= d) x += y;
        }
        if (lane == 31) warp_tot[w] = x;
        __syncthreads();
        if (w == 0) {
            int t = warp_tot[lane], u = t;
            #pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, u, d);
                if (lane >= d) u += y;
            }
            warp_tot[lane] = u - t;
        }
        __syncthreads();
        const int excl = carry + warp_tot[w] + x - v;
        if (i < n_blocks) block_sums[i] = excl;
        __syncthreads();
        if (threadIdx.x == 1023) carry = excl + v;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        block_sums[n_blocks] = carry;
        *total = carry;
    }
}

// ------------------------------------------------------------------------------------------------ augmentation

struct AugmentParams {
    const uint64_t* own; const uint64_t* opp; const float* pi; const int8_t* z;
    const int32_t* episode_idx; const int32_t* move_idx;
    const int32_t* rank; const int32_t* block_offsets;
    uint64_t* out_own; uint64_t* out_opp; float* out_pi; int8_t* out_z;
    int32_t* out_episode; int32_t* out_move; uint8_t* out_transf;
    int64_t S;
    int rows, cols, action_size, n_twins, board_actions;   // board_actions: entries of pi that are cells (0 = columns)
};

__global__ void __launch_bounds__(AZ_SAMPLE_WARPS * 32) k_augment_apply(AugmentParams p) {
    __shared__ float s_pi[AZ_SAMPLE_WARPS][96];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int A = p.action_size;
    for (int64_t s = (int64_t)blockIdx.x * AZ_SAMPLE_WARPS + w; s < p.S; s += (int64_t)gridDim.x * AZ_SAMPLE_WARPS) {
        const uint64_t own = p.own[s], opp = p.opp[s];
        const int8_t z = p.z[s];
        const int32_t ep = p.episode_idx[s], mv = p.move_idx[s];
        const int32_t r = p.rank[s];
        __syncwarp();
        for (int a = lane; a < A; a += 32) {
            const float v = p.pi[s * A + a];
            s_pi[w][a] = v;
            p.out_pi[s * A + a] = v;                             // the original sample stays at index s
        }
        if (lane == 0) {
            p.out_own[s] = own; p.out_opp[s] = opp; p.out_z[s] = z;
            p.out_episode[s] = ep; p.out_move[s] = mv; p.out_transf[s] = 0;
        }
        __syncwarp();
        if (r < 0) continue;
        const int64_t first = p.S + ((int64_t)p.block_offsets[s / AZ_SCAN_BLOCK] + r) * p.n_twins;
        for (int k = 0; k < p.n_twins; ++k) {
            const int code = k + 1;
            const int64_t o = first + k;
            // board: lane handles cells `lane` and `lane + 32` of the 8-wide canvas
            uint64_t t_own = 0, t_opp = 0;
            #pragma unroll
            for (int h = 0; h < 2; ++h) {
                const int cell = lane + 32 * h, i = cell >> 3, j = cell & 7;
                bool b_own = false, b_opp = false;
                if (i < p.rows && j < p.cols) {
                    int si, sj;
                    az_sym_source(code, i, j, p.rows, p.cols, si, sj);
                    const int sb = si * 8 + sj;
                    b_own = (own >> sb) & 1ULL;
                    b_opp = (opp >> sb) & 1ULL;
                }
                t_own |= (uint64_t)__ballot_sync(0xffffffffu, b_own) << (32 * h);
                t_opp |= (uint64_t)__ballot_sync(0xffffffffu, b_opp) << (32 * h);
            }
            // pi: board entries follow the cells (index i*cols + j), the tail (pass) is kept (othello.py:424-428);
            // a column policy is reversed whole (connect4.py:445)
            for (int a = lane; a < A; a += 32) {
                float v;
                if (p.board_actions == 0) v = s_pi[w][A - 1 - a];
                else if (a >= p.board_actions) v = s_pi[w][a];
                else {
                    int si, sj;
                    az_sym_source(code, a / p.cols, a % p.cols, p.rows, p.cols, si, sj);
                    v = s_pi[w][si * p.cols + sj];
                }
                p.out_pi[o * A + a] = v;
            }
            if (lane == 0) {
                p.out_own[o] = t_own; p.out_opp[o] = t_opp; p.out_z[o] = z;
                p.out_episode[o] = ep; p.out_move[o] = mv; p.out_transf[o] = (uint8_t)code;
            }
        }
    }
}

// ------------------------------------------------------------------------------------------------ mini-batches

// input[b] = float grid of sample index[b] (own = +1, opp = -1), pi_out[b] = its pi, z_out[b] = its outcome
// (trainer.py:309-318: the three torch tensors of one batch).
__global__ void __launch_bounds__(AZ_SAMPLE_WARPS * 32) k_samples_batch(
    const uint64_t* __restrict__ own, const uint64_t* __restrict__ opp, const float* __restrict__ pi,
    const int8_t* __restrict__ z, const int64_t* __restrict__ index, int64_t S, int64_t B, int rows, int cols, int A,
    float* __restrict__ input, float* __restrict__ pi_out, float* __restrict__ z_out, int32_t* __restrict__ bad) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int cells = rows * cols;
    for (int64_t b = (int64_t)blockIdx.x * AZ_SAMPLE_WARPS + w; b < B; b += (int64_t)gridDim.x * AZ_SAMPLE_WARPS) {
        const int64_t s = index[b];
        if (s < 0 || s >= S) {                                   // refuse, never read out of bounds
            if (lane == 0) atomicExch(bad, 1);
            continue;
        }
        const uint64_t o = own[s], q = opp[s];
        for (int c = lane; c < cells; c += 32) {
            const int bit = (c / cols) * 8 + (c % cols);
            input[b * cells + c] = (float)((int)((o >> bit) & 1ULL) - (int)((q >> bit) & 1ULL));
        }
        for (int a = lane; a < A; a += 32) pi_out[b * A + a] = pi[s * A + a];
        if (lane == 0) z_out[b] = (float)z[s];
    }
}

// ------------------------------------------------------------------------------------------------ C ABI

static int az_sample_grid(int64_t items) {
    int64_t need = (items + AZ_SAMPLE_WARPS - 1) / AZ_SAMPLE_WARPS;
    int64_t cap = (int64_t)AZ_NUM_SMS_B200 * 8;
    if (need < 1) need = 1;
    return (int)(need < cap ? need : cap);
}

static bool az_valid_sample_game(int game, int rows, int cols, int action_size) {
    if (game == AZ_GAME_OTHELLO) return az_valid_n(cols) && rows == cols && action_size == cols * cols + 1;
    if (game == AZ_GAME_CONNECT4)
        return cols >= 4 && cols <= 8 && rows >= 4 && rows <= 8 && action_size == cols;
    if (game == AZ_GAME_TICTACTOE) return rows == 3 && cols == 3 && action_size == 9;
    return false;
}

extern "C" int64_t az_augment_scratch_bytes(int64_t S) {
    if (S < 0) return -1;
    const int64_t n_blocks = (S + AZ_SCAN_BLOCK - 1) / AZ_SCAN_BLOCK;
    return (S + n_blocks + 1) * (int64_t)sizeof(int32_t) + 16;
}

extern "C" int az_augment_twins(int game) {
    return (game == AZ_GAME_OTHELLO || game == AZ_GAME_TICTACTOE) ? 7 : game == AZ_GAME_CONNECT4 ? 1 : AZ_ERR_ARG;
}

extern "C" int az_augment_plan(const int32_t* move_idx, int64_t S, int min_move_idx, void* scratch, int64_t* total,
                               void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (S < 0 || S > (1LL << 30) || !scratch || !total || (S > 0 && !move_idx)) return AZ_ERR_ARG;
    const int n_blocks = (int)((S + AZ_SCAN_BLOCK - 1) / AZ_SCAN_BLOCK);
    int32_t* rank = (int32_t*)scratch;
    int32_t* block_sums = rank + S;
    if (n_blocks > 0) {
        k_augment_rank<<<n_blocks, AZ_SCAN_BLOCK, 0, stream>>>(move_idx, S, min_move_idx, rank, block_sums);
        AZ_CHECK_LAUNCH("k_augment_rank");
    }
    k_augment_block_offsets<<<1, 1024, 0, stream>>>(block_sums, n_blocks, total);
    AZ_CHECK_LAUNCH("k_augment_block_offsets");
    return AZ_OK;
}

extern "C" int az_augment_apply(const uint64_t* own, const uint64_t* opp, const float* pi, const int8_t* z,
                                const int32_t* episode_idx, const int32_t* move_idx, int64_t S, int game, int rows,
                                int cols, int action_size, const void* scratch, uint64_t* out_own, uint64_t* out_opp,
                                float* out_pi, int8_t* out_z, int32_t* out_episode, int32_t* out_move,
                                uint8_t* out_transf, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (S < 0 || !az_valid_sample_game(game, rows, cols, action_size) || !scratch) return AZ_ERR_ARG;
    if (S == 0) return AZ_OK;
    if (!own || !opp || !pi || !z || !episode_idx || !move_idx || !out_own || !out_opp || !out_pi || !out_z ||
        !out_episode || !out_move || !out_transf)
        return AZ_ERR_ARG;
    AugmentParams p;
    p.own = own; p.opp = opp; p.pi = pi; p.z = z; p.episode_idx = episode_idx; p.move_idx = move_idx;
    p.rank = (const int32_t*)scratch; p.block_offsets = p.rank + S;
    p.out_own = out_own; p.out_opp = out_opp; p.out_pi = out_pi; p.out_z = out_z;
    p.out_episode = out_episode; p.out_move = out_move; p.out_transf = out_transf;
    p.S = S; p.rows = rows; p.cols = cols; p.action_size = action_size;
    p.n_twins = az_augment_twins(game);
    p.board_actions = game == AZ_GAME_CONNECT4 ? 0 : cols * cols;       // TicTacToe: all 9 entries are cells
    k_augment_apply<<<az_sample_grid(S), AZ_SAMPLE_WARPS * 32, 0, stream>>>(p);
    AZ_CHECK_LAUNCH("k_augment_apply");
    return AZ_OK;
}

extern "C" int az_samples_batch(const uint64_t* own, const uint64_t* opp, const float* pi, const int8_t* z,
                                const int64_t* index, int64_t S, int64_t B, int rows, int cols, int action_size,
                                float* input, float* pi_out, float* z_out, int32_t* bad_index, void* stream_) {
    cudaStream_t stream = (cudaStream_t)stream_;
    if (S < 0 || B < 0 || rows < 1 || rows > 8 || cols < 1 || cols > 8 || action_size < 1 || action_size > 96 ||
        !bad_index)
        return AZ_ERR_ARG;
    if (B == 0) return AZ_OK;
    if (!own || !opp || !pi || !z || !index || !input || !pi_out || !z_out) return AZ_ERR_ARG;
    k_samples_batch<<<az_sample_grid(B), AZ_SAMPLE_WARPS * 32, 0, stream>>>(own, opp, pi, z, index, S, B, rows, cols,
                                                                          action_size, input, pi_out, z_out,
                                                                          bad_index);
    AZ_CHECK_LAUNCH("k_samples_batch");
    return AZ_OK;
}

// ------------------------------------------------------------------------------------------------ history -> samples

// The self-play history [T][G] (one row per ply, written by az_tree_move) as the training memory the reference builds
// (trainer.py:241-265): samples ordered by episode, then move; z = winner * player (Sample.normalize, trainer.py:74-78).
// A game is live for a prefix of the plies, so sample (g, t) sits at offset[g] + t.  One warp per (ply, game) entry:
// 17 B of position, A floats of pi (lane j moves entries j, j + 32, j + 64), 10 B of indices.
__global__ void __launch_bounds__(AZ_SAMPLE_WARPS * 32) k_samples_pack(
        const uint64_t* __restrict__ h_own, const uint64_t* __restrict__ h_opp, const int8_t* __restrict__ h_player,
        const uint8_t* __restrict__ h_live, const float* __restrict__ h_pi, const int8_t* __restrict__ winner,
        const int64_t* __restrict__ offset, int T, int64_t G, int A, uint64_t* __restrict__ own, uint64_t* __restrict__ opp,
        float* __restrict__ pi, int8_t* __restrict__ z, int32_t* __restrict__ episode, int32_t* __restrict__ move) {
    const int64_t e = (int64_t)blockIdx.x * AZ_SAMPLE_WARPS + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (e >= (int64_t)T * G) return;
    if (!h_live[e]) return;
    const int64_t t = e / G, g = e - t * G;
    const int64_t i = offset[g] + t;
    const float* src = h_pi + e * A;
    float* dst = pi + i * A;
    for (int a = lane; a < A; a += 32) dst[a] = src[a];
    if (lane == 0) {
        const int8_t pl = h_player[e];
        own[i] = h_own[e];
        opp[i] = h_opp[e];
        z[i] = (int8_t)(winner[g] * pl);
        episode[i] = (int32_t)g;
        move[i] = (int32_t)t;
    }
}

extern "C" int az_samples_pack(const uint64_t* h_own, const uint64_t* h_opp, const int8_t* h_player, const uint8_t* h_live,
                               const float* h_pi, const int8_t* winner, const int64_t* offset, int T, int64_t G,
                               int action_size, uint64_t* own, uint64_t* opp, float* pi, int8_t* z, int32_t* episode_idx,
                               int32_t* move_idx, void* stream) {
    if (T < 0 || G < 0 || action_size < 1 || action_size > AZ_CANVAS_ACTIONS) return AZ_ERR_ARG;
    if (T == 0 || G == 0) return AZ_OK;
    if (!h_own || !h_opp || !h_player || !h_live || !h_pi || !winner || !offset || !own || !opp || !pi || !z ||
        !episode_idx || !move_idx)
        return AZ_ERR_ARG;
    const int64_t entries = (int64_t)T * G;
    const unsigned grid = (unsigned)((entries + AZ_SAMPLE_WARPS - 1) / AZ_SAMPLE_WARPS);
    k_samples_pack<<<grid, AZ_SAMPLE_WARPS * 32, 0, (cudaStream_t)stream>>>(h_own, h_opp, h_player, h_live, h_pi, winner,
                                                                           offset, T, G, action_size, own, opp, pi, z,
                                                                           episode_idx, move_idx);
    AZ_CHECK_LAUNCH("k_samples_pack");
    return AZ_OK;
}
