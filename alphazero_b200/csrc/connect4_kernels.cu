// Connect4 board kernels + their C-ABI entry points (include/az_b200.h).  sm_100a only.
//
// Same shape as the Othello board kernels: register-resident integer work, one thread per board (a warp = 32
// boards, SoA arrays read/written fully coalesced), grid-stride loops over grids sized from the SM count.
// Replaces the NumPy code of alphazero/games/connect4.py:111-258 (see connect4_bits.cuh for the line map).
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "connect4_bits.cuh"
#include "rules.cuh"

static inline bool az_valid_c4(int width, int height) { return width >= 4 && width <= 8 && height >= 4 && height <= 8; }

__global__ void __launch_bounds__(256) k_connect4_legal(const uint64_t* __restrict__ own,
                                                        const uint64_t* __restrict__ opp, int width, int64_t G,
                                                        uint64_t* __restrict__ legal) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x)
        legal[g] = az_c4_legal_cols(own[g], opp[g], width);
}

__global__ void __launch_bounds__(256) k_connect4_apply(const uint64_t* __restrict__ own,
                                                        const uint64_t* __restrict__ opp,
                                                        const int8_t* __restrict__ player,
                                                        const uint8_t* __restrict__ action, int width, uint64_t bmask,
                                                        int64_t G, uint64_t* out_own, uint64_t* out_opp,
                                                        int8_t* out_player, uint8_t* __restrict__ status) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        uint64_t o = own[g], p = opp[g];
        int pl = player[g];
        const int a = action[g];
        // connect4.py:145-155: out-of-range column or full column -> ValueError in the reference
        const bool ok = a < width && ((az_c4_legal_cols(o, p, width) >> a) & 1ULL);
        if (ok) {
            az_c4_apply(o, p, a, bmask);
            pl = -pl;
        }
        out_own[g] = o; out_opp[g] = p; out_player[g] = (int8_t)pl;
        if (status) status[g] = ok ? 0 : 1;
    }
}

__global__ void __launch_bounds__(256) k_connect4_terminal(const uint64_t* __restrict__ own,
                                                           const uint64_t* __restrict__ opp,
                                                           const int8_t* __restrict__ player, uint64_t bmask,
                                                           int64_t G, uint8_t* __restrict__ terminal,
                                                           int8_t* __restrict__ score, int8_t* __restrict__ winner) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        uint64_t o = own[g], p = opp[g];
        if (terminal) terminal[g] = az_c4_is_terminal(o, p, bmask) ? 1 : 0;
        if (score) score[g] = (int8_t)az_disc_diff(o, p);
        if (winner) winner[g] = (int8_t)az_c4_winner(o, p, player[g]);
    }
}

template <typename T>
__device__ __forceinline__ T az_c4_cell_value(int v);
template <>
__device__ __forceinline__ float az_c4_cell_value<float>(int v) { return (float)v; }
template <>
__device__ __forceinline__ __nv_bfloat16 az_c4_cell_value<__nv_bfloat16>(int v) { return __float2bfloat16((float)v); }

// One thread per output cell, row-major (height, width) like the reference's grid.
template <typename T>
__global__ void __launch_bounds__(256) k_connect4_encode(const uint64_t* __restrict__ own,
                                                         const uint64_t* __restrict__ opp, int width, int height,
                                                         int64_t G, T* __restrict__ out) {
    const int cells = width * height;
    const int64_t total = G * cells;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t g = i / cells;
        int k = (int)(i - g * cells);
        int bit = (k / width) * 8 + (k % width);
        int v = (int)((own[g] >> bit) & 1ULL) - (int)((opp[g] >> bit) & 1ULL);
        out[i] = az_c4_cell_value<T>(v);
    }
}

__global__ void __launch_bounds__(128) k_connect4_playouts(const uint64_t* __restrict__ own,
                                                           const uint64_t* __restrict__ opp,
                                                           const int8_t* __restrict__ player, int width, int height,
                                                           int64_t n_roots, int64_t per_root, uint64_t seed,
                                                           uint64_t first_id, uint64_t* __restrict__ final_black,
                                                           uint64_t* __restrict__ final_white,
                                                           int8_t* __restrict__ winner, uint8_t* __restrict__ plies) {
    using R = az_rules<AZ_GAME_CONNECT4>;
    const az_dims d = R::dims(width, height);
    const int64_t total = n_roots * per_root;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t g = t / per_root;
        uint64_t o = own[g], p = opp[g];
        int pl = player[g];
        int np = az_random_playout_g<AZ_GAME_CONNECT4>(o, p, pl, d, first_id + (uint64_t)t, seed);
        if (final_black) final_black[t] = pl > 0 ? o : p;
        if (final_white) final_white[t] = pl > 0 ? p : o;
        if (winner) winner[t] = (int8_t)R::winner(o, p, pl, d);
        if (plies) plies[t] = (uint8_t)np;
    }
}

// ------------------------------------------------------------------------------------------------ perft

// counters[0] = next frontier size, counters[1] = finished games met (one leaf each), counters[2] = DFS leaves.
__global__ void __launch_bounds__(256) k_c4_perft_expand(const uint64_t* __restrict__ own,
                                                         const uint64_t* __restrict__ opp, int width, uint64_t bmask,
                                                         uint64_t F, uint64_t cap, uint64_t* __restrict__ nown,
                                                         uint64_t* __restrict__ nopp,
                                                         unsigned long long* __restrict__ counters) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < F; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t o = own[i], p = opp[i];
        if (az_c4_is_terminal(o, p, bmask)) {
            atomicAdd(&counters[1], 1ULL);
            continue;
        }
        uint64_t legal = az_c4_legal_cols(o, p, width);
        unsigned long long slot = atomicAdd(&counters[0], (unsigned long long)__popcll(legal));
        while (legal) {
            int a = __ffsll((long long)legal) - 1;
            legal &= legal - 1;
            uint64_t co = o, cp = p;
            az_c4_apply(co, cp, a, bmask);
            if (slot < cap) { nown[slot] = co; nopp[slot] = cp; }
            ++slot;
        }
    }
}

#define AZ_C4_PERFT_MAX_DFS 8

__device__ unsigned long long az_c4_perft_dfs(uint64_t own, uint64_t opp, int width, uint64_t bmask, int depth) {
    // iterative depth-first walk with an explicit stack of (position, untried columns)
    uint64_t s_own[AZ_C4_PERFT_MAX_DFS], s_opp[AZ_C4_PERFT_MAX_DFS];
    uint32_t s_moves[AZ_C4_PERFT_MAX_DFS];
    if (depth == 0 || az_c4_is_terminal(own, opp, bmask)) return 1ULL;
    unsigned long long total = 0;
    int sp = 0;
    s_own[0] = own; s_opp[0] = opp; s_moves[0] = (uint32_t)az_c4_legal_cols(own, opp, width); sp = 1;
    while (sp > 0) {
        uint32_t m = s_moves[sp - 1];
        if (m == 0u) { --sp; continue; }
        int a = __ffs((int)m) - 1;
        s_moves[sp - 1] = m & (m - 1u);
        uint64_t o = s_own[sp - 1], p = s_opp[sp - 1];
        az_c4_apply(o, p, a, bmask);
        const int left = depth - sp;          // plies left below the child
        if (left == 0 || az_c4_is_terminal(o, p, bmask)) { total += 1ULL; continue; }
        s_own[sp] = o; s_opp[sp] = p; s_moves[sp] = (uint32_t)az_c4_legal_cols(o, p, width); ++sp;
    }
    return total;
}

__global__ void __launch_bounds__(128) k_c4_perft_dfs(const uint64_t* __restrict__ own,
                                                      const uint64_t* __restrict__ opp, int width, uint64_t bmask,
                                                      uint64_t F, int depth, unsigned long long* __restrict__ counters) {
    unsigned long long local = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < F; i += (uint64_t)gridDim.x * blockDim.x)
        local += az_c4_perft_dfs(own[i], opp[i], width, bmask, depth);
    for (int off = 16; off > 0; off >>= 1) local += __shfl_down_sync(0xffffffffu, local, off);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&counters[2], local);
}

// ------------------------------------------------------------------------------------------------ C ABI

extern "C" int az_connect4_legal(const uint64_t* own, const uint64_t* opp, int width, int height, int64_t G,
                                 uint64_t* legal, void* stream) {
    if (!az_valid_c4(width, height) || G < 0 || (G > 0 && (!own || !opp || !legal))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_connect4_legal<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, width, G, legal);
    AZ_CHECK_LAUNCH("k_connect4_legal");
    return AZ_OK;
}

extern "C" int az_connect4_apply(const uint64_t* own, const uint64_t* opp, const int8_t* player,
                                 const uint8_t* action, int width, int height, int64_t G, uint64_t* out_own,
                                 uint64_t* out_opp, int8_t* out_player, uint8_t* status, void* stream) {
    if (!az_valid_c4(width, height) || G < 0 ||
        (G > 0 && (!own || !opp || !player || !action || !out_own || !out_opp || !out_player)))
        return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_connect4_apply<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(
        own, opp, player, action, width, az_c4_board_mask(width, height), G, out_own, out_opp, out_player, status);
    AZ_CHECK_LAUNCH("k_connect4_apply");
    return AZ_OK;
}

extern "C" int az_connect4_terminal_score(const uint64_t* own, const uint64_t* opp, const int8_t* player, int width,
                                          int height, int64_t G, uint8_t* terminal, int8_t* score, int8_t* winner,
                                          void* stream) {
    if (!az_valid_c4(width, height) || G < 0 || (G > 0 && (!own || !opp || (winner && !player)))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_connect4_terminal<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(
        own, opp, player, az_c4_board_mask(width, height), G, terminal, score, winner);
    AZ_CHECK_LAUNCH("k_connect4_terminal");
    return AZ_OK;
}

extern "C" int az_connect4_encode(const uint64_t* own, const uint64_t* opp, int width, int height, int64_t G,
                                  void* out, int dtype, void* stream) {
    if (!az_valid_c4(width, height) || G < 0 || (dtype != 0 && dtype != 1) || (G > 0 && (!own || !opp || !out)))
        return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    int grid = az_grid_for(G * width * height, 256);
    if (dtype == 0)
        k_connect4_encode<float><<<grid, 256, 0, (cudaStream_t)stream>>>(own, opp, width, height, G, (float*)out);
    else
        k_connect4_encode<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(own, opp, width, height, G,
                                                                                (__nv_bfloat16*)out);
    AZ_CHECK_LAUNCH("k_connect4_encode");
    return AZ_OK;
}

extern "C" int az_connect4_playouts(const uint64_t* own, const uint64_t* opp, const int8_t* player, int width,
                                    int height, int64_t n_roots, int64_t per_root, uint64_t seed, uint64_t first_id,
                                    uint64_t* final_black, uint64_t* final_white, int8_t* winner, uint8_t* plies,
                                    void* stream) {
    if (!az_valid_c4(width, height) || n_roots < 0 || per_root < 0 || (n_roots > 0 && (!own || !opp || !player)))
        return AZ_ERR_ARG;
    int64_t total = n_roots * per_root;
    if (total == 0) return AZ_OK;
    int64_t need = (total + 127) / 128;
    int64_t cap = (int64_t)AZ_NUM_SMS_B200 * 16;
    k_connect4_playouts<<<(int)(need < cap ? need : cap), 128, 0, (cudaStream_t)stream>>>(
        own, opp, player, width, height, n_roots, per_root, seed, first_id, final_black, final_white, winner, plies);
    AZ_CHECK_LAUNCH("k_connect4_playouts");
    return AZ_OK;
}

extern "C" int az_connect4_perft(uint64_t own_h, uint64_t opp_h, int width, int height, int depth, int dfs_depth,
                                 uint64_t* scratch, size_t scratch_u64, uint64_t* count_h, void* stream_) {
    if (!az_valid_c4(width, height) || depth < 0 || dfs_depth < 0 || dfs_depth > AZ_C4_PERFT_MAX_DFS - 1 || !scratch ||
        !count_h || scratch_u64 < 8 + 4 * 64)
        return AZ_ERR_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const uint64_t bmask = az_c4_board_mask(width, height);
    const uint64_t cap = (scratch_u64 - 8) / 4;
    unsigned long long* counters = (unsigned long long*)scratch;
    uint64_t* bufs[4] = {scratch + 8, scratch + 8 + cap, scratch + 8 + 2 * cap, scratch + 8 + 3 * cap};
    unsigned long long host_counters[3] = {0, 0, 0};
    AZ_CUDA_TRY(cudaMemsetAsync(counters, 0, 8 * sizeof(uint64_t), stream));
    AZ_CUDA_TRY(cudaMemcpyAsync(bufs[0], &own_h, 8, cudaMemcpyHostToDevice, stream));
    AZ_CUDA_TRY(cudaMemcpyAsync(bufs[1], &opp_h, 8, cudaMemcpyHostToDevice, stream));
    AZ_CUDA_TRY(cudaStreamSynchronize(stream));
    uint64_t F = 1;
    int cur = 0;
    int levels = depth > dfs_depth ? depth - dfs_depth : 0;
    for (int l = 0; l < levels && F > 0; ++l) {
        AZ_CUDA_TRY(cudaMemsetAsync(counters, 0, sizeof(uint64_t), stream));
        k_c4_perft_expand<<<az_grid_for((int64_t)F, 256), 256, 0, stream>>>(
            bufs[2 * cur], bufs[2 * cur + 1], width, bmask, F, cap, bufs[2 * (1 - cur)], bufs[2 * (1 - cur) + 1], counters);
        AZ_CHECK_LAUNCH("k_c4_perft_expand");
        AZ_CUDA_TRY(cudaMemcpyAsync(host_counters, counters, 8, cudaMemcpyDeviceToHost, stream));
        AZ_CUDA_TRY(cudaStreamSynchronize(stream));
        F = host_counters[0];
        if (F > cap) return AZ_ERR_CAPACITY;
        cur = 1 - cur;
    }
    int rest = depth - levels;
    if (F > 0) {
        int64_t need = ((int64_t)F + 127) / 128;
        int64_t gcap = (int64_t)AZ_NUM_SMS_B200 * 16;
        k_c4_perft_dfs<<<(int)(need < gcap ? need : gcap), 128, 0, stream>>>(bufs[2 * cur], bufs[2 * cur + 1], width, bmask,
                                                                            F, rest, counters);
        AZ_CHECK_LAUNCH("k_c4_perft_dfs");
    }
    AZ_CUDA_TRY(cudaMemcpyAsync(host_counters, counters, 24, cudaMemcpyDeviceToHost, stream));
    AZ_CUDA_TRY(cudaStreamSynchronize(stream));
    *count_h = host_counters[1] + host_counters[2];
    return AZ_OK;
}
