// Othello board kernels + their C-ABI entry points (include/az_b200.h).  sm_100a only.
//
// All of these are register-resident integer work (16 B in, <= 17 B out per board): the bound is the INT32 ALU
// pipe, not HBM.  One thread owns one board, so a warp is a batch of 32 boards; loads/stores of the SoA
// arrays are therefore fully coalesced (32 x 8 B per array per warp).  Grids are sized as a multiple of the
// SM count with a grid-stride loop.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <string.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "othello_bits.cuh"
#include "philox.cuh"
#include "playout.cuh"

// ------------------------------------------------------------------------------------------------ movegen

__global__ void __launch_bounds__(256) k_othello_legal(const uint64_t* __restrict__ own,
                                                       const uint64_t* __restrict__ opp, uint64_t bmask,
                                                       int64_t G, uint64_t* __restrict__ legal) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x)
        legal[g] = az_legal_mask(own[g], opp[g], bmask);
}

__global__ void __launch_bounds__(256) k_othello_flips(const uint64_t* __restrict__ own,
                                                       const uint64_t* __restrict__ opp,
                                                       const uint8_t* __restrict__ action, uint64_t bmask,
                                                       int64_t G, uint64_t* __restrict__ flips) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        int a = action[g];
        uint64_t o = own[g], p = opp[g];
        uint64_t mv = (a < 64) ? (1ULL << a) : 0ULL;
        mv &= bmask & ~(o | p);
        flips[g] = mv ? az_flips(mv, o, p) : 0ULL;
    }
}

__global__ void __launch_bounds__(256) k_othello_apply(const uint64_t* __restrict__ own,
                                                       const uint64_t* __restrict__ opp,
                                                       const int8_t* __restrict__ player,
                                                       const uint8_t* __restrict__ action, uint64_t bmask,
                                                       int64_t G, uint64_t* out_own, uint64_t* out_opp,
                                                       int8_t* out_player, uint8_t* __restrict__ status) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        uint64_t o = own[g], p = opp[g];
        int pl = player[g];
        int a = action[g];
        uint64_t legal = az_legal_mask(o, p, bmask);
        bool ok;
        if (a == AZ_PASS) ok = (legal == 0ULL);                       // pass only when nothing else (othello.py:158-165)
        else ok = (a < 64) && ((legal >> a) & 1ULL);
        if (ok) {
            az_apply(o, p, a);
            pl = -pl;
        }
        out_own[g] = o; out_opp[g] = p; out_player[g] = (int8_t)pl;
        if (status) status[g] = ok ? 0 : 1;
    }
}

__global__ void __launch_bounds__(256) k_othello_terminal(const uint64_t* __restrict__ own,
                                                          const uint64_t* __restrict__ opp,
                                                          const int8_t* __restrict__ player, uint64_t bmask,
                                                          int64_t G, uint8_t* __restrict__ terminal,
                                                          int8_t* __restrict__ score, int8_t* __restrict__ winner) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        uint64_t o = own[g], p = opp[g];
        if (terminal) terminal[g] = az_is_terminal(o, p, bmask) ? 1 : 0;
        if (score) score[g] = (int8_t)az_disc_diff(o, p);
        if (winner) winner[g] = (int8_t)az_winner(o, p, player[g]);
    }
}

template <typename T>
__device__ __forceinline__ T az_cell_value(int v);
template <>
__device__ __forceinline__ float az_cell_value<float>(int v) { return (float)v; }
template <>
__device__ __forceinline__ __nv_bfloat16 az_cell_value<__nv_bfloat16>(int v) { return __float2bfloat16((float)v); }

// One thread per output cell: consecutive threads write consecutive cells (coalesced stores).
template <typename T>
__global__ void __launch_bounds__(256) k_othello_encode(const uint64_t* __restrict__ own,
                                                        const uint64_t* __restrict__ opp, int n, int64_t G,
                                                        T* __restrict__ out) {
    const int cells = n * n;
    const int64_t total = G * cells;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        int64_t g = i / cells;
        int k = (int)(i - g * cells);
        int bit = (k / n) * 8 + (k % n);
        int v = (int)((own[g] >> bit) & 1ULL) - (int)((opp[g] >> bit) & 1ULL);
        out[i] = az_cell_value<T>(v);
    }
}

// ------------------------------------------------------------------------------------------------ perft

// counters[0] = size of the next frontier, counters[1] = finished games met so far (each counts as one leaf),
// counters[2] = leaves counted by the DFS kernel.
__global__ void __launch_bounds__(256) k_perft_expand(const uint64_t* __restrict__ own,
                                                      const uint64_t* __restrict__ opp, uint64_t bmask,
                                                      uint64_t F, uint64_t cap, uint64_t* __restrict__ nown,
                                                      uint64_t* __restrict__ nopp,
                                                      unsigned long long* __restrict__ counters) {
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < F; i += (uint64_t)gridDim.x * blockDim.x) {
        uint64_t o = own[i], p = opp[i];
        uint64_t legal = az_legal_mask(o, p, bmask);
        if (legal == 0ULL) {
            if (az_legal_mask(p, o, bmask) == 0ULL) {
                atomicAdd(&counters[1], 1ULL);
            } else {
                unsigned long long slot = atomicAdd(&counters[0], 1ULL);
                if (slot < cap) { nown[slot] = p; nopp[slot] = o; }
            }
            continue;
        }
        int cnt = __popcll(legal);
        unsigned long long slot = atomicAdd(&counters[0], (unsigned long long)cnt);
        while (legal) {
            int a = __ffsll((long long)legal) - 1;
            legal &= legal - 1;
            uint64_t co = o, cp = p;
            az_apply(co, cp, a);
            if (slot < cap) { nown[slot] = co; nopp[slot] = cp; }
            ++slot;
        }
    }
}

#define AZ_PERFT_MAX_DFS 8

__device__ unsigned long long az_perft_dfs(uint64_t own, uint64_t opp, uint64_t bmask, int depth) {
    // Iterative depth-first walk; a pass is a ply, a finished game is one leaf, the last ply is counted by popc.
    uint64_t s_own[AZ_PERFT_MAX_DFS], s_opp[AZ_PERFT_MAX_DFS], s_moves[AZ_PERFT_MAX_DFS];
    unsigned long long total = 0;
    int sp = 0;
    int d = depth;          // plies left at the node on top
    uint64_t o = own, p = opp;
    bool enter = true;      // true: (o,p,d) is a node to open; false: pop
    while (true) {
        if (enter) {
            if (d == 0) { total += 1; enter = false; continue; }
            uint64_t legal = az_legal_mask(o, p, bmask);
            if (legal == 0ULL) {
                if (d == 1 || az_legal_mask(p, o, bmask) == 0ULL) { total += 1; enter = false; continue; }
                uint64_t t = o; o = p; p = t;   // forced pass costs a ply
                d -= 1;
                // the parent frame stays as it is: the pass child is its only child
                s_own[sp] = 0; s_opp[sp] = 0; s_moves[sp] = 0; ++sp;
                continue;
            }
            if (d == 1) { total += (unsigned long long)__popcll(legal); enter = false; continue; }
            s_own[sp] = o; s_opp[sp] = p; s_moves[sp] = legal; ++sp;
            enter = false;
            // fallthrough to the pop logic below, which takes the first move of the frame just pushed
        }
        // pop: find the deepest frame with a move left
        while (sp > 0 && s_moves[sp - 1] == 0ULL) --sp;
        if (sp == 0) break;
        uint64_t m = s_moves[sp - 1];
        int a = __ffsll((long long)m) - 1;
        s_moves[sp - 1] = m & (m - 1);
        o = s_own[sp - 1]; p = s_opp[sp - 1];
        az_apply(o, p, a);
        d = depth - sp;     // frames on the stack = plies already spent
        enter = true;
    }
    return total;
}

__global__ void __launch_bounds__(128) k_perft_dfs(const uint64_t* __restrict__ own,
                                                   const uint64_t* __restrict__ opp, uint64_t bmask, uint64_t F,
                                                   int depth, unsigned long long* __restrict__ counters) {
    unsigned long long local = 0;
    for (uint64_t i = blockIdx.x * (uint64_t)blockDim.x + threadIdx.x; i < F; i += (uint64_t)gridDim.x * blockDim.x)
        local += az_perft_dfs(own[i], opp[i], bmask, depth);
    // warp reduction, one atomic per warp
    for (int off = 16; off > 0; off >>= 1) local += __shfl_down_sync(0xffffffffu, local, off);
    if ((threadIdx.x & 31) == 0 && local) atomicAdd(&counters[2], local);
}

// ------------------------------------------------------------------------------------------------ playouts

__global__ void __launch_bounds__(128) k_othello_playouts(const uint64_t* __restrict__ own,
                                                          const uint64_t* __restrict__ opp,
                                                          const int8_t* __restrict__ player, uint64_t bmask,
                                                          int64_t n_roots, int64_t per_root, uint64_t seed,
                                                          uint64_t first_id, uint64_t* __restrict__ final_black,
                                                          uint64_t* __restrict__ final_white,
                                                          int8_t* __restrict__ winner, uint8_t* __restrict__ plies) {
    const int64_t total = n_roots * per_root;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        int64_t g = t / per_root;
        uint64_t o = own[g], p = opp[g];
        int pl = player[g];
        int np = az_random_playout(o, p, pl, bmask, first_id + (uint64_t)t, seed);
        uint64_t black = pl > 0 ? o : p, white = pl > 0 ? p : o;
        if (final_black) final_black[t] = black;
        if (final_white) final_white[t] = white;
        if (winner) {
            int d = __popcll(black) - __popcll(white);
            winner[t] = (int8_t)(d > 0 ? 1 : (d < 0 ? -1 : 0));
        }
        if (plies) plies[t] = (uint8_t)np;
    }
}

// ------------------------------------------------------------------------------------------------ C ABI

static thread_local char g_last_error[512] = "";

int az_set_cuda_error(cudaError_t e, const char* where) {
    snprintf(g_last_error, sizeof(g_last_error), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
    return AZ_ERR_CUDA;
}

extern "C" int az_abi_version(void) { return 10; }
// a non-CUDA failure with a message for az_last_cuda_error (argument / capacity errors that deserve an explanation)
int az_set_error_message(int code, const char* msg) {
    snprintf(g_last_error, sizeof(g_last_error), "%s", msg);
    return code;
}
extern "C" const char* az_last_cuda_error(void) { return g_last_error; }

extern "C" int az_othello_legal(const uint64_t* own, const uint64_t* opp, int n, int64_t G, uint64_t* legal,
                                void* stream) {
    if (!az_valid_n(n) || G < 0 || (G > 0 && (!own || !opp || !legal))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_othello_legal<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, az_board_mask(n), G, legal);
    AZ_CHECK_LAUNCH("k_othello_legal");
    return AZ_OK;
}

extern "C" int az_othello_flips(const uint64_t* own, const uint64_t* opp, const uint8_t* action, int n, int64_t G,
                                uint64_t* flips, void* stream) {
    if (!az_valid_n(n) || G < 0 || (G > 0 && (!own || !opp || !action || !flips))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_othello_flips<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, action, az_board_mask(n), G,
                                                                          flips);
    AZ_CHECK_LAUNCH("k_othello_flips");
    return AZ_OK;
}

extern "C" int az_othello_apply(const uint64_t* own, const uint64_t* opp, const int8_t* player,
                                const uint8_t* action, int n, int64_t G, uint64_t* out_own, uint64_t* out_opp,
                                int8_t* out_player, uint8_t* status, void* stream) {
    if (!az_valid_n(n) || G < 0 || (G > 0 && (!own || !opp || !player || !action || !out_own || !out_opp || !out_player)))
        return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_othello_apply<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, player, action,
                                                                          az_board_mask(n), G, out_own, out_opp,
                                                                          out_player, status);
    AZ_CHECK_LAUNCH("k_othello_apply");
    return AZ_OK;
}

extern "C" int az_othello_terminal_score(const uint64_t* own, const uint64_t* opp, const int8_t* player, int n,
                                         int64_t G, uint8_t* terminal, int8_t* score, int8_t* winner,
                                         void* stream) {
    if (!az_valid_n(n) || G < 0 || (G > 0 && (!own || !opp || (winner && !player)))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_othello_terminal<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, player, az_board_mask(n), G,
                                                                             terminal, score, winner);
    AZ_CHECK_LAUNCH("k_othello_terminal");
    return AZ_OK;
}

extern "C" int az_othello_encode(const uint64_t* own, const uint64_t* opp, int n, int64_t G, void* out, int dtype,
                                 void* stream) {
    if (!az_valid_n(n) || G < 0 || (dtype != 0 && dtype != 1) || (G > 0 && (!own || !opp || !out))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    int grid = az_grid_for(G * n * n, 256);
    if (dtype == 0)
        k_othello_encode<float><<<grid, 256, 0, (cudaStream_t)stream>>>(own, opp, n, G, (float*)out);
    else
        k_othello_encode<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(own, opp, n, G, (__nv_bfloat16*)out);
    AZ_CHECK_LAUNCH("k_othello_encode");
    return AZ_OK;
}

extern "C" int az_othello_playouts(const uint64_t* own, const uint64_t* opp, const int8_t* player, int n,
                                   int64_t n_roots, int64_t per_root, uint64_t seed, uint64_t first_id,
                                   uint64_t* final_black, uint64_t* final_white, int8_t* winner, uint8_t* plies,
                                   void* stream) {
    if (!az_valid_n(n) || n_roots < 0 || per_root < 0 || (n_roots > 0 && (!own || !opp || !player))) return AZ_ERR_ARG;
    int64_t total = n_roots * per_root;
    if (total == 0) return AZ_OK;
    // 128-thread CTAs, capped at 148 SMs x 16 resident CTAs; the loop is divergent (games end at different
    // plies) so more, smaller CTAs keep the schedulers fed.
    int64_t need = (total + 127) / 128;
    int64_t cap = (int64_t)AZ_NUM_SMS_B200 * 16;
    int grid = (int)(need < cap ? need : cap);
    k_othello_playouts<<<grid, 128, 0, (cudaStream_t)stream>>>(own, opp, player, az_board_mask(n), n_roots, per_root,
                                                               seed, first_id, final_black, final_white, winner,
                                                               plies);
    AZ_CHECK_LAUNCH("k_othello_playouts");
    return AZ_OK;
}

extern "C" int az_othello_perft(uint64_t own_h, uint64_t opp_h, int n, int depth, int dfs_depth, uint64_t* scratch,
                                size_t scratch_u64, uint64_t* count_h, void* stream_) {
    if (!az_valid_n(n) || depth < 0 || dfs_depth < 0 || dfs_depth > AZ_PERFT_MAX_DFS - 1 || !scratch || !count_h ||
        scratch_u64 < 8 + 4 * 64)
        return AZ_ERR_ARG;
    cudaStream_t stream = (cudaStream_t)stream_;
    const uint64_t bmask = az_board_mask(n);
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
        k_perft_expand<<<az_grid_for((int64_t)F, 256), 256, 0, stream>>>(bufs[2 * cur], bufs[2 * cur + 1], bmask, F, cap,
                                                                        bufs[2 * (1 - cur)], bufs[2 * (1 - cur) + 1],
                                                                        counters);
        AZ_CHECK_LAUNCH("k_perft_expand");
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
        k_perft_dfs<<<(int)(need < gcap ? need : gcap), 128, 0, stream>>>(bufs[2 * cur], bufs[2 * cur + 1], bmask, F,
                                                                         rest, counters);
        AZ_CHECK_LAUNCH("k_perft_dfs");
    }
    AZ_CUDA_TRY(cudaMemcpyAsync(host_counters, counters, 24, cudaMemcpyDeviceToHost, stream));
    AZ_CUDA_TRY(cudaStreamSynchronize(stream));
    *count_h = host_counters[1] + host_counters[2];
    return AZ_OK;
}
