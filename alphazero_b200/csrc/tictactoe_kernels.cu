// TicTacToe board kernels, the fused TicTacToeNet forward and their C-ABI entry points (include/az_b200.h).
// sm_100a only.  Same shape as the other board kernels: one thread per board, SoA arrays, grid-stride loops.
// Replaces alphazero/games/tictactoe.py:112-184 (board) and :302-316 (network forward) for batches of positions.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "rules.cuh"

__global__ void __launch_bounds__(256) k_tictactoe_legal(const uint64_t* __restrict__ own,
                                                         const uint64_t* __restrict__ opp, int64_t G,
                                                         uint64_t* __restrict__ legal) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x)
        legal[g] = az_ttt_legal(own[g], opp[g]);
}

__global__ void __launch_bounds__(256) k_tictactoe_apply(const uint64_t* __restrict__ own,
                                                         const uint64_t* __restrict__ opp,
                                                         const int8_t* __restrict__ player,
                                                         const uint8_t* __restrict__ action, int64_t G,
                                                         uint64_t* out_own, uint64_t* out_opp, int8_t* out_player,
                                                         uint8_t* __restrict__ status) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        uint64_t o = own[g], p = opp[g];
        int pl = player[g];
        const int a = action[g];
        // tictactoe.py:139-143, 159-160: a cell outside the board or an occupied cell -> ValueError in the reference
        const bool ok = a < 64 && ((az_ttt_legal(o, p) >> a) & 1ULL);
        if (ok) {
            az_ttt_apply(o, p, a);
            pl = -pl;
        }
        out_own[g] = o; out_opp[g] = p; out_player[g] = (int8_t)pl;
        if (status) status[g] = ok ? 0 : 1;
    }
}

__global__ void __launch_bounds__(256) k_tictactoe_terminal(const uint64_t* __restrict__ own,
                                                            const uint64_t* __restrict__ opp,
                                                            const int8_t* __restrict__ player, int64_t G,
                                                            uint8_t* __restrict__ terminal, int8_t* __restrict__ score,
                                                            int8_t* __restrict__ winner) {
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t o = own[g], p = opp[g];
        const bool term = az_ttt_is_terminal(o, p);
        if (terminal) terminal[g] = term ? 1 : 0;
        if (score) score[g] = (int8_t)az_ttt_score(o, p);
        if (winner) winner[g] = (int8_t)(term ? az_ttt_winner(o, p, player[g]) : 0);
    }
}

template <typename T>
__device__ __forceinline__ T az_ttt_cell_value(int v);
template <>
__device__ __forceinline__ float az_ttt_cell_value<float>(int v) { return (float)v; }
template <>
__device__ __forceinline__ __nv_bfloat16 az_ttt_cell_value<__nv_bfloat16>(int v) { return __float2bfloat16((float)v); }

template <typename T>
__global__ void __launch_bounds__(256) k_tictactoe_encode(const uint64_t* __restrict__ own,
                                                          const uint64_t* __restrict__ opp, int64_t G,
                                                          T* __restrict__ out) {
    const int64_t total = G * 9;
    for (int64_t i = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t g = i / 9;
        const int k = (int)(i - g * 9);
        const int bit = (k / 3) * 8 + (k % 3);
        out[i] = az_ttt_cell_value<T>((int)((own[g] >> bit) & 1ULL) - (int)((opp[g] >> bit) & 1ULL));
    }
}

__global__ void __launch_bounds__(128) k_tictactoe_playouts(const uint64_t* __restrict__ own,
                                                            const uint64_t* __restrict__ opp,
                                                            const int8_t* __restrict__ player, int64_t n_roots,
                                                            int64_t per_root, uint64_t seed, uint64_t first_id,
                                                            uint64_t* __restrict__ final_black,
                                                            uint64_t* __restrict__ final_white,
                                                            int8_t* __restrict__ winner, uint8_t* __restrict__ plies) {
    using R = az_rules<AZ_GAME_TICTACTOE>;
    const az_dims d = R::dims(3, 3);
    const int64_t total = n_roots * per_root;
    for (int64_t t = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; t < total; t += (int64_t)gridDim.x * blockDim.x) {
        const int64_t g = t / per_root;
        uint64_t o = own[g], p = opp[g];
        int pl = player[g];
        const int np = az_random_playout_g<AZ_GAME_TICTACTOE>(o, p, pl, d, first_id + (uint64_t)t, seed);
        if (final_black) final_black[t] = pl > 0 ? o : p;
        if (final_white) final_white[t] = pl > 0 ? p : o;
        if (winner) winner[t] = (int8_t)R::winner(o, p, pl, d);
        if (plies) plies[t] = (uint8_t)np;
    }
}

// ------------------------------------------------------------------------------------------------ TicTacToeNet

// weights (float32, eval-mode BatchNorm folded into fc1 / fc2 by the host): W1[9][9] b1[9] W2[9][9] b2[9]
// Wp[9][9] bp[9] Wv[9] bv -- 280 values, row = output unit.
#define AZ_TTT_NET_FLOATS 280

// One thread per position, everything in registers / constant-indexed shared memory: 9 -> 9 -> 9 -> (9 + 1) units,
// 261 FMAs.  probs = softmax over the 9 cells (log_softmax + exp of predict, base.py:350-355), value = tanh.
__global__ void __launch_bounds__(128) k_tictactoe_net(const uint64_t* __restrict__ own,
                                                       const uint64_t* __restrict__ opp, int64_t G,
                                                       const float* __restrict__ weights, float* __restrict__ probs,
                                                       float* __restrict__ value) {
    __shared__ float w[AZ_TTT_NET_FLOATS];
    for (int i = threadIdx.x; i < AZ_TTT_NET_FLOATS; i += blockDim.x) w[i] = weights[i];
    __syncthreads();
    const float *W1 = w, *b1 = w + 81, *W2 = w + 90, *b2 = w + 171, *Wp = w + 180, *bp = w + 261, *Wv = w + 270;
    const float bv = w[279];
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        const uint64_t o = own[g], p = opp[g];
        float x[9], h1[9], h2[9], z[9];
        #pragma unroll
        for (int k = 0; k < 9; ++k) {
            const int bit = (k / 3) * 8 + (k % 3);
            x[k] = (float)((int)((o >> bit) & 1ULL) - (int)((p >> bit) & 1ULL));
        }
        #pragma unroll
        for (int j = 0; j < 9; ++j) {
            float a = b1[j];
            #pragma unroll
            for (int k = 0; k < 9; ++k) a = fmaf(W1[j * 9 + k], x[k], a);
            h1[j] = fmaxf(a, 0.0f);
        }
        #pragma unroll
        for (int j = 0; j < 9; ++j) {
            float a = b2[j];
            #pragma unroll
            for (int k = 0; k < 9; ++k) a = fmaf(W2[j * 9 + k], h1[k], a);
            h2[j] = fmaxf(a, 0.0f);
        }
        float zmax = -3.0e38f, v = bv;
        #pragma unroll
        for (int j = 0; j < 9; ++j) {
            float a = bp[j];
            #pragma unroll
            for (int k = 0; k < 9; ++k) a = fmaf(Wp[j * 9 + k], h2[k], a);
            z[j] = a;
            zmax = fmaxf(zmax, a);
            v = fmaf(Wv[j], h2[j], v);
        }
        float sum = 0.0f;
        #pragma unroll
        for (int j = 0; j < 9; ++j) { z[j] = expf(z[j] - zmax); sum += z[j]; }
        const float inv = 1.0f / sum;
        #pragma unroll
        for (int j = 0; j < 9; ++j) probs[g * 9 + j] = z[j] * inv;
        value[g] = tanhf(v);
    }
}

// ------------------------------------------------------------------------------------------------ C ABI

extern "C" int az_tictactoe_legal(const uint64_t* own, const uint64_t* opp, int64_t G, uint64_t* legal, void* stream) {
    if (G < 0 || (G > 0 && (!own || !opp || !legal))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_tictactoe_legal<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, G, legal);
    AZ_CHECK_LAUNCH("k_tictactoe_legal");
    return AZ_OK;
}

extern "C" int az_tictactoe_apply(const uint64_t* own, const uint64_t* opp, const int8_t* player,
                                  const uint8_t* action, int64_t G, uint64_t* out_own, uint64_t* out_opp,
                                  int8_t* out_player, uint8_t* status, void* stream) {
    if (G < 0 || (G > 0 && (!own || !opp || !player || !action || !out_own || !out_opp || !out_player)))
        return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_tictactoe_apply<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, player, action, G, out_own,
                                                                            out_opp, out_player, status);
    AZ_CHECK_LAUNCH("k_tictactoe_apply");
    return AZ_OK;
}

extern "C" int az_tictactoe_terminal_score(const uint64_t* own, const uint64_t* opp, const int8_t* player, int64_t G,
                                           uint8_t* terminal, int8_t* score, int8_t* winner, void* stream) {
    if (G < 0 || (G > 0 && (!own || !opp || (winner && !player)))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_tictactoe_terminal<<<az_grid_for(G, 256), 256, 0, (cudaStream_t)stream>>>(own, opp, player, G, terminal, score,
                                                                               winner);
    AZ_CHECK_LAUNCH("k_tictactoe_terminal");
    return AZ_OK;
}

extern "C" int az_tictactoe_encode(const uint64_t* own, const uint64_t* opp, int64_t G, void* out, int dtype,
                                   void* stream) {
    if (G < 0 || (dtype != 0 && dtype != 1) || (G > 0 && (!own || !opp || !out))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    const int grid = az_grid_for(G * 9, 256);
    if (dtype == 0) k_tictactoe_encode<float><<<grid, 256, 0, (cudaStream_t)stream>>>(own, opp, G, (float*)out);
    else k_tictactoe_encode<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(own, opp, G, (__nv_bfloat16*)out);
    AZ_CHECK_LAUNCH("k_tictactoe_encode");
    return AZ_OK;
}

extern "C" int az_tictactoe_playouts(const uint64_t* own, const uint64_t* opp, const int8_t* player, int64_t n_roots,
                                     int64_t per_root, uint64_t seed, uint64_t first_id, uint64_t* final_black,
                                     uint64_t* final_white, int8_t* winner, uint8_t* plies, void* stream) {
    if (n_roots < 0 || per_root < 0 || (n_roots > 0 && (!own || !opp || !player))) return AZ_ERR_ARG;
    const int64_t total = n_roots * per_root;
    if (total == 0) return AZ_OK;
    k_tictactoe_playouts<<<az_grid_for(total, 128), 128, 0, (cudaStream_t)stream>>>(
        own, opp, player, n_roots, per_root, seed, first_id, final_black, final_white, winner, plies);
    AZ_CHECK_LAUNCH("k_tictactoe_playouts");
    return AZ_OK;
}

extern "C" int az_tictactoe_net_floats(void) { return AZ_TTT_NET_FLOATS; }

extern "C" int az_tictactoe_net(const uint64_t* own, const uint64_t* opp, int64_t G, const float* weights,
                                float* probs, float* value, void* stream) {
    if (G < 0 || !weights || (G > 0 && (!own || !opp || !probs || !value))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_tictactoe_net<<<az_grid_for(G, 128), 128, 0, (cudaStream_t)stream>>>(own, opp, G, weights, probs, value);
    AZ_CHECK_LAUNCH("k_tictactoe_net");
    return AZ_OK;
}
