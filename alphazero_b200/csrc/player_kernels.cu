// Move rules of the reference's search-free players for a batch of games (alphazero/players.py:76-123).
//
// GreedyPlayer.get_move (players.py:98-123): every legal move is played on a clone of the board, the move's value is
// -clone.get_score() (get_score is from the NEXT mover's viewpoint), the best value wins and ties are broken
// uniformly at random (utils.fair_max).  One thread per game: the positions are two 64-bit words, a move costs a few
// hundred integer instructions, nothing is read from memory but the position.
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "rules.cuh"

template <int GAME>
__global__ void __launch_bounds__(256) k_greedy_move(const uint64_t* __restrict__ own, const uint64_t* __restrict__ opp,
                                                     const uint8_t* __restrict__ mask, int n, int height, int64_t G,
                                                     uint64_t seed, uint32_t ctr, int tie_random,
                                                     uint8_t* __restrict__ action, uint64_t* __restrict__ ties) {
    using R = az_rules<GAME>;
    const az_dims d = R::dims(n, height);
    for (int64_t g = blockIdx.x * (int64_t)blockDim.x + threadIdx.x; g < G; g += (int64_t)gridDim.x * blockDim.x) {
        if (mask && !mask[g]) { action[g] = 255; if (ties) ties[g] = 0ULL; continue; }
        const uint64_t o = own[g], p = opp[g];
        const uint64_t legal = R::legal(o, p, d);
        if (legal == 0ULL) {                       // Othello: the pass is the only move (othello.py:187-188)
            action[g] = (uint8_t)R::nth_action(legal, 0);
            if (ties) ties[g] = 0ULL;
            continue;
        }
        int best = -1000;
        uint64_t best_set = 0ULL;
        for (uint64_t rest = legal; rest; rest &= rest - 1ULL) {
            const int a = __ffsll((long long)rest) - 1;
            uint64_t no = o, np = p;
            R::apply(no, np, a, d);                // sides swap: `no` is now the next mover's discs
            const int v = -R::score(no, np);
            if (v > best) { best = v; best_set = 0ULL; }
            if (v == best) best_set |= 1ULL << a;
        }
        int k = 0;
        const int cnt = __popcll(best_set);
        if (tie_random && cnt > 1) {               // a unique maximum consumes no random number (utils.py:28-34)
            const az_u32x4 r = az_philox4x32_10((uint32_t)g, (uint32_t)((uint64_t)g >> 32), ctr, 0x67726479u,
                                                (uint32_t)seed, (uint32_t)(seed >> 32));
            k = (int)az_bounded(r.x, (uint32_t)cnt);
        }
        action[g] = (uint8_t)az_nth_set_bit(best_set, k);
        if (ties) ties[g] = best_set;
    }
}

extern "C" int az_greedy_move(int game, const uint64_t* own, const uint64_t* opp, const uint8_t* mask, int n, int height,
                              int64_t G, uint64_t seed, uint32_t ctr, int tie_random, uint8_t* action, uint64_t* ties,
                              void* stream) {
    if (G < 0 || (G > 0 && (!own || !opp || !action))) return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    const int grid = az_grid_for(G, 256);
    cudaStream_t s = (cudaStream_t)stream;
    switch (game) {
        case AZ_GAME_OTHELLO:
            if (!az_valid_n(n)) return AZ_ERR_ARG;
            k_greedy_move<AZ_GAME_OTHELLO><<<grid, 256, 0, s>>>(own, opp, mask, n, n, G, seed, ctr, tie_random, action, ties);
            break;
        case AZ_GAME_CONNECT4:
            if (n < 4 || n > 8 || height < 4 || height > 8) return AZ_ERR_ARG;
            k_greedy_move<AZ_GAME_CONNECT4><<<grid, 256, 0, s>>>(own, opp, mask, n, height, G, seed, ctr, tie_random, action, ties);
            break;
        case AZ_GAME_TICTACTOE:
            k_greedy_move<AZ_GAME_TICTACTOE><<<grid, 256, 0, s>>>(own, opp, mask, 3, 3, G, seed, ctr, tie_random, action, ties);
            break;
        default:
            return AZ_ERR_ARG;
    }
    AZ_CHECK_LAUNCH("k_greedy_move");
    return AZ_OK;
}
