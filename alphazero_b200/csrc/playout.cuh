// Uniform random playout shared by az_othello_playouts and the rollout-mode tree kernels.
// Replaces MCT.rollout + OthelloBoard.get_random_move (alphazero/mcts.py:173-180, games/othello.py:191-194).
#pragma once
#include "othello_bits.cuh"
#include "philox.cuh"

// One uniform random playout to the end of the game from (own, opp, player); returns the plies played.
// Random stream: philox4x32_10(ctr = {id_lo, id_hi, ply, 0}, key = seed).x per ply with a choice to make.
__device__ __forceinline__ int az_random_playout(uint64_t& own, uint64_t& opp, int& player, uint64_t bmask,
                                                 uint64_t id, uint64_t seed) {
    int ply = 0;
    while (true) {
        uint64_t legal = az_legal_mask(own, opp, bmask);
        if (legal == 0ULL) {
            if (az_legal_mask(opp, own, bmask) == 0ULL) break;     // neither side can place: game over
            uint64_t t = own; own = opp; opp = t;                  // forced pass
        } else {
            int cnt = __popcll(legal);
            az_u32x4 r = az_philox4x32_10((uint32_t)id, (uint32_t)(id >> 32), (uint32_t)ply, 0u,
                                          (uint32_t)seed, (uint32_t)(seed >> 32));
            int a = az_nth_set_bit(legal, (int)az_bounded(r.x, (uint32_t)cnt));
            uint64_t mv = 1ULL << a;
            uint64_t f = az_flips(mv, own, opp);
            uint64_t no = opp ^ f;
            opp = own ^ (f | mv);
            own = no;
        }
        player = -player;
        ++ply;
    }
    return ply;
}

