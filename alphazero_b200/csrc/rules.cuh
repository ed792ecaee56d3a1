// Game rules behind one compile-time interface, so that the tree / playout kernels are written once and
// instantiated per game (the reference does the same through the Board base class, alphazero/base.py:95-225).
//
//   az_rules<AZ_GAME_OTHELLO>   alphazero/games/othello.py   actions 0..63 = cell r*8+c, 64 = pass
//   az_rules<AZ_GAME_CONNECT4>  alphazero/games/connect4.py  actions 0..width-1 = column, no pass
//   az_rules<AZ_GAME_TICTACTOE> alphazero/games/tictactoe.py actions = cell r*8+c (r, c < 3), no pass
//
// `legal()` returns a 64-bit set of action ids (Othello: placement cells; Connect4: playable columns);
// children are created in ascending action order.  An empty set means "pass only" for Othello and cannot occur at
// a non-terminal Connect4 position.
#pragma once
#include "connect4_bits.cuh"
#include "othello_bits.cuh"
#include "tictactoe_bits.cuh"
#include "philox.cuh"

// member-function flavour of AZ_HD (a class member cannot be `static inline` twice on the host build)
#ifdef __CUDACC__
#define AZ_HDM __host__ __device__ __forceinline__
#else
#define AZ_HDM inline
#endif

struct az_dims {
    int rows, cols;
    uint64_t bmask;
};

template <int GAME>
struct az_rules;

template <>
struct az_rules<AZ_GAME_OTHELLO> {
    static AZ_HDM az_dims dims(int n, int /*height*/) { return az_dims{n, n, az_board_mask(n)}; }
    static AZ_HDM int action_size(const az_dims& d) { return d.rows * d.cols + 1; }
    static AZ_HDM uint64_t legal(uint64_t own, uint64_t opp, const az_dims& d) { return az_legal_mask(own, opp, d.bmask); }
    static AZ_HDM int n_actions(uint64_t legal) { return legal ? az_popc64(legal) : 1; }
    static AZ_HDM int nth_action(uint64_t legal, int j) { return legal ? az_nth_set_bit(legal, j) : AZ_PASS; }
    static AZ_HDM bool is_legal(uint64_t legal, int a) {
        return a == AZ_PASS ? legal == 0ULL : (a < 64 && ((legal >> a) & 1ULL));
    }
    static AZ_HDM void apply(uint64_t& own, uint64_t& opp, int a, const az_dims&) { az_apply(own, opp, a); }
    static AZ_HDM bool is_terminal(uint64_t own, uint64_t opp, const az_dims& d) { return az_is_terminal(own, opp, d.bmask); }
    static AZ_HDM int winner(uint64_t own, uint64_t opp, int player, const az_dims&) { return az_winner(own, opp, player); }
    // get_score(): sum(player * grid) from the side to move's view (othello.py:133-135)
    static AZ_HDM int score(uint64_t own, uint64_t opp) { return az_disc_diff(own, opp); }
    // index of a canvas action in the network's policy vector (r*n+c, pass last; othello.py:391, 408-409)
    static AZ_HDM int net_index(int a, const az_dims& d) { return a == AZ_PASS ? d.rows * d.cols : (a >> 3) * d.cols + (a & 7); }
};

template <>
struct az_rules<AZ_GAME_CONNECT4> {
    static AZ_HDM az_dims dims(int width, int height) { return az_dims{height, width, az_c4_board_mask(width, height)}; }
    static AZ_HDM int action_size(const az_dims& d) { return d.cols; }
    static AZ_HDM uint64_t legal(uint64_t own, uint64_t opp, const az_dims& d) { return az_c4_legal_cols(own, opp, d.cols); }
    static AZ_HDM int n_actions(uint64_t legal) { return legal ? az_popc64(legal) : 1; }
    static AZ_HDM int nth_action(uint64_t legal, int j) { return legal ? az_nth_set_bit(legal, j) : 0; }
    static AZ_HDM bool is_legal(uint64_t legal, int a) { return a < 8 && ((legal >> a) & 1ULL); }
    static AZ_HDM void apply(uint64_t& own, uint64_t& opp, int a, const az_dims& d) { az_c4_apply(own, opp, a, d.bmask); }
    static AZ_HDM bool is_terminal(uint64_t own, uint64_t opp, const az_dims& d) { return az_c4_is_terminal(own, opp, d.bmask); }
    static AZ_HDM int winner(uint64_t own, uint64_t opp, int player, const az_dims&) { return az_c4_winner(own, opp, player); }
    static AZ_HDM int score(uint64_t own, uint64_t opp) { return az_disc_diff(own, opp); }      // connect4.py:137-139
    static AZ_HDM int net_index(int a, const az_dims& d) { return a < d.cols ? a : 0; }
};

template <>
struct az_rules<AZ_GAME_TICTACTOE> {
    static AZ_HDM az_dims dims(int, int) { return az_dims{3, 3, AZ_TTT_BOARD}; }
    static AZ_HDM int action_size(const az_dims&) { return 9; }
    static AZ_HDM uint64_t legal(uint64_t own, uint64_t opp, const az_dims&) { return az_ttt_legal(own, opp); }
    static AZ_HDM int n_actions(uint64_t legal) { return legal ? az_popc64(legal) : 1; }
    static AZ_HDM int nth_action(uint64_t legal, int j) { return legal ? az_nth_set_bit(legal, j) : 0; }
    static AZ_HDM bool is_legal(uint64_t legal, int a) { return a < 64 && ((legal >> a) & 1ULL); }
    static AZ_HDM void apply(uint64_t& own, uint64_t& opp, int a, const az_dims&) { az_ttt_apply(own, opp, a); }
    static AZ_HDM bool is_terminal(uint64_t own, uint64_t opp, const az_dims&) { return az_ttt_is_terminal(own, opp); }
    static AZ_HDM int winner(uint64_t own, uint64_t opp, int player, const az_dims&) { return az_ttt_winner(own, opp, player); }
    static AZ_HDM int score(uint64_t own, uint64_t opp) { return az_ttt_score(own, opp); }       // tictactoe.py:127-133
    // tictactoe.py:325, 340: index 3*r + c
    static AZ_HDM int net_index(int a, const az_dims&) { return (a >> 3) * 3 + (a & 7); }
};

#ifdef __CUDACC__
// Warp-cooperative rule queries for kernels that give a whole warp to one game (every lane holds the same position;
// the warp must be converged).  Othello spreads the eight ray directions over lanes (othello_bits.cuh); Connect4's
// rules are a handful of instructions, so every lane just evaluates them.
template <int GAME>
struct az_warp_rules;

template <>
struct az_warp_rules<AZ_GAME_OTHELLO> {
    // legal actions of the side to move; *terminal = neither side has a placement (othello.py:212-219)
    static __device__ __forceinline__ uint64_t legal(uint64_t own, uint64_t opp, const az_dims& d, int lane, bool* terminal) {
        uint64_t other;
        const uint64_t l = az_legal_warp(own, opp, d.bmask, lane, &other);
        if (terminal) *terminal = l == 0ULL && other == 0ULL;
        return l;
    }
    static __device__ __forceinline__ void apply(uint64_t& own, uint64_t& opp, int a, const az_dims&, int lane) {
        az_apply_warp(own, opp, a, lane);
    }
};

template <>
struct az_warp_rules<AZ_GAME_CONNECT4> {
    static __device__ __forceinline__ uint64_t legal(uint64_t own, uint64_t opp, const az_dims& d, int, bool* terminal) {
        if (terminal) *terminal = az_c4_is_terminal(own, opp, d.bmask);
        return az_c4_legal_cols(own, opp, d.cols);
    }
    static __device__ __forceinline__ void apply(uint64_t& own, uint64_t& opp, int a, const az_dims& d, int) {
        az_c4_apply(own, opp, a, d.bmask);
    }
};

template <>
struct az_warp_rules<AZ_GAME_TICTACTOE> {
    static __device__ __forceinline__ uint64_t legal(uint64_t own, uint64_t opp, const az_dims&, int, bool* terminal) {
        if (terminal) *terminal = az_ttt_is_terminal(own, opp);
        return az_ttt_legal(own, opp);
    }
    static __device__ __forceinline__ void apply(uint64_t& own, uint64_t& opp, int a, const az_dims&, int) {
        az_ttt_apply(own, opp, a);
    }
};

// One uniform random playout to the end of the game from (own, opp, player); returns the plies played.
// Random stream: philox4x32_10(ctr = {id_lo, id_hi, ply, 0}, key = seed).x per ply with a choice to make; the k-th
// legal action in ascending order is played with k = mulhi(x, #legal).  A side without a move passes (Othello).
// Replaces MCT.rollout + Board.get_random_move (alphazero/mcts.py:173-180, games/othello.py:191-194,
// games/connect4.py:164-168).
template <int GAME>
__device__ __forceinline__ int az_random_playout_g(uint64_t& own, uint64_t& opp, int& player, const az_dims& d,
                                                   uint64_t id, uint64_t seed) {
    using R = az_rules<GAME>;
    int ply = 0;
    while (!R::is_terminal(own, opp, d)) {
        uint64_t legal = R::legal(own, opp, d);
        int a;
        if (legal == 0ULL) {
            a = R::nth_action(legal, 0);       // Othello: forced pass, no draw consumed
        } else {
            int cnt = __popcll(legal);
            az_u32x4 r = az_philox4x32_10((uint32_t)id, (uint32_t)(id >> 32), (uint32_t)ply, 0u, (uint32_t)seed,
                                          (uint32_t)(seed >> 32));
            a = az_nth_set_bit(legal, (int)az_bounded(r.x, (uint32_t)cnt));
        }
        R::apply(own, opp, a, d);
        player = -player;
        ++ply;
    }
    return ply;
}
#endif
