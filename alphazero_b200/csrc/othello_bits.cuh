// Othello bitboard primitives (device + host inline), sm_100a build.
//
// Board layout: bit index = r*8 + c on an 8-wide canvas, r = row (0 = top row of the reference's grid),
// c = column.  Boards with n in {4,6,8} live in the top-left n x n corner; `az_board_mask(n)` selects it.
// State is side-relative: `own` = discs of the side to move, `opp` = the other side, `player` = +1 (black,
// moves first) or -1 (white), exactly the reference's ids (alphazero/games/othello.py:54-58, 102-109).
//
// Replaces the ray walks of the reference:
//   __get_flips / is_legal_move   alphazero/games/othello.py:141-174
//   get_moves                     alphazero/games/othello.py:176-189   (-> az_legal_mask; 0 means "pass only")
//   play_move                     alphazero/games/othello.py:196-210   (-> az_flips + az_apply)
//   is_game_over / get_winner     alphazero/games/othello.py:212-229   (-> az_is_terminal / az_disc_diff)
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define AZ_HD __host__ __device__ __forceinline__
#else
#define AZ_HD static inline
#endif

#define AZ_NOT_COL0 0xFEFEFEFEFEFEFEFEULL
#define AZ_NOT_COL7 0x7F7F7F7F7F7F7F7FULL
#define AZ_PASS 64  // action index of the pass move on the 8-wide canvas (reference: (n, n))

// Direction ids: 0 E, 1 W, 2 S, 3 N, 4 SE, 5 SW, 6 NE, 7 NW.  One step of every disc in x.
template <int D>
AZ_HD uint64_t az_shift(uint64_t x) {
    if (D == 0) return (x << 1) & AZ_NOT_COL0;
    if (D == 1) return (x >> 1) & AZ_NOT_COL7;
    if (D == 2) return (x << 8);
    if (D == 3) return (x >> 8);
    if (D == 4) return (x << 9) & AZ_NOT_COL0;
    if (D == 5) return (x << 7) & AZ_NOT_COL7;
    if (D == 6) return (x >> 7) & AZ_NOT_COL0;
    return (x >> 9) & AZ_NOT_COL7;
}

AZ_HD uint64_t az_shift_dyn(uint64_t x, int d) {
    switch (d) {
        case 0: return az_shift<0>(x);
        case 1: return az_shift<1>(x);
        case 2: return az_shift<2>(x);
        case 3: return az_shift<3>(x);
        case 4: return az_shift<4>(x);
        case 5: return az_shift<5>(x);
        case 6: return az_shift<6>(x);
        default: return az_shift<7>(x);
    }
}

AZ_HD uint64_t az_board_mask(int n) {
    uint64_t row = (n >= 8) ? 0xFFULL : ((1ULL << n) - 1ULL);
    uint64_t m = 0;
    for (int r = 0; r < n; ++r) m |= row << (8 * r);
    return m;
}

// Dumb7fill along one direction: opponent discs contiguous to `gen` (excluding gen).  A line holds <= 6 of them.
template <int D>
AZ_HD uint64_t az_fill(uint64_t gen, uint64_t opp) {
    uint64_t x = az_shift<D>(gen) & opp;
    x |= az_shift<D>(x) & opp;
    x |= az_shift<D>(x) & opp;
    x |= az_shift<D>(x) & opp;
    x |= az_shift<D>(x) & opp;
    x |= az_shift<D>(x) & opp;
    return x;
}

template <int D>
AZ_HD uint64_t az_legal_dir(uint64_t own, uint64_t opp, uint64_t empty) {
    return az_shift<D>(az_fill<D>(own, opp)) & empty;
}

// Placement squares of the side owning `own` (bit set <=> reference is_legal_move((r,c)) is True).
// 0 <=> the reference's get_moves() returns [pass_move].
AZ_HD uint64_t az_legal_mask(uint64_t own, uint64_t opp, uint64_t board_mask) {
    uint64_t empty = ~(own | opp) & board_mask;
    uint64_t m = az_legal_dir<0>(own, opp, empty);
    m |= az_legal_dir<1>(own, opp, empty);
    m |= az_legal_dir<2>(own, opp, empty);
    m |= az_legal_dir<3>(own, opp, empty);
    m |= az_legal_dir<4>(own, opp, empty);
    m |= az_legal_dir<5>(own, opp, empty);
    m |= az_legal_dir<6>(own, opp, empty);
    m |= az_legal_dir<7>(own, opp, empty);
    return m;
}

template <int D>
AZ_HD uint64_t az_flips_dir(uint64_t mv, uint64_t own, uint64_t opp) {
    uint64_t x = az_fill<D>(mv, opp);
    return (az_shift<D>(x) & own) ? x : 0ULL;
}

// Discs flipped when the side owning `own` plays on the single-bit square `mv` (0 if mv flips nothing).
AZ_HD uint64_t az_flips(uint64_t mv, uint64_t own, uint64_t opp) {
    uint64_t f = az_flips_dir<0>(mv, own, opp);
    f |= az_flips_dir<1>(mv, own, opp);
    f |= az_flips_dir<2>(mv, own, opp);
    f |= az_flips_dir<3>(mv, own, opp);
    f |= az_flips_dir<4>(mv, own, opp);
    f |= az_flips_dir<5>(mv, own, opp);
    f |= az_flips_dir<6>(mv, own, opp);
    f |= az_flips_dir<7>(mv, own, opp);
    return f;
}

AZ_HD uint64_t az_flips_dir_dyn(uint64_t mv, uint64_t own, uint64_t opp, int d) {
    uint64_t x = az_shift_dyn(mv, d) & opp;
    x |= az_shift_dyn(x, d) & opp;
    x |= az_shift_dyn(x, d) & opp;
    x |= az_shift_dyn(x, d) & opp;
    x |= az_shift_dyn(x, d) & opp;
    x |= az_shift_dyn(x, d) & opp;
    return (az_shift_dyn(x, d) & own) ? x : 0ULL;
}

// Plays action `a` (0..63 placement, AZ_PASS pass) for the side to move and swaps sides.
// The caller guarantees legality (checked variants live in the kernels).
AZ_HD void az_apply(uint64_t& own, uint64_t& opp, int a) {
    if (a != AZ_PASS) {
        uint64_t mv = 1ULL << a;
        uint64_t f = az_flips(mv, own, opp);
        own ^= f | mv;
        opp ^= f;
    }
    uint64_t t = own; own = opp; opp = t;
}

// Terminal <=> neither colour has a placement (othello.py:212-219), independent of the side to move.
AZ_HD bool az_is_terminal(uint64_t own, uint64_t opp, uint64_t board_mask) {
    return az_legal_mask(own, opp, board_mask) == 0ULL && az_legal_mask(opp, own, board_mask) == 0ULL;
}

AZ_HD int az_popc64(uint64_t x) {
#ifdef __CUDA_ARCH__
    return __popcll(x);
#else
    return __builtin_popcountll(x);
#endif
}

// get_score(): sum(player * grid) = #own - #opp from the side to move's view (othello.py:133-135).
AZ_HD int az_disc_diff(uint64_t own, uint64_t opp) { return az_popc64(own) - az_popc64(opp); }

// Absolute winner id (+1 black / -1 white / 0 draw), no empties bonus (othello.py:221-229).
AZ_HD int az_winner(uint64_t own, uint64_t opp, int player) {
    int d = az_disc_diff(own, opp);
    return d == 0 ? 0 : (d > 0 ? player : -player);
}

// Index of the k-th (0-based) set bit of m; m must have more than k bits set (popcount descent, no loop).
AZ_HD int az_nth_set_bit(uint64_t m, int k) {
    uint32_t w = (uint32_t)m;
    int pos = 0;
    int c = az_popc64((uint64_t)w);
    if (k >= c) { k -= c; pos = 32; w = (uint32_t)(m >> 32); }
    c = az_popc64((uint64_t)(w & 0xFFFFu));
    if (k >= c) { k -= c; pos += 16; w >>= 16; }
    c = az_popc64((uint64_t)(w & 0xFFu));
    if (k >= c) { k -= c; pos += 8; w >>= 8; }
    c = az_popc64((uint64_t)(w & 0xFu));
    if (k >= c) { k -= c; pos += 4; w >>= 4; }
    c = az_popc64((uint64_t)(w & 0x3u));
    if (k >= c) { k -= c; pos += 2; w >>= 2; }
    c = (int)(w & 1u);
    if (k >= c) { pos += 1; }
    return pos;
}
