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
    if (n >= 8) return ~0ULL;
    const uint64_t row = (1ULL << n) - 1ULL;                       // n low columns of a canvas row ...
    return (row * 0x0101010101010101ULL) & ((1ULL << (8 * n)) - 1ULL);   // ... in each of the n low rows
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

#ifdef __CUDACC__
// ------------------------------------------------------------------------------------------------ warp-cooperative forms
// The tree kernels give one warp to one game, so every lane would repeat the same 8-direction ray computation.
// These variants give one DIRECTION to a lane instead and OR the eight results with `redux.sync`: ~1/5 of the
// instructions of az_flips / az_legal_mask.  All lanes must be converged and hold the same (own, opp).
//
// Lanes 0..3 walk E, S, SE, SW with left shifts by 1, 8, 9, 7; lanes 4..7 walk the opposite directions W, N, NW, NE
// as the SAME left shifts on the bit-reversed board (brev maps cell (r, c) to (7-r, 7-c)), so the code is uniform.
// Lanes 8..15 repeat the pattern for a second position (used for "the other side's moves"); lanes 16..31 idle along.
struct az_dirlane {
    uint32_t s;        // left shift of this lane's direction
    uint64_t keep;     // cells that may receive a shifted disc (column wrap guard)
    bool rev;          // lane works on the bit-reversed board
};

__device__ __forceinline__ az_dirlane az_dirlane_of(int lane) {
    const uint32_t k = (uint32_t)lane & 3u;
    az_dirlane d;
    d.s = (0x7981u >> (4u * k)) & 15u;                          // 1, 8, 9, 7
    // column guard, the same byte pattern in both words: all cells (S), not column 7 (SW), not column 0 (E, SE)
    const uint32_t w = k == 1u ? 0xFFFFFFFFu : (k == 3u ? (uint32_t)AZ_NOT_COL7 : (uint32_t)AZ_NOT_COL0);
    d.keep = ((uint64_t)w << 32) | w;
    d.rev = (lane & 4) != 0;
    return d;
}

// OR of x over the lanes with `mine` set, delivered to every lane: two `redux.sync` (low / high word) instead of three
// dependent 64-bit shuffle steps plus a broadcast.
__device__ __forceinline__ uint64_t az_or_lanes(uint64_t x, bool mine) {
    const uint32_t lo = __reduce_or_sync(0xffffffffu, mine ? (uint32_t)x : 0u);
    const uint32_t hi = __reduce_or_sync(0xffffffffu, mine ? (uint32_t)(x >> 32) : 0u);
    return ((uint64_t)hi << 32) | lo;
}

// Discs flipped when the side owning `own` plays the single-bit square `mv` (same value in every lane).
__device__ __forceinline__ uint64_t az_flips_warp(uint64_t mv, uint64_t own, uint64_t opp, int lane) {
    const az_dirlane d = az_dirlane_of(lane);
    const uint64_t o = d.rev ? __brevll(own) : own, p = d.rev ? __brevll(opp) : opp;
    uint64_t x = ((d.rev ? __brevll(mv) : mv) << d.s) & d.keep & p;
#pragma unroll
    for (int i = 0; i < 5; ++i) x |= (x << d.s) & d.keep & p;
    x = ((x << d.s) & d.keep & o) ? x : 0ULL;
    if (d.rev) x = __brevll(x);
    return az_or_lanes(x, lane < 8);            // one position only: lanes 8.. would duplicate lanes 0..7
}

// Placement squares of the side owning `own` (result in every lane) and, through other_out, of the other side.
__device__ __forceinline__ uint64_t az_legal_warp(uint64_t own, uint64_t opp, uint64_t board_mask, int lane,
                                                  uint64_t* other_out) {
    const az_dirlane d = az_dirlane_of(lane);
    const bool second = (lane & 8) != 0;        // lanes 8..15: the roles of own / opp swapped
    uint64_t o = second ? opp : own, p = second ? own : opp;
    uint64_t empty = ~(own | opp) & board_mask;
    if (d.rev) { o = __brevll(o); p = __brevll(p); empty = __brevll(empty); }
    uint64_t x = (o << d.s) & d.keep & p;
#pragma unroll
    for (int i = 0; i < 5; ++i) x |= (x << d.s) & d.keep & p;
    x = (x << d.s) & d.keep & empty;
    if (d.rev) x = __brevll(x);
    // lanes 0..7 hold the directions of legal(own, opp), lanes 8..15 those of legal(opp, own)
    if (other_out) *other_out = az_or_lanes(x, (lane & 24) == 8);
    return az_or_lanes(x, lane < 8);
}

// Plays action `a` for the side to move and swaps sides; warp-cooperative twin of az_apply.
__device__ __forceinline__ void az_apply_warp(uint64_t& own, uint64_t& opp, int a, int lane) {
    if (a != AZ_PASS) {
        const uint64_t mv = 1ULL << a;
        const uint64_t f = az_flips_warp(mv, own, opp, lane);
        own ^= f | mv;
        opp ^= f;
    }
    const uint64_t t = own; own = opp; opp = t;
}
#endif
