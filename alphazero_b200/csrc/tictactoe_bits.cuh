// TicTacToe rules on the engine's 8-wide bitboard canvas (bit r*8 + c, r, c < 3), side-relative (own / opp).
// Replaces the NumPy code of alphazero/games/tictactoe.py:112-184:
//   get_moves (:145-150)      -> az_ttt_legal: the empty cells, ascending canvas order = the reference's row-major order
//   play_move (:157-164)      -> az_ttt_apply
//   is_game_over (:166-171)   -> az_ttt_is_terminal: an alignment of three or a full board
//   get_winner (:173-184)     -> az_ttt_winner: +1 is looked for first, then -1, else draw
//   get_score (:127-133)      -> az_ttt_score: "inf" (127) when the side to move has two on an alignment whose third
//                                cell is free
#pragma once
#include <stdint.h>

#include "othello_bits.cuh"

#define AZ_TTT_BOARD 0x070707ULL
#define AZ_TTT_SCORE_INF 127

// the 8 alignments in the order of get_alignments_sums (rows, columns, diagonal, anti-diagonal)
AZ_HD uint64_t az_ttt_line(int i) {
    return i < 3 ? (0x07ULL << (8 * i)) : i < 6 ? (0x010101ULL << (i - 3)) : i == 6 ? 0x040201ULL : 0x010204ULL;
}

AZ_HD bool az_ttt_has_line(uint64_t x) {
    bool any = false;
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint64_t m = az_ttt_line(i);
        any = any || ((x & m) == m);
    }
    return any;
}

AZ_HD uint64_t az_ttt_legal(uint64_t own, uint64_t opp) { return ~(own | opp) & AZ_TTT_BOARD; }

AZ_HD void az_ttt_apply(uint64_t& own, uint64_t& opp, int a) {
    const uint64_t o = own | (1ULL << a);
    own = opp;          // the other side moves next
    opp = o;
}

AZ_HD bool az_ttt_is_terminal(uint64_t own, uint64_t opp) {
    return az_ttt_has_line(own) || az_ttt_has_line(opp) || ((own | opp) & AZ_TTT_BOARD) == AZ_TTT_BOARD;
}

// absolute winner id (+1 / -1 / 0); `player` = side to move (owner of `own`)
AZ_HD int az_ttt_winner(uint64_t own, uint64_t opp, int player) {
    const uint64_t black = player > 0 ? own : opp, white = player > 0 ? opp : own;
    return az_ttt_has_line(black) ? 1 : (az_ttt_has_line(white) ? -1 : 0);
}

AZ_HD int az_ttt_score(uint64_t own, uint64_t opp) {
    bool win = false;
    #pragma unroll
    for (int i = 0; i < 8; ++i) {
        const uint64_t m = az_ttt_line(i);
        win = win || (az_popc64(own & m) == 2 && (opp & m) == 0ULL);
    }
    return win ? AZ_TTT_SCORE_INF : 0;
}
