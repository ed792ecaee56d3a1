// Connect4 bitboard primitives (device + host inline), sm_100a build.
//
// Same canvas as Othello: bit index = r*8 + c, r = row of the reference's grid (0 = TOP row, discs fall towards
// row height-1), c = column; a width x height board (4 <= width, height <= 8) lives in the top-left corner.
// State is side-relative: `own` = discs of the side to move, `opp` = the other side, player = +1 / -1
// (alphazero/games/connect4.py:58-63).  An action is the column index 0..width-1; there is no pass move.
//
// Replaces the reference's array code:
//   __count_free_rows / get_moves / is_legal_move   alphazero/games/connect4.py:111-113, 145-162
//   play_move                                       alphazero/games/connect4.py:170-182
//   __check_alignment / __get_game_status           alphazero/games/connect4.py:184-244
//   is_game_over / get_winner                       alphazero/games/connect4.py:246-258
#pragma once
#include <stdint.h>

#include "othello_bits.cuh"

#define AZ_COL0 0x0101010101010101ULL

// Cells of a width x height board on the canvas.
AZ_HD uint64_t az_c4_board_mask(int width, int height) {
    uint64_t row = (width >= 8) ? 0xFFULL : ((1ULL << width) - 1ULL);
    uint64_t m = 0;
    for (int r = 0; r < height; ++r) m |= row << (8 * r);
    return m;
}

// Playable columns as a bit mask over column indices (bit c set <=> free_rows[c] > 0, connect4.py:145-149):
// a column is playable iff its top cell (row 0 = bits 0..width-1) is empty.
AZ_HD uint64_t az_c4_legal_cols(uint64_t own, uint64_t opp, int width) {
    uint64_t top = (width >= 8) ? 0xFFULL : ((1ULL << width) - 1ULL);
    return ~(own | opp) & top;
}

// Canvas bit of the cell a disc dropped in column `col` lands on: row free_rows[col]-1, i.e. the lowest empty cell
// (connect4.py:176-180).  The column must be playable.
AZ_HD int az_c4_drop_bit(uint64_t own, uint64_t opp, int col, uint64_t board_mask) {
    uint64_t empties = (AZ_COL0 << col) & board_mask & ~(own | opp);
#ifdef __CUDA_ARCH__
    return 63 - __clzll((long long)empties);
#else
    return 63 - __builtin_clzll(empties);
#endif
}

// Drops a disc of the side to move in column `col` and swaps sides (connect4.py:170-182); legality is the caller's.
AZ_HD void az_c4_apply(uint64_t& own, uint64_t& opp, int col, uint64_t board_mask) {
    uint64_t mv = 1ULL << az_c4_drop_bit(own, opp, col, board_mask);
    uint64_t t = own | mv;
    own = opp;
    opp = t;
}

// True iff `m` holds four aligned discs (rows, columns, both diagonals; connect4.py:184-237).  The column masks
// keep a horizontal/diagonal run from wrapping around the 8-wide canvas.
AZ_HD bool az_c4_has_four(uint64_t m) {
    uint64_t v = m & (m >> 8) & (m >> 16) & (m >> 24);                                  // column
    uint64_t h = m & (m >> 1) & (m >> 2) & (m >> 3) & 0x1F1F1F1F1F1F1F1FULL;            // row, start col <= 4
    uint64_t d = m & (m >> 9) & (m >> 18) & (m >> 27) & 0x1F1F1F1F1F1F1F1FULL;          // (r+1, c+1) diagonal
    uint64_t a = m & (m >> 7) & (m >> 14) & (m >> 21) & 0xF8F8F8F8F8F8F8F8ULL;          // (r+1, c-1) diagonal
    return (v | h | d | a) != 0ULL;
}

// __get_game_status is not None (connect4.py:213-244): an alignment of either colour, or a full board.
AZ_HD bool az_c4_is_terminal(uint64_t own, uint64_t opp, uint64_t board_mask) {
    return az_c4_has_four(own) || az_c4_has_four(opp) || ((own | opp) & board_mask) == board_mask;
}

// Absolute winner id of a finished game (+1 / -1 / 0 draw), connect4.py:248-258.  With legal play only the side
// that just moved (`opp`) can own an alignment; a hand-made grid holding alignments of both colours is resolved
// in favour of `opp` here while the reference returns whichever its scan order meets first.
AZ_HD int az_c4_winner(uint64_t own, uint64_t opp, int player) {
    if (az_c4_has_four(opp)) return -player;
    if (az_c4_has_four(own)) return player;
    return 0;
}
