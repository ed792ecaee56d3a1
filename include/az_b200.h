/* az_b200.h -- C ABI of libaz_b200.so: the B200 (sm_100a) hot path of t0m1ab/alphazero.
 *
 * The reference has no FFI layer (it is pure Python); these entry points are what a
 * maintainer would bind with ctypes to replace the per-object Python loops named beside each function.
 * Conventions:
 *   - every pointer is a DEVICE pointer owned by the caller unless the name ends in `_h` (host value);
 *   - `stream` is a cudaStream_t passed as void*; functions only enqueue work unless stated otherwise;
 *   - return value: 0 = ok, <0 = error (AZ_ERR_*); the Python layer turns them into ValueError/RuntimeError
 *     under the same conditions as the reference raises;
 *   - boards are side-relative bitboards: own/opp = discs of the side to move / the other side, bit r*8+c on an
 *     8-wide canvas, player = +1 (black, moves first) or -1 (white); action a in 0..63 = cell (a/8, a%8),
 *     64 = pass (the reference's (n, n));
 *   - no allocation, no global state: all storage is caller-provided.
 */
#ifndef AZ_B200_H
#define AZ_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define AZ_OK 0
#define AZ_ERR_ARG (-1)      /* bad argument (n not in {4,6,8}, null pointer, ...) */
#define AZ_ERR_CUDA (-2)     /* a CUDA runtime call failed; az_last_cuda_error() has the text */
#define AZ_ERR_CAPACITY (-3) /* caller-provided scratch too small */

#define AZ_ACTION_PASS 64
#define AZ_CANVAS_ACTIONS 65 /* 64 cells + pass on the 8-wide canvas */

/* ABI version, bumped when a signature changes. */
int az_abi_version(void);
/* Text of the last CUDA error seen by this library on the calling thread ("" if none). */
const char* az_last_cuda_error(void);

/* ---------------------------------------------------------------- Othello board kernels ---------------- */

/* legal[g] = placement squares of the side to move; 0 <=> get_moves() == [pass_move].
 * Replaces OthelloBoard.get_moves / is_legal_move, alphazero/games/othello.py:155-189. */
int az_othello_legal(const uint64_t* own, const uint64_t* opp, int n, int64_t G, uint64_t* legal, void* stream);

/* flips[g] = discs flipped if the side to move plays action[g] (0 for pass or for a non-flipping square).
 * Replaces OthelloBoard.__get_flips over the 8 DIRECTIONS, alphazero/games/othello.py:141-153, 206-208. */
int az_othello_flips(const uint64_t* own, const uint64_t* opp, const uint8_t* action, int n, int64_t G,
                     uint64_t* flips, void* stream);

/* Plays action[g] and swaps sides: (own,opp,player) -> (out_own,out_opp,out_player) (in-place allowed).
 * status[g] = 0 ok, 1 illegal move (board left unchanged) -- the reference raises ValueError there.
 * Replaces OthelloBoard.play_move, alphazero/games/othello.py:196-210. */
int az_othello_apply(const uint64_t* own, const uint64_t* opp, const int8_t* player, const uint8_t* action,
                     int n, int64_t G, uint64_t* out_own, uint64_t* out_opp, int8_t* out_player,
                     uint8_t* status, void* stream);

/* terminal[g] = 1 iff neither colour has a placement; score[g] = #own - #opp (get_score());
 * winner[g] = absolute id +1/-1/0 by disc difference (meaningful when terminal).
 * Replaces is_game_over / get_winner / get_score, alphazero/games/othello.py:133-135, 212-229. */
int az_othello_terminal_score(const uint64_t* own, const uint64_t* opp, const int8_t* player, int n, int64_t G,
                              uint8_t* terminal, int8_t* score, int8_t* winner, void* stream);

/* perft from one position (host values), convention of SURVEY.md 4: a pass is a ply, a finished game counts as
 * one leaf.  Breadth-first on the device down to `depth - dfs_depth`, then one thread per frontier board
 * walks the last dfs_depth plies.  scratch: device buffer of scratch_u64 uint64 used for the frontiers
 * (needs 4 * max_frontier entries).  SYNCHRONOUS (reads frontier sizes back); result in *count_h. */
int az_othello_perft(uint64_t own_h, uint64_t opp_h, int n, int depth, int dfs_depth, uint64_t* scratch,
                     size_t scratch_u64, uint64_t* count_h, void* stream);

/* Uniform random playouts to the end of the game, one thread per playout.  Playout i of root g (global id
 * id = first_id + g*per_root + i) picks, at ply p, the k-th legal placement in ascending bit order with
 * k = mulhi(philox4x32_10(ctr = {id_lo, id_hi, p, 0}, key = {seed_lo, seed_hi}).x, #legal); a side without
 * placement passes.  Outputs per playout: final black/white bitboards, absolute winner, plies played.
 * Replaces MCT.rollout + get_random_move, alphazero/mcts.py:173-180, games/othello.py:191-194
 * (distribution-equal to the reference, stream-equal to oracle/az_oracle.c). */
int az_othello_playouts(const uint64_t* own, const uint64_t* opp, const int8_t* player, int n, int64_t n_roots,
                        int64_t per_root, uint64_t seed, uint64_t first_id, uint64_t* final_black,
                        uint64_t* final_white, int8_t* winner, uint8_t* plies, void* stream);

/* Encodes player*grid for the network: out[g, r*n+c] = +1 own, -1 opp, 0 empty (row-major n x n).
 * dtype: 0 = float32, 1 = bfloat16.  Replaces the torch.tensor(board.player * board.grid) of
 * PolicyValueNetwork.evaluate, alphazero/base.py:363. */
int az_othello_encode(const uint64_t* own, const uint64_t* opp, int n, int64_t G, void* out, int dtype,
                      void* stream);

/* AZ_TREE_SECTION */

#ifdef __cplusplus
}
#endif
#endif /* AZ_B200_H */
