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

#define AZ_GAME_OTHELLO 0  /* alphazero/games/othello.py */
#define AZ_GAME_CONNECT4 1 /* alphazero/games/connect4.py */
#define AZ_GAME_TICTACTOE 2 /* alphazero/games/tictactoe.py */

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

/* ---------------------------------------------------------------- Connect4 board kernels --------------- */
/* Same canvas (bit r*8+c, row 0 = top of the reference's grid), width x height boards with 4 <= width, height <= 8;
 * an action is a column index 0..width-1, there is no pass. */

/* legal[g] = bit mask over COLUMN indices of the playable columns (free_rows[c] > 0).
 * Replaces Connect4Board.get_moves / is_legal_move, alphazero/games/connect4.py:145-162. */
int az_connect4_legal(const uint64_t* own, const uint64_t* opp, int width, int height, int64_t G, uint64_t* legal,
                      void* stream);

/* Drops a disc of the side to move in column action[g] and swaps sides (in-place allowed).
 * status[g] = 0 ok, 1 illegal (column full or out of range; board unchanged) -- the reference raises ValueError.
 * Replaces Connect4Board.play_move, alphazero/games/connect4.py:170-182. */
int az_connect4_apply(const uint64_t* own, const uint64_t* opp, const int8_t* player, const uint8_t* action,
                      int width, int height, int64_t G, uint64_t* out_own, uint64_t* out_opp, int8_t* out_player,
                      uint8_t* status, void* stream);

/* terminal[g] = 1 iff a colour has four aligned discs or the board is full; score[g] = #own - #opp (get_score());
 * winner[g] = absolute id of the aligned colour, 0 for a draw (meaningful when terminal).
 * Replaces __get_game_status / is_game_over / get_winner / get_score, alphazero/games/connect4.py:137-139, 184-258. */
int az_connect4_terminal_score(const uint64_t* own, const uint64_t* opp, const int8_t* player, int width,
                               int height, int64_t G, uint8_t* terminal, int8_t* score, int8_t* winner,
                               void* stream);

/* Uniform random playouts, same random stream convention as az_othello_playouts (k-th playable column in ascending
 * order).  Replaces MCT.rollout + Connect4Board.get_random_move, mcts.py:173-180, connect4.py:164-168. */
int az_connect4_playouts(const uint64_t* own, const uint64_t* opp, const int8_t* player, int width, int height,
                         int64_t n_roots, int64_t per_root, uint64_t seed, uint64_t first_id,
                         uint64_t* final_black, uint64_t* final_white, int8_t* winner, uint8_t* plies, void* stream);

/* Network input player*grid, row-major [G][height][width] (Connect4Net.forward then VIEWS these height*width values
 * as (width, height), connect4.py:399 -- a reinterpretation, not a transpose, reproduced by keeping this order).
 * dtype: 0 = float32, 1 = bfloat16. */
int az_connect4_encode(const uint64_t* own, const uint64_t* opp, int width, int height, int64_t G, void* out,
                       int dtype, void* stream);

/* perft from one position (host values): a finished game counts as one leaf.  Same scheme and scratch contract as
 * az_othello_perft.  SYNCHRONOUS. */
int az_connect4_perft(uint64_t own_h, uint64_t opp_h, int width, int height, int depth, int dfs_depth,
                      uint64_t* scratch, size_t scratch_u64, uint64_t* count_h, void* stream);

/* ---------------------------------------------------------------- MCTS tree kernels -------------------- */

/* One tree node, 32 bytes, 32-byte aligned: a child block is a contiguous run of these, so a warp reads a whole
 * block with two coalesced 128-bit loads per lane.  Replaces alphazero.mcts.Node (mcts.py:8-46).
 *   Q  running mean reward of the edge into the node, from the viewpoint of the player who moved into it,
 *      float64 updated as (N*Q + r)/(N+1) exactly as mcts.py:220;
 *   P  prior (float32, as the reference's np.float32 until Dirichlet noise);  N visit count. */
typedef struct az_node {
    double Q;
    float P;
    uint32_t N;
    uint32_t first_child; /* index of the first child inside the game's pool half; 0 = no child block */
    uint8_t n_children;
    uint8_t move;         /* action that led here (Othello: 0..63, 64 = pass; Connect4: column) */
    uint8_t flags;        /* AZ_NODE_* */
    uint8_t reserved0;
    uint32_t vloss;       /* throughput mode (az_tree.n_leaves >= 1): descents in flight through this node */
    float W;              /* throughput mode: float32 sum of the backed-up rewards */
} az_node;

#define AZ_NODE_EVAL 1u     /* children allocated with priors (reference: node._children_probs is not None) */
#define AZ_NODE_EXPANDED 2u /* children are part of the tree (reference: len(node.children) > 0) */
#define AZ_NODE_NOISED 4u   /* Dirichlet noise applied to the children's priors (node.has_dirichlet_noise) */

/* leaf_kind values (per game, per step) */
#define AZ_LEAF_IDLE 0
#define AZ_LEAF_NN 1        /* evaluate with the network, expand, back up */
#define AZ_LEAF_PRIME 2     /* fresh root: evaluate + expand only, no visit (mcts.py:231-233) */
#define AZ_LEAF_ROLLOUT 3   /* random playout from the leaf board, back up the winner */
#define AZ_LEAF_DRAW 4      /* finished game at the leaf: back up 0 */
#define AZ_LEAF_BLACK 5     /* finished game, black (+1) won */
#define AZ_LEAF_WHITE 6     /* finished game, white (-1) won */

#define AZ_STATUS_OVERFLOW 1u /* node pool of this game is full: children could not be allocated */
#define AZ_STATUS_DEPTH 2u    /* path longer than max_depth */
#define AZ_STATUS_ILLEGAL 4u  /* az_tree_advance got an illegal action for this game */

#define AZ_MODE_ROLLOUT 0
#define AZ_MODE_NEURAL 1

/* Descriptor of G concurrent game trees.  All pointers are device pointers owned by the caller. */
typedef struct az_tree {
    az_node* nodes;        /* [G][2][capacity]  per-game pool, two halves (re-rooting compacts into the other) */
    uint8_t* half;         /* [G] active half */
    uint32_t* n_alloc;     /* [G] nodes used in the active half (root is index 0) */
    uint64_t* root_own;    /* [G] root position, side-relative */
    uint64_t* root_opp;    /* [G] */
    int8_t* root_player;   /* [G] +1 / -1 */
    uint8_t* active;       /* [G] 1 = the game takes part in search / advance */
    uint32_t* sims_done;   /* [G] simulations backed up since the search started (MCT.n_rollouts) */
    uint32_t* rng_ctr;     /* [G] per-game random event counter */
    uint32_t* status;      /* [G] AZ_STATUS_* bits, sticky */
    uint32_t* leaf_node;   /* [G] */
    uint8_t* leaf_kind;    /* [G] AZ_LEAF_* */
    uint64_t* leaf_own;    /* [G] leaf position, side-relative */
    uint64_t* leaf_opp;    /* [G] */
    int8_t* leaf_player;   /* [G] */
    uint32_t* path;        /* [G][max_depth] node indices root..leaf */
    uint32_t* path_len;    /* [G] */
    void* enc;             /* [G][rows*cols] network input, player*grid row-major (written by select) */
    float* probs;          /* [G][A] network probabilities exp(log_softmax); A = n*n+1 (pass last) / width */
    float* value;          /* [G] neural: network value, side to move's view; rollout: absolute winner id */
    uint64_t seed;
    int32_t G;
    int32_t n;             /* Othello: board size 4/6/8; Connect4: width (columns) */
    int32_t capacity;      /* nodes per half */
    int32_t max_depth;
    int32_t mode;          /* AZ_MODE_* */
    int32_t tie_random;    /* 0: ties -> lowest action index (parity mode); 1: uniform among ties (fair_max) */
    int32_t enc_dtype;     /* 0 float32, 1 bfloat16 */
    float dirichlet_alpha; /* <= 0: no noise */
    float dirichlet_eps;
    int32_t game;          /* AZ_GAME_* */
    int32_t height;        /* Connect4: rows; ignored for Othello */
    int32_t logits_ld;     /* row pitch of `logits` in elements (>= A + 1) */
    const void* logits;    /* optional, neural mode: bf16 [G][logits_ld] raw network outputs (policy logits 0 .. A-1, value
                              logit at A).  When not NULL az_tree_step takes priors = softmax(logits[0 .. A)) and
                              value = tanh(logits[A]) -- the arithmetic of az_net_heads, same bits -- from here and
                              ignores probs / value: the evaluator then needs no output-transform launch. */
    int32_t n_leaves;      /* 0: parity mode, the reference's one leaf per game per step (mcts.py:254-269).
                              L >= 1: throughput mode (neural only) -- up to L leaves in flight per game and step under
                              virtual loss; every per-leaf array (leaf_*, path, path_len, enc, probs, value, logits) then
                              has G*L entries, slot l*G + g = descent l of game g.  L = 1 is bit-identical to parity. */
    int32_t reserved;
} az_tree;

/* Fresh trees (reference: MCT() / player.reset()) for the games with mask[g] != 0 (mask NULL = all):
 * n_alloc = 1, zeroed root node, leaf_kind idle, sims_done 0.  Root boards are NOT touched. */
int az_tree_reset(const az_tree* t, const uint8_t* mask, void* stream);

/* One batched step of MCT.__search_iter (mcts.py:226-252) for every active game, one warp per game:
 *   do_finish: consume the previous leaf: children + priors from probs (get_normalized_probs,
 *              othello.py:384-402), back_propagate (mcts.py:197-223), sims_done++;
 *   do_select: if sims_done < n_sim: root priming / Dirichlet noise / select_node (mcts.py:127-171), then write the
 *              leaf (kind, board, path) and its network encoding.
 * Between two calls the caller fills probs/value for the leaves (network, az_eval_hash or az_tree_rollout). */
int az_tree_step(const az_tree* t, int n_sim, int do_finish, int do_select, void* stream);

/* Rollout-mode leaf evaluation: value[g] = winner of one uniform random playout from the leaf board for every
 * game whose leaf_kind is AZ_LEAF_ROLLOUT (MCT.rollout, mcts.py:173-180). */
int az_tree_rollout(const az_tree* t, void* stream);

/* Root statistics as the reference exposes them through root.children (mcts.py:88-116): for every canvas action a,
 * N[g][a] = visits of the child (-1 if the root has no such child or is not expanded), Q, P likewise;
 * root_N[g] = root visits.  Outputs may be NULL. */
int az_tree_root_stats(const az_tree* t, int32_t* N, double* Q, float* P, uint32_t* root_N, void* stream);

/* MCT.get_action_probs (mcts.py:95-116) + the sampling of MCTSPlayer.get_move (players.py:184-189):
 * temp[g] == 0 -> argmax of visits (ties per tie_random), pi one-hot; else pi ~ N^(1/temp), action sampled.
 * action[g] = chosen canvas action; pi[g][n*n+1] in network order (to_neural_output, othello.py:404-412). */
int az_tree_choose(const az_tree* t, const float* temp, uint8_t* action, float* pi, void* stream);

/* Plays action[g] (255 = none) on the root board of every active game and re-roots its tree like
 * MCT.change_root (mcts.py:118-125): keep the child's subtree if the root was expanded, else a fresh root.
 * The kept subtree is compacted breadth-first into the other pool half.  Illegal actions set
 * AZ_STATUS_ILLEGAL and leave the game untouched. */
int az_tree_advance(const az_tree* t, const uint8_t* action, void* stream);

/* One ply of batched self-play for every active game (AlphaZeroTrainer.self_play's inner loop, trainer.py:236-262, for
 * all games at once): records the position and the policy target of the ply, chooses the move like az_tree_choose with
 * temperature `temp` for all games (schedulers.py), plays it and re-roots like az_tree_advance, and retires the games
 * that are over (active[g] = 0, winner[game] = absolute id of the winner, othello.py:221-229).
 *   slot_game  int32 [G] game id of every slot of this tree batch, or NULL (slot = game); the arrays below are indexed
 *              by game id
 *   h_own, h_opp, h_player, h_live   this ply's row of the history: root position before the move, 1 in h_live
 *              (identity mapping: 0 for inactive slots; with slot_game the caller pre-zeroes the row)
 *   h_pi       this ply's policy targets, [games][action_size] in network order
 *   counters   int64 [8], caller-zeroed: [0] live games after the ply, [1] the largest re-rooted tree among them
 *              (nodes), [2] sum of sims_done of the moved games, [3] moves played, [4] the largest tree before the move */
int az_tree_move(const az_tree* t, float temp, const int32_t* slot_game, uint64_t* h_own, uint64_t* h_opp,
                 int8_t* h_player, uint8_t* h_live, float* h_pi, int8_t* winner, int64_t* counters, void* stream);

/* Deterministic hash evaluator (test double for PolicyValueNetwork.evaluate, see tests/hash_eval.py):
 * probs[g][a] = q/4096 for a < action_size, value[g] in {-256..255}/256 from (own, opp, salt). */
int az_eval_hash(const uint64_t* own, const uint64_t* opp, int action_size, int64_t G, uint64_t salt, float* probs,
                 float* value, void* stream);

/* ---------------------------------------------------------------- search-free players ------------------ */

/* GreedyPlayer.get_move for G positions (alphazero/players.py:98-123): every legal move is played on a copy, its value
 * is -get_score() of the resulting board (the score of the side that just moved), the best value wins; ties are broken
 * uniformly at random (utils.fair_max, utils.py:28-34) with Philox4x32-10 keyed by (seed; game index, ctr) when
 * tie_random != 0, else the lowest action index wins.  game: AZ_GAME_*; n / height as in az_tree (Othello: board size;
 * Connect4: width, height).  mask (optional): games with mask[g] == 0 get action 255.  action: uint8 [G] on the
 * canvas action space (Othello pass = 64 when there is no placement).  ties (optional): uint64 [G], the set of
 * maximising actions (0 for a pass / masked game). */
int az_greedy_move(int game, const uint64_t* own, const uint64_t* opp, const uint8_t* mask, int n, int height, int64_t G,
                   uint64_t seed, uint32_t ctr, int tie_random, uint8_t* action, uint64_t* ties, void* stream);

/* ---------------------------------------------------------------- fused network kernels ---------------- */

/* Bytes of the packed trunk weights expected by az_net_trunk (built by alphazero_b200/fused_net.py:
 * pack_trunk_weights), BatchNorm folded, bf16 in the tcgen05 K-major operand layout:
 *   [conv2 | conv3 | conv4: 3 kernel rows x [ci chunk 4][(kernel column, co) 96][ci % 8]]   3 x 18 432 bytes
 *   [conv1: [k chunk 2][co 32][8], taps 0..8 then zeros]                                       1 024 bytes
 *   [biases of conv1..conv4: float32 [4][32]]                                                    512 bytes
 *   [conv2 | conv3 | conv4 once more with kernel row and column exchanged:
 *    3 kernel columns x [ci chunk 4][(kernel row, co) 96][ci % 8]]                          3 x 18 432 bytes
 * (the last section is what k_net_trunk_rows stages by one bulk copy). */
int az_net_wblob_bytes(void);

/* Fused conv1..conv4 (+ folded BatchNorm + ReLU) of OthelloNet.forward / Connect4Net.forward
 * (alphazero/games/othello.py:366-374, connect4.py:397-403) for G positions given as side-relative bitboards in IMAGE
 * layout: bit i*8 + j = cell (i, j) of the rows x cols image the first convolution sees (the network input
 * player*grid, base.py:363, is formed on the fly).  Othello: the image is the board, (rows, cols) = (8, 8) or (6, 6).
 * Connect4 7x6: Connect4Net views the (height, width) grid as a (width, height) image, (rows, cols) = (7, 6); make
 * the bitboards with az_connect4_image_bits.
 * features: bf16 [G][(rows-4)*(cols-4) * 32], position-major / channel-minor (fc1's columns are permuted to match);
 * with features_packed != 0 the same values in the packed operand layout az_net_fcstack reads:
 * [ceil(G/128)][K/64][8][128][8] bf16, element (g, k) at tile g/128, chunk k/64, plane (k%64)/8, row g%128, k%8.
 * tcgen05 implicit GEMM, activations resident in shared memory.
 * debug (optional): bf16 [ceil(G / boards_per_group)][768][32] dump of the activations after conv<debug_layer>
 * (az_net_trunk_boards_per_group gives boards_per_group).
 * 8 x 8 images in batches of 1024 positions and more are computed by k_net_trunk_rows (csrc/trunk_rows_kernels.cu: tile =
 * one board row of 14 boards, kernel column = a 16-byte shift of the A operand, kernel row = the N dimension, one TMEM
 * accumulator per output tile; the same results within bf16 rounding, 18.5 against 22.6 us per 4096 positions) or, when
 * there are more groups of 14 boards than SMs, by k_net_trunk_rows2<8, 8> (csrc/trunk_rows2_kernels.cu: the same formulation
 * with two CTAs per SM, bit-identical to k_net_trunk_rows, 16.5 us; bit 2 of features_packed or AZ_B200_TRUNK_ROWS1=1
 * keeps k_net_trunk_rows).  k_net_trunk_rows2<7, 6> (groups of 18 boards) takes Connect4's 7 x 6 image above 2664
 * positions (debug dump as for 8 x 8 with 7 lanes per board; bit 1 or 2 keeps the first kernel).  Its
 * debug dump is bf16 [ceil(G/14)][1024][32], row = board_row * 128 + (g % 14) * 9 + column.  Bit 1 of features_packed
 * (or AZ_B200_TRUNK_V1=1 in the environment) keeps the first kernel for such a batch too. */
int az_net_trunk(const uint64_t* own, const uint64_t* opp, int rows, int cols, int64_t G, const void* wblob,
                 void* features, int features_packed, void* debug, int debug_layer, void* stream);

/* Boards per group az_net_trunk uses for G images of `rows` rows on the current device (> 0; the layout of its
 * optional debug dump depends on it: group = g / boards, row = (g % boards) * (rows+1)*8 + (1+i)*8 + j). */
int az_net_trunk_boards_per_group(int rows, int64_t G);

/* Image-layout bitboards for az_net_trunk from canvas bitboards of a width x height Connect4 grid: image cell (i, j)
 * (i < width, j < height) = flat element i*height + j of the row-major grid -- the .view(-1, 1, width, height) of
 * Connect4Net.forward, alphazero/games/connect4.py:399 (a reinterpretation, not a transpose). */
int az_connect4_image_bits(const uint64_t* own, const uint64_t* opp, int width, int height, int64_t G,
                           uint64_t* own_img, uint64_t* opp_img, void* stream);

/* probs[g][0 .. A) = softmax(logits[g][0 .. A)), value[g] = tanh(logits[g][A]); A = action_size <= 96; logits bf16,
 * row stride ld.  Replaces log_softmax + exp + tanh of the network forward / PolicyValueNetwork.predict
 * (othello.py:382, connect4.py:411, base.py:350-355). */
int az_net_heads(const void* logits, int ld, int action_size, int64_t G, float* probs, float* value, void* stream);

/* The two output layers and predict's exp(log_softmax) / tanh in one kernel: logits = h2 x Wh^T + bh (tcgen05 GEMM,
 * fp32 logits that never reach memory), probs[g][0 .. A) = softmax(logits[0 .. A)), value[g] = tanh(logits[A]).
 * h2: bf16 [G][K] (K a multiple of 32, <= 512).  wblob: az_net_headgemm_blob_bytes(K, NH) bytes =
 * bf16 [K/8][NH][8] with blob[kc][n][j] = Wh[n][kc*8 + j] (rows 0 .. A-1 = fc_probs, row A = fc_value, the rest zero),
 * then float bias[NH]; NH a multiple of 16 with A + 1 <= NH <= 96.
 * Replaces fc_probs / fc_value + log_softmax / tanh (othello.py:361-362, 379-382; connect4.py:389-390, 408-411) and
 * PolicyValueNetwork.predict (base.py:350-355). */
int64_t az_net_headgemm_blob_bytes(int K, int NH);
int az_net_headgemm(const void* h2, int K, const void* wblob, int NH, int action_size, int64_t G, float* probs,
                    float* value, void* stream);

/* The fully connected stack fc1 + BN + ReLU -> fc2 + BN + ReLU -> fc_probs || fc_value of OthelloNet / Connect4Net
 * (othello.py:352-362, 374-382; connect4.py:381-390, 403-411; eval-mode BatchNorm1d folded, dropout off) as ONE
 * tcgen05 kernel over thread-block clusters of RT x NS CTAs (RT row tiles of 128 positions x NS column slices,
 * RT*NS <= 8; operands by multicast bulk async copies, h1 / h2 exchanged through `h1` / `h2` scratch that stays in
 * the L2).  Replaces the three library GEMMs of the first fused evaluator.
 *   x       bf16 packed operand layout [T][K1/64][8][128][8] (what az_net_trunk writes with features_packed = 1),
 *           T = ceil(G/128) rounded up to a multiple of RT
 *   wblob   az_net_fcstack_plan(...)[13] bytes: per layer the weights [NS][K/kch][kch/8][N/NS][8] bf16
 *           (blob[c][kc][p][n][j] = W[c*N/NS + n][kc*kch + 8p + j]; heads: slice c holds outputs c*hpc .. c*hpc+hpc-1,
 *           zero padded to a multiple of 16 rows), then float32 biases in the same column order
 *   h1, h2  scratch, T*128*N1*2 and T*128*N2*2 bytes
 *   logits  bf16 [G][ld], columns 0 .. n_out-1 written (n_out = action_size + 1: policy logits, then the value logit)
 * az_net_fcstack_plan fills out[15]: K chunk per layer [0..2], columns per CTA per layer [3..5], hpc [6], byte offsets
 * of the weights [7..9] and biases [10..12], blob bytes [13], kernel variant [14]; AZ_ERR_ARG when there is no mapping
 * for the shape (needs N1/NS <= 256, N2/NS <= 128, ceil(n_out/NS) <= 32, K1 % 64 == 0 or K1 in {16, 32}).
 * Variant 2 (k_net_fcstack2; chosen for RT = 1, NS = 4, K1 % 64 == 0, K1 <= 512, N1 % 512 == 0, N2 % 128 == 0,
 * n_out <= 80 -- OthelloNet): fc1 in two passes of PW = N1/8 columns, so that the h1 exchange of pass 0 runs under the
 * MMAs of pass 1; the heads split over K with the fp32 partial sums reduced through distributed shared memory, so
 * that h2 never leaves the SM (`h2` unused).  Its blob: fc1 [8][K1/64][8][PW][8] (pass-slice 2c + h = columns
 * (2c + h) PW ..), fc2 [4][N1/64][8][N2/4][8], heads [4][N2/32][n3][8] (K slice c, n3 = n_out rounded up to 16, zero
 * padded), then the biases in natural column order; `ld` must be even.
 * Variant 3 (k_net_fcstack3; RT = 2, NS = 4, the shapes of variant 2 with 1536 n3 + 256 <= 256 N2): the same schedule
 * on CTA pairs (tcgen05.mma.cta_group::2, M = 256: two row tiles share a column slice's weights, each CTA stages half
 * of its columns); the four pairs of a row-tile pair exchange h1 and the heads' partial sums through `h1` / `h2` and
 * gpu-scope counters in the last 256 bytes of every row tile's 128 x N2 x 2 bytes of `h2`.  THOSE COUNTERS MUST BE ZERO
 * before the first launch (allocate `h2` zeroed and let nothing else write it; the kernel leaves them zero); a launch
 * that finds them dirty traps after ~2 s instead of hanging.  Its blob: every column group of variant 2 once more in
 * two halves (fc1 [16][K1/64][8][PW/2][8], fc2 [8][N1/64][8][N2/8][8], heads [4][2][N2/32][n3/2][8]).  Measured slower
 * than variant 2 (profiles/r2_fcstack_experiments.txt); the evaluator uses RT = 1. */
int az_net_fcstack_plan(int K1, int N1, int N2, int n_out, int RT, int NS, int64_t* out);
/* Measurement hook: while buf != NULL every az_net_fcstack launch writes per-CTA SM-clock stamps of its pipeline phases
 * to buf ([grid][32] int64, device memory; layout in csrc/fcstack_kernels.cu / tools/fcstack_probe.py). */
int az_net_fcstack_debug(void* buf);
int az_net_fcstack(const void* x, const void* wblob, void* h1, void* h2, void* logits, int ld, int n_out, int64_t G,
                   int K1, int N1, int N2, int RT, int NS, void* stream);
/* az_net_fcstack for callers that cannot allocate device memory (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Same arguments and results as az_net_fcstack minus h1 / h2: the scratch is a static device array inside the library
 * (6 MB = 2048 positions at N1 + N2 = 1536, held by every GPU that loads the library until the process exits); the
 * call allocates and frees nothing.  Shapes whose h1 + h2 (ceil(G/128) rounded up to RT, x 128 x (N1 + N2) x 2 bytes)
 * exceed the array return AZ_ERR_CAPACITY with the sizes in az_last_cuda_error().  All calls on a device share the
 * array (it keeps the last call's contents; the only words that have to start at zero, the counters of plan variant
 * 3, are cleared by a memset node in front of the launch): calls that overlap on the GPU --
 * different streams, or graphs replayed concurrently -- race; serialise them or use az_net_fcstack. */
int az_net_fcstack_static(const void* x, const void* wblob, void* logits, int ld, int n_out, int64_t G, int K1, int N1,
                          int N2, int RT, int NS, void* stream);

/* ---------------------------------------------------------------- TicTacToe ----------------------------- */
/* 3 x 3 board on the same canvas (bit r*8+c, r, c < 3); an action is the canvas cell, there is no pass; the policy
 * vector index of cell (r, c) is 3*r + c (alphazero/games/tictactoe.py:325, 340).  The tree kernels take
 * az_tree.game = AZ_GAME_TICTACTOE with n = height = 3. */

/* legal[g] = mask of the empty cells.  Replaces TicTacToeBoard.get_moves / is_legal_move, tictactoe.py:139-150. */
int az_tictactoe_legal(const uint64_t* own, const uint64_t* opp, int64_t G, uint64_t* legal, void* stream);

/* Places a piece of the side to move on action[g]; status 1 = occupied cell or not a cell (the reference raises
 * ValueError), board unchanged.  Replaces TicTacToeBoard.play_move, tictactoe.py:157-164. */
int az_tictactoe_apply(const uint64_t* own, const uint64_t* opp, const int8_t* player, const uint8_t* action,
                       int64_t G, uint64_t* out_own, uint64_t* out_opp, int8_t* out_player, uint8_t* status,
                       void* stream);

/* terminal = an alignment of three or a full board; score = 127 (the reference's float("inf")) when the side to move
 * has two on an alignment whose third cell is free, else 0; winner = absolute id (0 when live or drawn).
 * Replaces is_game_over / get_score / get_winner, tictactoe.py:127-133, 166-184. */
int az_tictactoe_terminal_score(const uint64_t* own, const uint64_t* opp, const int8_t* player, int64_t G,
                                uint8_t* terminal, int8_t* score, int8_t* winner, void* stream);

/* Network input player*grid, [G][3][3]; dtype 0 = float32, 1 = bfloat16 (base.py:363). */
int az_tictactoe_encode(const uint64_t* own, const uint64_t* opp, int64_t G, void* out, int dtype, void* stream);

/* Uniform random playouts (same random stream convention as az_othello_playouts).  Replaces MCT.rollout +
 * TicTacToeBoard.get_random_move, mcts.py:173-180, tictactoe.py:152-155. */
int az_tictactoe_playouts(const uint64_t* own, const uint64_t* opp, const int8_t* player, int64_t n_roots,
                          int64_t per_root, uint64_t seed, uint64_t first_id, uint64_t* final_black,
                          uint64_t* final_white, int8_t* winner, uint8_t* plies, void* stream);

/* TicTacToeNet.forward + predict (tictactoe.py:302-316, base.py:350-355) for G positions given as bitboards, fp32:
 * probs[g][9] = softmax, value[g] = tanh.  weights: az_tictactoe_net_floats() float32 values
 * W1[9][9] b1[9] W2[9][9] b2[9] Wp[9][9] bp[9] Wv[9] bv with the eval-mode BatchNorm1d layers folded into fc1 / fc2. */
int az_tictactoe_net_floats(void);
int az_tictactoe_net(const uint64_t* own, const uint64_t* opp, int64_t G, const float* weights, float* probs,
                     float* value, void* stream);

/* ---------------------------------------------------------------- training samples ---------------------- */

/* A sample of the self-play memory (alphazero/trainer.py:30-78, after Sample.normalize): own/opp = bitboards of
 * the side to move / the other side (state * player), pi float32[action_size] in network order, z = outcome *
 * player, episode_idx, move_idx.  All arrays are SoA over S samples.
 *
 * Symmetry augmentation of AlphaZeroTrainer.self_play (trainer.py:275-284; strategies games/registers.py:37-50):
 * every sample with move_idx >= min_move_idx (the reference uses 2) gets az_augment_twins(game) twins --
 * Othello: reflect_h, rot90, reflect_h+rot90, rot180, reflect_h+rot180, rot270, reflect_h+rot270 (the reference's
 * append order); Connect4: reflect_h only.  Boards follow np.flip(axis=1) / np.rot90(k) (trainer.py:91,111), pi
 * follows reflect/rotate_neural_output (othello.py:414-450: board part transformed, pass entry kept;
 * connect4.py:437-445: columns reversed).
 *
 *   az_augment_plan   ranks the eligible samples into `scratch` (az_augment_scratch_bytes(S) bytes) and writes their
 *                     number E to *total (device int64).
 *   az_augment_apply  writes the S originals to out[0 .. S) (transf = 0) and the twins of the e-th eligible sample to
 *                     out[S + e*T + k], k = 0 .. T-1, transf = k + 1 -- the order of the reference's memory list.
 *                     The out arrays hold S + E*T samples. */
int64_t az_augment_scratch_bytes(int64_t S);
int az_augment_twins(int game);
int az_augment_plan(const int32_t* move_idx, int64_t S, int min_move_idx, void* scratch, int64_t* total,
                    void* stream);
int az_augment_apply(const uint64_t* own, const uint64_t* opp, const float* pi, const int8_t* z,
                     const int32_t* episode_idx, const int32_t* move_idx, int64_t S, int game, int rows, int cols,
                     int action_size, const void* scratch, uint64_t* out_own, uint64_t* out_opp, float* out_pi,
                     int8_t* out_z, int32_t* out_episode, int32_t* out_move, uint8_t* out_transf, void* stream);

/* The self-play history of G games over T plies (rows written by az_tree_move: position before the move, policy target,
 * live flag) as the training memory of AlphaZeroTrainer.self_play (trainer.py:241-265): sample (game g, ply t) at index
 * offset[g] + t -- offset = exclusive prefix sum of the games' lengths, so the samples are ordered by episode, then
 * move -- with z = winner[g] * player (Sample.normalize, trainer.py:74-78), episode_idx = g, move_idx = t. */
int az_samples_pack(const uint64_t* h_own, const uint64_t* h_opp, const int8_t* h_player, const uint8_t* h_live,
                    const float* h_pi, const int8_t* winner, const int64_t* offset, int T, int64_t G, int action_size,
                    uint64_t* own, uint64_t* opp, float* pi, int8_t* z, int32_t* episode_idx, int32_t* move_idx,
                    void* stream);

/* One mini-batch of AlphaZeroTrainer._batch_generator (trainer.py:288-322): for b < B, input[b] = the float32
 * rows x cols grid of sample index[b] (own = +1, opp = -1), pi_out[b] = its pi, z_out[b][0] = its outcome.
 * An index outside [0, S) sets *bad_index (device int32, caller-zeroed) to 1 and leaves row b untouched. */
int az_samples_batch(const uint64_t* own, const uint64_t* opp, const float* pi, const int8_t* z,
                     const int64_t* index, int64_t S, int64_t B, int rows, int cols, int action_size, float* input,
                     float* pi_out, float* z_out, int32_t* bad_index, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AZ_B200_H */
