// Batched MCTS over thousands of concurrent games: SoA/AoS node pools in HBM, one warp per game.  sm_100a only.
//
// What the reference does with a pointer tree of Python objects, one simulation at a time
// (alphazero/mcts.py:8-269), happens here for G games per launch:
//   * a game's nodes live in its own pool slice; the children of a node are one contiguous block of 32-byte
//     records, so selecting among them is two coalesced 128-bit loads per lane (lane j <-> child j), a PUCT/UCT
//     evaluation in float64 (bit-identical to the reference's NumPy arithmetic) and a warp shuffle arg-max;
//   * the board is replayed in registers along the path (16 B of state, ~400 integer instructions per ply);
//   * the leaf is handed to the evaluator through flat [G] buffers (board, network encoding) and the result
//     (probs, value) is consumed by the next launch, which also backs up along the recorded path.
// There is exactly one warp per game and at most one leaf in flight per game, which is the reference's sequential
// order (visit counts are identical under a deterministic evaluator); no atomics are needed because a pool is
// only ever touched by its own warp.
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <math.h>
#include <stdint.h>

#include "../../include/az_b200.h"
#include "common.cuh"
#include "othello_bits.cuh"
#include "philox.cuh"
#include "playout.cuh"
#include "rules.cuh"
#include "net_heads.cuh"

static_assert(sizeof(az_node) == 32, "az_node must be 32 bytes");

#define AZ_FULL 0xffffffffu
#define AZ_WARPS_PER_CTA 4

struct NodeRec {
    double Q;
    float P;
    uint32_t N;
    uint32_t first_child;
    uint32_t meta;  // n_children | move << 8 | flags << 16
    uint32_t vloss; // descents in flight through this node (throughput mode; always 0 in parity mode)
    float W;        // float32 sum of the backed-up rewards (throughput mode)
    __device__ __forceinline__ int n_children() const { return (int)(meta & 0xffu); }
    __device__ __forceinline__ int move() const { return (int)((meta >> 8) & 0xffu); }
    __device__ __forceinline__ uint32_t flags() const { return (meta >> 16) & 0xffu; }
};

__device__ __forceinline__ NodeRec az_load_node(const az_node* p) {
    const uint4* q = reinterpret_cast<const uint4*>(p);
    uint4 a = q[0], b = q[1];
    NodeRec r;
    r.Q = __hiloint2double((int)a.y, (int)a.x);
    r.P = __uint_as_float(a.z);
    r.N = a.w;
    r.first_child = b.x;
    r.meta = b.y;
    r.vloss = b.z;
    r.W = __uint_as_float(b.w);
    return r;
}

__device__ __forceinline__ void az_store_node(az_node* p, double Q, float P, uint32_t N, uint32_t first_child,
                                              uint32_t meta) {
    uint4* q = reinterpret_cast<uint4*>(p);
    uint4 a, b;
    a.x = (uint32_t)__double2loint(Q);
    a.y = (uint32_t)__double2hiint(Q);
    a.z = __float_as_uint(P);
    a.w = N;
    b.x = first_child;
    b.y = meta;
    b.z = 0u;
    b.w = 0u;
    q[0] = a;
    q[1] = b;
}

__device__ __forceinline__ uint32_t az_make_meta(int n_children, int move, uint32_t flags) {
    return (uint32_t)n_children | ((uint32_t)move << 8) | (flags << 16);
}

// The selection only follows the winner's links (N, child block, move / flags): its Q and P are not broadcast.
__device__ __forceinline__ NodeRec az_bcast_node(const NodeRec& r, int src) {
    NodeRec o;
    o.Q = 0.0;
    o.P = 0.f;
    o.N = __shfl_sync(AZ_FULL, r.N, src);
    o.first_child = __shfl_sync(AZ_FULL, r.first_child, src);
    o.meta = __shfl_sync(AZ_FULL, r.meta, src);
    return o;
}

__device__ __forceinline__ double az_warp_sum(double v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(AZ_FULL, v, off);
    return v;
}
__device__ __forceinline__ float az_warp_sum(float v) {
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(AZ_FULL, v, off);
    return v;
}

// Random 32-bit draw for (game, event, lane/sub-event).
__device__ __forceinline__ az_u32x4 az_draw4(uint64_t seed, uint32_t g, uint32_t event, uint32_t sub, uint32_t tag) {
    return az_philox4x32_10(g, event, sub, tag, (uint32_t)seed, (uint32_t)(seed >> 32));
}

// Writes the network encoding of a side-relative position: +1 own, -1 opp, 0 empty, row-major n x n.
// Out of line with scalar arguments: only the generic (torch module) evaluator asks for it, the fused network path reads
// the leaf bitboards, and its runtime divisions would otherwise sit inside the tree step's instruction stream.
__device__ __noinline__ void az_encode_leaf(void* enc, int enc_dtype, int rows, int cols, int g, int lane, uint64_t own,
                                            uint64_t opp) {
    const int cells = rows * cols;
    for (int k = lane; k < cells; k += 32) {
        int bit = (k / cols) * 8 + (k % cols);
        float v = (float)((int)((own >> bit) & 1ULL) - (int)((opp >> bit) & 1ULL));
        if (enc_dtype == 0) reinterpret_cast<float*>(enc)[(size_t)g * cells + k] = v;
        else reinterpret_cast<__nv_bfloat16*>(enc)[(size_t)g * cells + k] = __float2bfloat16(v);
    }
}

// ------------------------------------------------------------------------------------------------ finish

// Allocates the children block of `leaf` for all legal moves of (own, opp) and fills the priors.
// neural: P = probs[idx]/sum over the legal moves in float32, uniform when sum < 1e-6
//         (OthelloNet.get_normalized_probs, othello.py:384-402);  rollout: P = 0 (mcts.py:152-154).
// Returns the first child index (0 on pool overflow).
// What az_alloc_children needs from memory, loaded by the caller ahead of time (k_tree_step issues these together with
// the other per-game loads, so that their latencies overlap instead of forming a chain through the expansion).
struct AllocPre {
    uint32_t base;      // n_alloc[g]
    uint32_t meta;      // word 5 of the leaf's node record
    float row[3];       // network probabilities of actions lane, lane + 32, lane + 64 (0 beyond the action size)
};

// probs[idx] out of a row held lane-wise; every lane of the warp must call it.
__device__ __forceinline__ float az_row_gather(const float (&row)[3], int idx) {
    const float v0 = __shfl_sync(AZ_FULL, row[0], idx & 31), v1 = __shfl_sync(AZ_FULL, row[1], idx & 31),
                v2 = __shfl_sync(AZ_FULL, row[2], idx & 31);
    const int s = idx >> 5;
    return s == 0 ? v0 : (s == 1 ? v1 : v2);
}

template <int GAME>
__device__ __forceinline__ uint32_t az_alloc_children(const az_tree& t, const az_dims& d, az_node* nodes, int g,
                                                      int lane, uint32_t leaf, uint64_t own, uint64_t opp,
                                                      bool with_priors, uint32_t extra_flags, int* k_out,
                                                      const AllocPre* pre = nullptr) {
    using R = az_rules<GAME>;
    const uint64_t legal = az_warp_rules<GAME>::legal(own, opp, d, lane, nullptr);
    const int k = R::n_actions(legal);
    *k_out = k;
    const uint32_t base = pre ? pre->base : t.n_alloc[g];
    if (base + (uint32_t)k > (uint32_t)t.capacity) {
        if (lane == 0) t.status[g] |= AZ_STATUS_OVERFLOW;
        return 0u;
    }
    // up to 33 children: lane j owns child j and (rarely) child j + 32
    int a0 = AZ_PASS, a1 = AZ_PASS;
    float p0 = 0.f, p1 = 0.f;
    const bool has0 = lane < k, has1 = lane + 32 < k;
    if (has0) a0 = R::nth_action(legal, lane);
    if (has1) a1 = R::nth_action(legal, lane + 32);
    if (with_priors) {
        if (pre) {
            const float q0 = az_row_gather(pre->row, has0 ? R::net_index(a0, d) : 0);
            if (has0) p0 = q0;
            if (k > 32) {                                       // warp-uniform
                const float q1 = az_row_gather(pre->row, has1 ? R::net_index(a1, d) : 0);
                if (has1) p1 = q1;
            }
        } else {
            const float* pr = t.probs + (size_t)g * R::action_size(d);
            if (has0) p0 = pr[R::net_index(a0, d)];
            if (has1) p1 = pr[R::net_index(a1, d)];
        }
        float sum = az_warp_sum(p0 + p1);
        if (sum < 1e-6f) {
            p0 = p1 = (float)(1.0 / (double)k);
        } else {
            // only lanes that hold a child divide: a zero numerator (idle lanes; every lane's second slot when k <= 32)
            // would send the whole warp through the division's slow path, twice per simulation
            if (has0) p0 = __fdiv_rn(p0, sum);
            if (has1) p1 = __fdiv_rn(p1, sum);
        }
    }
    if (has0) az_store_node(&nodes[base + lane], 0.0, p0, 0u, 0u, az_make_meta(0, a0, 0u));
    if (has1) az_store_node(&nodes[base + lane + 32], 0.0, p1, 0u, 0u, az_make_meta(0, a1, 0u));
    __syncwarp();  // every lane has read n_alloc[g] before lane 0 bumps it
    if (lane == 0) {
        // leaf record: keep Q/P/N, set the child block and flags (second 16 bytes only)
        uint32_t* w = reinterpret_cast<uint32_t*>(&nodes[leaf]);
        uint32_t meta = pre ? pre->meta : w[5];
        uint32_t flags = ((meta >> 16) & 0xffu) | AZ_NODE_EVAL | extra_flags;
        w[4] = base;
        w[5] = az_make_meta(k, (int)((meta >> 8) & 0xffu), flags);
        t.n_alloc[g] = base + (uint32_t)k;
    }
    __syncwarp();
    return base;
}

// MCT.back_propagate (mcts.py:197-223): reward for the leaf, sign alternating towards the root; every node on the
// path gets Q = (N*Q + r)/(N+1) in float64 and N += 1.  Lane j updates path node j (nodes are distinct).
// `path_lane` = path[lane] and `rec_lane` = the first 16 bytes (Q, P, N) of that node, loaded by the caller ahead of
// time (valid for lane < min(len, 32)): nothing between that load and here writes those bytes -- the expansion only
// touches the second half of the leaf's record and records that are not on the path.
__device__ __forceinline__ void az_backup(const az_tree& t, az_node* nodes, int g, int lane, int player_to_play,
                                          double outcome, int len, uint32_t path_lane, uint4 rec_lane) {
    double reward;
    if (fabs(outcome) < 1e-4) reward = 0.0;
    else reward = ((double)player_to_play * outcome > 0.0) ? -fabs(outcome) : fabs(outcome);
    const uint32_t* path = t.path + (size_t)g * t.max_depth;
#pragma unroll 1
    for (int j = lane; j < len; j += 32) {
        az_node* nd = &nodes[j < 32 ? path_lane : path[j]];
        uint4* q = reinterpret_cast<uint4*>(nd);
        uint4 a = j < 32 ? rec_lane : q[0];
        double Q = __hiloint2double((int)a.y, (int)a.x);
        uint32_t N = a.w;
        double r = ((len - 1 - j) & 1) ? -reward : reward;
        double Qn = __ddiv_rn(__dadd_rn(__dmul_rn((double)N, Q), r), (double)(N + 1u));
        a.x = (uint32_t)__double2loint(Qn);
        a.y = (uint32_t)__double2hiint(Qn);
        a.w = N + 1u;
        q[0] = a;
    }
    __syncwarp();
}

// Per-game scalars of the pending leaf, loaded together at kernel entry: the loads are independent, so their
// latencies overlap instead of forming a chain through az_finish.
struct LeafRegs {
    int kind, leaf_player;
    uint32_t leaf, path_len;
    uint64_t own, opp;
    float value;
    uint32_t path_lane;     // path[lane]
    AllocPre alloc;         // n_alloc, network row (first round of loads); leaf meta word (second round)
    uint4 rec_lane;         // first half of node path[lane] (second round)
};

// First round: everything that only depends on g.  With az_tree.logits the evaluator's raw bf16 row is turned into
// probabilities / value here (net_heads.cuh: the arithmetic of az_net_heads).
template <int GAME, int MODE>
__device__ __forceinline__ LeafRegs az_load_leaf(const az_tree& t, const az_dims& d, int g, int lane) {
    LeafRegs r;
    r.kind = t.leaf_kind[g];
    r.leaf = t.leaf_node[g];
    r.leaf_player = t.leaf_player[g];
    r.path_len = t.path_len[g];
    r.own = t.leaf_own[g];
    r.opp = t.leaf_opp[g];
    r.alloc.base = t.n_alloc[g];
    r.path_lane = lane < t.max_depth ? t.path[(size_t)g * t.max_depth + lane] : 0u;
    r.alloc.row[0] = r.alloc.row[1] = r.alloc.row[2] = 0.f;
    if (MODE == AZ_MODE_NEURAL) {
        const int A = az_rules<GAME>::action_size(d);
        if (t.logits) {
            const __nv_bfloat16* row = reinterpret_cast<const __nv_bfloat16*>(t.logits) + (size_t)g * t.logits_ld;
            float x[3];
#pragma unroll
            for (int i = 0; i < 3; ++i) x[i] = (lane + 32 * i) < A ? __bfloat162float(row[lane + 32 * i]) : 0.f;
            const __nv_bfloat16 v = row[A];
            az_heads_softmax(x, lane, A, r.alloc.row);
            r.value = az_heads_value(v);
        } else {
            const float* pr = t.probs + (size_t)g * A;
#pragma unroll
            for (int i = 0; i < 3; ++i) if (lane + 32 * i < A) r.alloc.row[i] = pr[lane + 32 * i];
            r.value = t.value[g];
        }
    } else {
        r.value = t.value[g];
    }
    return r;
}

// Second round: the two node reads whose addresses the first round produced.
__device__ __forceinline__ void az_load_leaf_nodes(LeafRegs& r, const az_node* nodes, int lane) {
    r.alloc.meta = 0u;
    r.rec_lane = make_uint4(0u, 0u, 0u, 0u);
    if (r.kind == AZ_LEAF_IDLE) return;
    r.alloc.meta = reinterpret_cast<const uint32_t*>(&nodes[r.leaf])[5];
    if (r.kind != AZ_LEAF_PRIME && lane < (int)r.path_len) r.rec_lane = reinterpret_cast<const uint4*>(&nodes[r.path_lane])[0];
}

template <int GAME>
__device__ __forceinline__ void az_finish(const az_tree& t, const az_dims& d, az_node* nodes, int g, int lane,
                                          const LeafRegs& lf) {
    const int kind = lf.kind;
    if (kind == AZ_LEAF_IDLE) return;
    const uint32_t leaf = lf.leaf;
    const int leaf_player = lf.leaf_player;
    if (kind == AZ_LEAF_NN || kind == AZ_LEAF_PRIME) {
        int k;
        az_alloc_children<GAME>(t, d, nodes, g, lane, leaf, lf.own, lf.opp, true, 0u, &k, &lf.alloc);
    }
    if (kind != AZ_LEAF_PRIME) {
        double outcome;
        if (kind == AZ_LEAF_NN) outcome = (double)((float)leaf_player * lf.value);     // base.py:366
        else if (kind == AZ_LEAF_ROLLOUT) outcome = (double)lf.value;                   // absolute winner id
        else outcome = kind == AZ_LEAF_DRAW ? 0.0 : (kind == AZ_LEAF_BLACK ? 1.0 : -1.0);
        az_backup(t, nodes, g, lane, leaf_player, outcome, (int)lf.path_len, lf.path_lane, lf.rec_lane);
        if (lane == 0) t.sims_done[g] += 1u;
    }
    if (lane == 0) t.leaf_kind[g] = AZ_LEAF_IDLE;
    __syncwarp();
}

// ------------------------------------------------------------------------------------------------ select

// fair_max's uniform pick among tied lanes (utils.py:28-34).  Out of line: ties are rare (a few unvisited children with
// equal priors), and the Philox rounds would otherwise sit in the middle of the selection loop's instruction stream.
__device__ __noinline__ int az_tie_pick(uint64_t seed, uint32_t g, uint32_t event, uint32_t ties) {
    az_u32x4 r = az_draw4(seed, g, event, 0u, 0x54494555u);
    int pick = (int)az_bounded(r.x, (uint32_t)__popc(ties));
    uint32_t m = ties;
    for (int i = 0; i < pick; ++i) m &= m - 1;
    return __ffs(m) - 1;
}

// Warp arg-max of `score` over the lanes with valid == true.  Ties: lowest child index (parity mode) or uniform
// among the tied children (fair_max, utils.py:28-34).  `idx` is the child index held by the lane.
// Returns the winning child index (same value in every lane) and, through src_lane, the lane that holds it.
//
// The maximum is taken with two `redux.sync` on an order-preserving 64-bit integer image of the double (high word, then
// low word among the lanes that hold the high-word maximum) instead of five dependent 64-bit shuffle steps: the
// selection loop is a chain of dependent warp collectives, so their count is what a simulation costs.  A lane's `idx`
// is `lane` or `lane + 32`, so the lowest tied child index is the lowest tied lane with idx < 32, else the lowest tied lane.
__device__ __forceinline__ int az_warp_argmax(double score, int idx, bool valid, bool tie_random, uint64_t seed,
                                              uint32_t g, uint32_t event, int* src_lane) {
    // NaN never occurs: Q, P finite; +inf only for unvisited children in UCT mode.  -0.0 is folded into +0.0 so that
    // equal doubles have equal images (score == best in the reference's comparison).
    uint64_t b = (uint64_t)__double_as_longlong(score);
    if (b == 0x8000000000000000ULL) b = 0ULL;
    b ^= (b >> 63) ? ~0ULL : 0x8000000000000000ULL;          // unsigned order == double order
    if (!valid) b = 0ULL;                                      // below the image of every double, -inf included
    const uint32_t hi = (uint32_t)(b >> 32), lo = (uint32_t)b;
    const uint32_t max_hi = __reduce_max_sync(AZ_FULL, hi);
    const bool cand = hi == max_hi;
    const uint32_t max_lo = __reduce_max_sync(AZ_FULL, cand ? lo : 0u);
    const bool tied = valid && cand && lo == max_lo;
    const uint32_t ties = __ballot_sync(AZ_FULL, tied);
    const uint32_t first_half = __ballot_sync(AZ_FULL, tied && idx < 32);
    int win_lane;
    if (!tie_random || __popc(ties) == 1) {
        win_lane = __ffs(first_half ? first_half : ties) - 1;
    } else {
        win_lane = az_tie_pick(seed, g, event, ties);
    }
    *src_lane = win_lane;
    // the winner's idx is its lane, or lane + 32 when it is not among the first-half ties: no shuffle needed
    return win_lane + (((first_half >> win_lane) & 1u) ? 0 : 32);
}

// Dirichlet noise on the priors of the root's children (mcts.py:234-240): P = (1-eps) P + eps * Dir(alpha)_j with
// Gamma(alpha) draws by Marsaglia-Tsang (alpha < 1 boosted by U^(1/alpha)), lane j <-> child j.
__device__ __noinline__ double az_gamma_draw(double alpha, uint64_t seed, uint32_t g, uint32_t event, uint32_t j) {
    uint32_t attempt = 0;
    az_u32x4 r = az_draw4(seed, g, event, j, 0x47414D30u);
    double boost = 1.0;
    if (alpha < 1.0) {
        boost = pow(az_uniform53(r.x, r.y), 1.0 / alpha);
        alpha += 1.0;
    }
    const double d = alpha - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
    while (true) {
        ++attempt;
        az_u32x4 a = az_draw4(seed, g, event, j, 0x47414D30u + attempt);
        az_u32x4 b = az_draw4(seed, g, event, j, 0x47424D30u + attempt);
        double u1 = az_uniform53(a.x, a.y), u2 = az_uniform53(a.z, a.w), u = az_uniform53(b.x, b.y);
        double x = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
        double v = 1.0 + c * x;
        if (v <= 0.0) continue;
        v = v * v * v;
        if (log(u) < 0.5 * x * x + d - d * v + d * log(v)) return boost * d * v;
    }
}

// Out of line (once per move per game; its float64 pow / log / cospi would triple the selection loop's code), with
// scalar arguments only so that the kernel's parameter struct never has to be materialised in local memory.
__device__ __noinline__ void az_root_noise(az_node* nodes, uint32_t first_child, int k, float alpha, float eps_f,
                                           uint64_t seed, int g, int lane, uint32_t event) {
    double g0 = 0.0, g1 = 0.0;
    if (lane < k) g0 = az_gamma_draw((double)alpha, seed, (uint32_t)g, event, (uint32_t)lane);
    if (lane + 32 < k) g1 = az_gamma_draw((double)alpha, seed, (uint32_t)g, event, (uint32_t)lane + 32u);
    const double sum = az_warp_sum(g0 + g1);
    const double eps = (double)eps_f;
    if (sum > 0.0) {
        for (int j = lane; j < k; j += 32) {
            float* P = &nodes[first_child + j].P;
            double noise = (j < 32 ? g0 : g1) / sum;
            *P = (float)((1.0 - eps) * (double)(*P) + eps * noise);
        }
    }
    if (lane == 0) {
        uint32_t* w = reinterpret_cast<uint32_t*>(&nodes[0]);
        w[5] |= (AZ_NODE_NOISED << 16);
    }
    __syncwarp();
}

// Picks the child of `cur` that maximises PUCT (neural, mcts.py:44-46) or UCT (rollout, mcts.py:38-42).
// Returns the child's pool index and its record (broadcast to all lanes).
// `pre` = record of child `lane` (valid when lane < k), loaded by the caller ahead of time.
template <int MODE>
__device__ __forceinline__ uint32_t az_pick_child(const az_tree& t, const az_node* nodes, int g, int lane,
                                                  const NodeRec& cur, uint32_t event, NodeRec* child_out,
                                                  const NodeRec& pre) {
    const int k = cur.n_children();
    const double parentN = (double)cur.N;
    // sqrt / log of the parent's visit count are shared by all children
    const double sqrtN = __dsqrt_rn(parentN);
    const double logN = (MODE == AZ_MODE_ROLLOUT && cur.N > 0u) ? log(parentN) : 0.0;
    NodeRec mine;
    mine.Q = 0.0; mine.P = 0.f; mine.N = 0u; mine.first_child = 0u; mine.meta = 0u;
    double best_score = -INFINITY;
    int best_j = -1;
#pragma unroll 1
    for (int j = lane; j < k; j += 32) {
        const NodeRec c = j < 32 ? pre : az_load_node(&nodes[cur.first_child + j]);
        double s;
        if (MODE == AZ_MODE_NEURAL) {
            // Q + c_puct * P * sqrt(N_parent) / (1 + N), c_puct = 1.0, evaluated left to right in float64
            s = __dadd_rn(c.Q, __ddiv_rn(__dmul_rn((double)c.P, sqrtN), (double)(1u + c.N)));
        } else {
            s = c.N == 0u ? INFINITY
                          : __dadd_rn(c.Q, __dmul_rn(1.4142135623730951, __dsqrt_rn(__ddiv_rn(logN, (double)c.N))));
        }
        if (best_j < 0 || s > best_score) { best_score = s; best_j = j; mine = c; }
    }
    int src;
    const int win = az_warp_argmax(best_score, best_j, best_j >= 0, t.tie_random != 0, t.seed, (uint32_t)g, event, &src);
    *child_out = az_bcast_node(mine, src);
    return cur.first_child + (uint32_t)win;
}

template <int GAME>
__device__ __forceinline__ int az_terminal_kind(uint64_t own, uint64_t opp, int player, const az_dims& d) {
    int w = az_rules<GAME>::winner(own, opp, player, d);
    return w == 0 ? AZ_LEAF_DRAW : (w > 0 ? AZ_LEAF_BLACK : AZ_LEAF_WHITE);
}

// MCT.__search_iter up to the evaluation (mcts.py:226-248): priming, noise, select_node; writes the leaf buffers.
struct RootRegs {       // per-game inputs of the selection, loaded with the first round of k_tree_step
    uint64_t own, opp;
    int player;
    uint32_t event;
};
__device__ __forceinline__ RootRegs az_load_root(const az_tree& t, int g) {
    RootRegs r;
    r.own = t.root_own[g]; r.opp = t.root_opp[g]; r.player = t.root_player[g]; r.event = t.rng_ctr[g];
    return r;
}

template <int GAME, int MODE>
__device__ __forceinline__ void az_select(const az_tree& t, const az_dims& d, az_node* nodes, int g, int lane,
                                          const RootRegs& root) {
    using W = az_warp_rules<GAME>;
    uint64_t own = root.own, opp = root.opp;
    int player = root.player;
    uint32_t event = root.event;
    uint32_t* path = t.path + (size_t)g * t.max_depth;
    NodeRec cur = az_load_node(&nodes[0]);
    uint32_t cur_idx = 0u;
    int kind = AZ_LEAF_IDLE;
    int len = 1;
    if (lane == 0) path[0] = 0u;

    if (MODE == AZ_MODE_NEURAL && !(cur.flags() & AZ_NODE_EVAL)) {
        kind = AZ_LEAF_PRIME;                                   // mcts.py:231-233
        len = 0;
    } else {
        if (MODE == AZ_MODE_NEURAL && t.dirichlet_alpha > 0.f && t.dirichlet_eps > 0.f &&
            (cur.flags() & AZ_NODE_EXPANDED) && !(cur.flags() & AZ_NODE_NOISED)) {
            az_root_noise(nodes, cur.first_child, cur.n_children(), t.dirichlet_alpha, t.dirichlet_eps, t.seed, g, lane,
                          event);                               // mcts.py:235-240
            event += 1u;
        }
        NodeRec pre;                 // child `lane` of cur, loaded one level ahead (while the move is applied)
        pre.Q = 0.0; pre.P = 0.f; pre.N = 0u; pre.first_child = 0u; pre.meta = 0u;
        bool pre_ok = false;
        while (true) {
            // invariant: cur is the root or a node with N > 0
            if (!(cur.flags() & AZ_NODE_EXPANDED)) {
                bool terminal;
                W::legal(own, opp, d, lane, &terminal);
                if (terminal) {                                 // mcts.py:146-147
                    kind = MODE == AZ_MODE_NEURAL ? az_terminal_kind<GAME>(own, opp, player, d) : AZ_LEAF_ROLLOUT;
                    break;
                }
                if (MODE == AZ_MODE_NEURAL) {
                    if (!(cur.flags() & AZ_NODE_EVAL)) { kind = AZ_LEAF_NN; break; }   // pool-overflow fallback
                    if (lane == 0) {
                        uint32_t* w = reinterpret_cast<uint32_t*>(&nodes[cur_idx]);
                        w[5] |= (AZ_NODE_EXPANDED << 16);       // children become part of the tree (mcts.py:155-160)
                    }
                } else {
                    int k;
                    uint32_t base = az_alloc_children<GAME>(t, d, nodes, g, lane, cur_idx, own, opp, false,
                                                            AZ_NODE_EXPANDED, &k);
                    if (base == 0u) { kind = AZ_LEAF_ROLLOUT; break; }                   // overflow: evaluate here
                    // mcts.py:163-165: uniform pick among the fresh children
                    az_u32x4 r = az_draw4(t.seed, (uint32_t)g, event, 0u, 0x45585044u);
                    event += 1u;
                    int j = (int)az_bounded(r.x, (uint32_t)k);
                    NodeRec c = az_load_node(&nodes[base + j]);
                    W::apply(own, opp, c.move(), d, lane);
                    player = -player;
                    cur_idx = base + (uint32_t)j;
                    if (len < t.max_depth) { if (lane == 0) path[len] = cur_idx; ++len; }
                    kind = AZ_LEAF_ROLLOUT;
                    break;
                }
            }
            if (!pre_ok && lane < cur.n_children()) pre = az_load_node(&nodes[cur.first_child + lane]);
            NodeRec child;
            uint32_t child_idx = az_pick_child<MODE>(t, nodes, g, lane, cur, event, &child, pre);
            event += 1u;
            // the chosen child's own children block (if it has one) is fetched now, under the move application
            pre_ok = child.first_child != 0u;
            if (pre_ok && lane < child.n_children()) pre = az_load_node(&nodes[child.first_child + lane]);
            W::apply(own, opp, child.move(), d, lane);
            player = -player;
            if (len >= t.max_depth) {
                if (lane == 0) t.status[g] |= AZ_STATUS_DEPTH;
                kind = MODE == AZ_MODE_NEURAL ? AZ_LEAF_NN : AZ_LEAF_ROLLOUT;
                break;
            }
            if (lane == 0) path[len] = child_idx;
            ++len;
            cur = child;
            cur_idx = child_idx;
            if (cur.N == 0u) {                                  // mcts.py:143-144: first visit = the leaf
                if (MODE == AZ_MODE_NEURAL) {
                    bool terminal;
                    W::legal(own, opp, d, lane, &terminal);
                    kind = terminal ? az_terminal_kind<GAME>(own, opp, player, d) : AZ_LEAF_NN;
                } else
                    kind = AZ_LEAF_ROLLOUT;
                break;
            }
        }
    }
    if (lane == 0) {
        t.leaf_node[g] = cur_idx;
        t.leaf_kind[g] = (uint8_t)kind;
        t.leaf_own[g] = own;
        t.leaf_opp[g] = opp;
        t.leaf_player[g] = (int8_t)player;
        t.path_len[g] = (uint32_t)len;
        t.rng_ctr[g] = event;
    }
    if (MODE == AZ_MODE_NEURAL && t.enc) az_encode_leaf(t.enc, t.enc_dtype, d.rows, d.cols, g, lane, own, opp);
}

#ifndef AZ_TREE_MIN_CTAS
#define AZ_TREE_MIN_CTAS 7    // resident CTAs per SM the register budget is sized for: 1024 CTAs / 148 SMs = one wave at 7
#endif
// MODE (AZ_MODE_NEURAL / AZ_MODE_ROLLOUT) is a template parameter so that each instantiation only carries its own mode's
// code: the kernel is bound by instruction fetch and dependent-issue latency, not by memory (ncu: `no_instructions` is the
// top warp stall), so everything that is not on the per-simulation path is compiled out or kept out of line.
// The block size is the launcher's choice (az_tree_step: any number of warps up to AZ_TREE_MIN_CTAS * AZ_WARPS_PER_CTA
// -- one warp per game, no block-level cooperation); the bound below gives the same 72-register budget as (128, 7).
template <int GAME, int MODE>
__global__ void __launch_bounds__(AZ_WARPS_PER_CTA * 32 * AZ_TREE_MIN_CTAS, 1) k_tree_step(az_tree t, int n_sim, int do_finish, int do_select) {
    // launched with az_launch_pdl: every warp waits for the evaluator's outputs (probs / value) before anything else --
    // also the warps of finished games, so that this grid never completes before its predecessor has -- and then lets
    // the next kernel (the network trunk of the following step) start its prologue beside the tail of this one
    az_pdl_wait();
    az_pdl_trigger();
    const int g = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= t.G) return;
    const az_dims d = az_rules<GAME>::dims(t.n, t.height);
    // first round: every load whose address only depends on g, issued before the first use of any of them
    const uint8_t active = t.active[g];
    const uint8_t half = t.half[g];
    const uint32_t sims_before = t.sims_done[g];
    LeafRegs lf;
    if (do_finish) lf = az_load_leaf<GAME, MODE>(t, d, g, lane);
    RootRegs root;
    if (do_select) root = az_load_root(t, g);
    if (!active) return;
    az_node* nodes = t.nodes + ((size_t)g * 2 + half) * (size_t)t.capacity;
    bool backed_up = false;
    if (do_finish) {
        az_load_leaf_nodes(lf, nodes, lane);       // second round
        az_finish<GAME>(t, d, nodes, g, lane, lf);
        backed_up = lf.kind != AZ_LEAF_IDLE && lf.kind != AZ_LEAF_PRIME;
    }
    if (do_select && sims_before + (backed_up ? 1u : 0u) < (uint32_t)n_sim) az_select<GAME, MODE>(t, d, nodes, g, lane, root);
}

// ------------------------------------------------------------------------------------------------ throughput mode
//
// L = az_tree.n_leaves > 1 leaves in flight per game and step (an EXTENSION: the reference evaluates one leaf at a time,
// mcts.py:254-269; with L = 1 this kernel is the reference's order and bit-identical to k_tree_step).  One warp still
// owns a game and makes its L descents one after the other -- descent l must see the virtual losses of descents < l,
// so they are a dependent chain anyway -- which keeps the search deterministic; what the mode buys is a leaf batch of
// G*L positions for the network.  Per node two more words (the spare 8 bytes of az_node):
//   vloss  descents in flight through the node: added along the path when a descent ends, removed by its backup.  A child
//          with v > 0 is scored as if it had v more visits that all lost: Q' = (W - v)/(N + v), N' = N + v; its parent's
//          visit count in the exploration term likewise.  A child with v == 0 is scored with the reference's float64
//          running mean exactly as in parity mode;
//   W      float32 sum of the rewards backed up through the node (the Q of the virtual-loss formula above).
// vloss, N and W are updated with atomics (red.global): today one warp owns a game, so they are uncontended, but the
// counters stay correct if the descents of a game are ever spread over warps.
// Leaf slot of (game g, descent l) in every per-leaf array: l*G + g, so the evaluator sees one flat batch of G*L rows.
// Two descents of one step can meet at the same not-yet-evaluated node (forced lines: a single child): the later one is
// dropped (no virtual loss was added yet, nothing to undo) and the game simply backs up fewer leaves that step; the host
// runs steps until every game has its n_sim simulations.

__device__ __forceinline__ void az_backup_multi(const az_tree& t, az_node* nodes, int lane, int player_to_play,
                                                double outcome, int len, const uint32_t* path) {
    double reward;
    if (fabs(outcome) < 1e-4) reward = 0.0;
    else reward = ((double)player_to_play * outcome > 0.0) ? -fabs(outcome) : fabs(outcome);
#pragma unroll 1
    for (int j = lane; j < len; j += 32) {
        az_node* nd = &nodes[path[j]];
        const uint4 a = reinterpret_cast<const uint4*>(nd)[0];
        const double Q = __hiloint2double((int)a.y, (int)a.x);
        const uint32_t N = a.w;
        const double r = ((len - 1 - j) & 1) ? -reward : reward;
        nd->Q = __ddiv_rn(__dadd_rn(__dmul_rn((double)N, Q), r), (double)(N + 1u));      // mcts.py:220
        atomicAdd(&nd->N, 1u);
        atomicAdd(&nd->W, (float)r);
        atomicSub(&nd->vloss, 1u);
    }
    __syncwarp();
}

// One descent (mcts.py:226-248 up to the evaluation) under the virtual losses of the descents already in flight.
// Returns 1 when a leaf was written to slot `s`, 0 when the descent was dropped; *only_one is set when no further descent
// may start this step (the root itself is being evaluated).
template <int GAME>
__device__ __forceinline__ int az_select_multi(const az_tree& t, const az_dims& d, az_node* nodes, int g, int lane, int s,
                                               const RootRegs& root, uint32_t& event, bool* only_one) {
    using W = az_warp_rules<GAME>;
    uint64_t own = root.own, opp = root.opp;
    int player = root.player;
    uint32_t* path = t.path + (size_t)s * t.max_depth;
    NodeRec cur = az_load_node(&nodes[0]);
    uint32_t cur_idx = 0u;
    int kind = AZ_LEAF_IDLE;
    int len = 1;
    if (lane == 0) path[0] = 0u;
    *only_one = false;
    if (!(cur.flags() & AZ_NODE_EVAL)) {
        kind = AZ_LEAF_PRIME;                                   // mcts.py:231-233
        len = 0;
        *only_one = true;
    } else {
        if (t.dirichlet_alpha > 0.f && t.dirichlet_eps > 0.f && (cur.flags() & AZ_NODE_EXPANDED) &&
            !(cur.flags() & AZ_NODE_NOISED)) {
            az_root_noise(nodes, cur.first_child, cur.n_children(), t.dirichlet_alpha, t.dirichlet_eps, t.seed, g, lane,
                          event);                               // mcts.py:235-240
            event += 1u;
        }
        while (true) {
            if (!(cur.flags() & AZ_NODE_EXPANDED)) {
                bool terminal;
                W::legal(own, opp, d, lane, &terminal);
                if (terminal) { kind = az_terminal_kind<GAME>(own, opp, player, d); break; }     // mcts.py:146-147
                if (!(cur.flags() & AZ_NODE_EVAL)) { kind = cur.vloss ? AZ_LEAF_IDLE : AZ_LEAF_NN; break; }   // overflow fallback
                if (lane == 0) {
                    uint32_t* w = reinterpret_cast<uint32_t*>(&nodes[cur_idx]);
                    w[5] |= (AZ_NODE_EXPANDED << 16);           // mcts.py:155-160
                }
            }
            // PUCT over the children, lane j <-> child j (and j + 32)
            const int k = cur.n_children();
            const double sqrtN = __dsqrt_rn((double)(cur.N + cur.vloss));
            NodeRec mine;
            mine.Q = 0.0; mine.P = 0.f; mine.N = 0u; mine.first_child = 0u; mine.meta = 0u; mine.vloss = 0u; mine.W = 0.f;
            double best_score = -INFINITY;
            int best_j = -1;
#pragma unroll 1
            for (int j = lane; j < k; j += 32) {
                const NodeRec c = az_load_node(&nodes[cur.first_child + j]);
                double sc;
                if (c.vloss == 0u) {
                    sc = __dadd_rn(c.Q, __ddiv_rn(__dmul_rn((double)c.P, sqrtN), (double)(1u + c.N)));     // mcts.py:44-46
                } else {
                    const double n_eff = (double)(c.N + c.vloss);
                    const double q_eff = __ddiv_rn((double)c.W - (double)c.vloss, n_eff);
                    sc = __dadd_rn(q_eff, __ddiv_rn(__dmul_rn((double)c.P, sqrtN), 1.0 + n_eff));
                }
                if (best_j < 0 || sc > best_score) { best_score = sc; best_j = j; mine = c; }
            }
            int src;
            const int win = az_warp_argmax(best_score, best_j, best_j >= 0, t.tie_random != 0, t.seed, (uint32_t)g, event, &src);
            event += 1u;
            NodeRec child = az_bcast_node(mine, src);
            child.vloss = __shfl_sync(AZ_FULL, mine.vloss, src);
            const uint32_t child_idx = cur.first_child + (uint32_t)win;
            W::apply(own, opp, child.move(), d, lane);
            player = -player;
            if (len >= t.max_depth) {
                if (lane == 0) t.status[g] |= AZ_STATUS_DEPTH;
                kind = AZ_LEAF_NN;
                break;
            }
            if (lane == 0) path[len] = child_idx;
            ++len;
            cur = child;
            cur_idx = child_idx;
            if (cur.N == 0u) {                                  // mcts.py:143-144: first visit = the leaf
                bool terminal;
                W::legal(own, opp, d, lane, &terminal);
                if (terminal) kind = az_terminal_kind<GAME>(own, opp, player, d);
                else kind = cur.vloss ? AZ_LEAF_IDLE : AZ_LEAF_NN;     // already being evaluated this step: drop
                break;
            }
        }
    }
    if (kind == AZ_LEAF_IDLE) return 0;
    __syncwarp();                                               // path[] written by lane 0 is read by all lanes
    if (kind != AZ_LEAF_PRIME)
        for (int j = lane; j < len; j += 32) atomicAdd(&nodes[path[j]].vloss, 1u);
    if (lane == 0) {
        t.leaf_node[s] = cur_idx;
        t.leaf_kind[s] = (uint8_t)kind;
        t.leaf_own[s] = own;
        t.leaf_opp[s] = opp;
        t.leaf_player[s] = (int8_t)player;
        t.path_len[s] = (uint32_t)len;
    }
    if (t.enc) az_encode_leaf(t.enc, t.enc_dtype, d.rows, d.cols, s, lane, own, opp);
    __syncwarp();
    return 1;
}

template <int GAME>
__global__ void __launch_bounds__(AZ_WARPS_PER_CTA * 32, AZ_TREE_MIN_CTAS) k_tree_step_multi(az_tree t, int n_sim, int do_finish, int do_select) {
    az_pdl_wait();
    az_pdl_trigger();
    const int g = blockIdx.x * AZ_WARPS_PER_CTA + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= t.G || !t.active[g]) return;
    const az_dims d = az_rules<GAME>::dims(t.n, t.height);
    const int L = t.n_leaves, A = az_rules<GAME>::action_size(d);
    az_node* nodes = t.nodes + ((size_t)g * 2 + t.half[g]) * (size_t)t.capacity;
    uint32_t sims = t.sims_done[g];
    if (do_finish) {
#pragma unroll 1
        for (int l = 0; l < L; ++l) {
            const int s = l * t.G + g;
            const int kind = t.leaf_kind[s];
            if (kind == AZ_LEAF_IDLE) continue;
            const uint32_t leaf = t.leaf_node[s];
            const int leaf_player = t.leaf_player[s];
            const int len = (int)t.path_len[s];
            float value = 0.f;
            if (kind == AZ_LEAF_NN || kind == AZ_LEAF_PRIME) {
                AllocPre pre;
                pre.base = t.n_alloc[g];
                pre.meta = reinterpret_cast<const uint32_t*>(&nodes[leaf])[5];
                pre.row[0] = pre.row[1] = pre.row[2] = 0.f;
                if (t.logits) {
                    const __nv_bfloat16* row = reinterpret_cast<const __nv_bfloat16*>(t.logits) + (size_t)s * t.logits_ld;
                    float x[3];
#pragma unroll
                    for (int i = 0; i < 3; ++i) x[i] = (lane + 32 * i) < A ? __bfloat162float(row[lane + 32 * i]) : 0.f;
                    const __nv_bfloat16 v = row[A];
                    az_heads_softmax(x, lane, A, pre.row);
                    value = az_heads_value(v);
                } else {
                    const float* pr = t.probs + (size_t)s * A;
#pragma unroll
                    for (int i = 0; i < 3; ++i) if (lane + 32 * i < A) pre.row[i] = pr[lane + 32 * i];
                    value = t.value[s];
                }
                int k;
                az_alloc_children<GAME>(t, d, nodes, g, lane, leaf, t.leaf_own[s], t.leaf_opp[s], true, 0u, &k, &pre);
            }
            if (kind != AZ_LEAF_PRIME) {
                double outcome;
                if (kind == AZ_LEAF_NN) outcome = (double)((float)leaf_player * value);        // base.py:366
                else outcome = kind == AZ_LEAF_DRAW ? 0.0 : (kind == AZ_LEAF_BLACK ? 1.0 : -1.0);
                az_backup_multi(t, nodes, lane, leaf_player, outcome, len, t.path + (size_t)s * t.max_depth);
                sims += 1u;
            }
            if (lane == 0) t.leaf_kind[s] = AZ_LEAF_IDLE;
            __syncwarp();
        }
        if (lane == 0) t.sims_done[g] = sims;
    }
    if (do_select) {
        const RootRegs root = az_load_root(t, g);
        uint32_t event = root.event;
        uint32_t in_flight = 0u;
#pragma unroll 1
        for (int l = 0; l < L; ++l) {
            if (sims + in_flight >= (uint32_t)n_sim) break;
            bool only_one;
            in_flight += (uint32_t)az_select_multi<GAME>(t, d, nodes, g, lane, l * t.G + g, root, event, &only_one);
            if (only_one) break;
        }
        if (lane == 0) t.rng_ctr[g] = event;
    }
}

// ------------------------------------------------------------------------------------------------ misc kernels

__global__ void __launch_bounds__(256) k_tree_reset(az_tree t, const uint8_t* __restrict__ mask) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= t.G || (mask && !mask[g])) return;
    az_node* nodes = t.nodes + ((size_t)g * 2 + t.half[g]) * (size_t)t.capacity;
    az_store_node(&nodes[0], 0.0, 0.f, 0u, 0u, az_make_meta(0, 0xff, 0u));
    t.n_alloc[g] = 1u;
    for (int l = 0; l < max(1, t.n_leaves); ++l) t.leaf_kind[l * t.G + g] = AZ_LEAF_IDLE;
    t.sims_done[g] = 0u;
    t.status[g] = 0u;
    t.path_len[g] = 0u;
}

// One thread per game: a uniform random playout from the leaf board (MCT.rollout, mcts.py:173-180).
template <int GAME>
__global__ void __launch_bounds__(128) k_tree_rollout(az_tree t) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= t.G || !t.active[g] || t.leaf_kind[g] != AZ_LEAF_ROLLOUT) return;
    const az_dims d = az_rules<GAME>::dims(t.n, t.height);
    uint64_t own = t.leaf_own[g], opp = t.leaf_opp[g];
    int player = t.leaf_player[g];
    const uint32_t ev = t.rng_ctr[g];
    const uint64_t id = ((uint64_t)(uint32_t)g << 32) | (uint64_t)ev, seed = t.seed ^ 0x524F4C4C4F555421ULL;
    if (GAME == AZ_GAME_OTHELLO) az_random_playout(own, opp, player, d.bmask, id, seed);
    else az_random_playout_g<GAME>(own, opp, player, d, id, seed);
    t.rng_ctr[g] = ev + 1u;
    t.value[g] = (float)az_rules<GAME>::winner(own, opp, player, d);
}

__global__ void __launch_bounds__(AZ_WARPS_PER_CTA * 32) k_tree_root_stats(az_tree t, int32_t* __restrict__ N,
                                                                         double* __restrict__ Q,
                                                                         float* __restrict__ P,
                                                                         uint32_t* __restrict__ root_N) {
    const int g = blockIdx.x * AZ_WARPS_PER_CTA + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= t.G) return;
    const az_node* nodes = t.nodes + ((size_t)g * 2 + t.half[g]) * (size_t)t.capacity;
    for (int a = lane; a < AZ_CANVAS_ACTIONS; a += 32) {
        if (N) N[(size_t)g * AZ_CANVAS_ACTIONS + a] = -1;
        if (Q) Q[(size_t)g * AZ_CANVAS_ACTIONS + a] = 0.0;
        if (P) P[(size_t)g * AZ_CANVAS_ACTIONS + a] = 0.f;
    }
    __syncwarp();
    NodeRec root = az_load_node(&nodes[0]);
    if (lane == 0 && root_N) root_N[g] = root.N;
    if (!(root.flags() & AZ_NODE_EXPANDED)) return;
    for (int j = lane; j < root.n_children(); j += 32) {
        NodeRec c = az_load_node(&nodes[root.first_child + j]);
        size_t o = (size_t)g * AZ_CANVAS_ACTIONS + c.move();
        if (N) N[o] = (int32_t)c.N;
        if (Q) Q[o] = c.Q;
        if (P) P[o] = c.P;
    }
}

// get_action_probs (mcts.py:95-116) + the categorical draw of MCTSPlayer.get_move (players.py:184-189) for game g by
// its warp: returns the chosen move (same value in every lane) and writes the policy row pi_row[0 .. A) (may be NULL).
template <int GAME>
__device__ __forceinline__ int az_choose_warp(const az_tree& t, const az_dims& d, int g, int lane, float T,
                                              float* __restrict__ pi_row) {
    using R = az_rules<GAME>;
    const int A = R::action_size(d);
    const az_node* nodes = t.nodes + ((size_t)g * 2 + t.half[g]) * (size_t)t.capacity;
    if (pi_row) for (int a = lane; a < A; a += 32) pi_row[a] = 0.f;
    __syncwarp();
    NodeRec root = az_load_node(&nodes[0]);
    const uint32_t event = t.rng_ctr[g];
    if (!(root.flags() & AZ_NODE_EXPANDED)) {                   // mcts.py:101-102: {pass_move: 1.}
        if (lane == 0 && pi_row) pi_row[A - 1] = 1.f;
        return AZ_PASS;
    }
    const int k = root.n_children();
    // lane j holds child j (and j+32 in the 33-children corner case)
    uint32_t n0 = 0, n1 = 0;
    int m0 = AZ_PASS, m1 = AZ_PASS;
    if (lane < k) { NodeRec c = az_load_node(&nodes[root.first_child + lane]); n0 = c.N; m0 = c.move(); }
    if (lane + 32 < k) { NodeRec c = az_load_node(&nodes[root.first_child + lane + 32]); n1 = c.N; m1 = c.move(); }
    int chosen_move;
    if (T == 0.f) {
        // fair_max over visit counts
        bool second = (lane + 32 < k) && n1 > n0;
        int src;
        az_warp_argmax((double)(second ? n1 : n0), second ? lane + 32 : lane, lane < k, t.tie_random != 0, t.seed,
                       (uint32_t)g, event, &src);
        chosen_move = __shfl_sync(AZ_FULL, second ? m1 : m0, src);
        if (lane == 0 && pi_row) pi_row[R::net_index(chosen_move, d)] = 1.f;
    } else {
        const double inv = 1.0 / (double)T;
        double w0 = lane < k ? pow((double)n0, inv) : 0.0;
        double w1 = lane + 32 < k ? pow((double)n1, inv) : 0.0;
        if (T == 1.f) { w0 = lane < k ? (double)n0 : 0.0; w1 = lane + 32 < k ? (double)n1 : 0.0; }
        const double sum = az_warp_sum(w0 + w1);
        const double p0 = w0 / sum, p1 = w1 / sum;
        if (pi_row) {
            if (lane < k) pi_row[R::net_index(m0, d)] = (float)p0;
            if (lane + 32 < k) pi_row[R::net_index(m1, d)] = (float)p1;
        }
        // inverse-CDF draw in child order: inclusive prefix sum of p0 over lanes, then the 33rd child
        double c = p0;
#pragma unroll
        for (int off = 1; off < 32; off <<= 1) {
            double o = __shfl_up_sync(AZ_FULL, c, off);
            if (lane >= off) c += o;
        }
        az_u32x4 r = az_draw4(t.seed, (uint32_t)g, event, 0u, 0x43484F4Fu);
        const double u = az_uniform53(r.x, r.y);
        const uint32_t hit = __ballot_sync(AZ_FULL, lane < k && p0 > 0.0 && u < c);
        if (hit) {
            chosen_move = __shfl_sync(AZ_FULL, m0, __ffs(hit) - 1);
        } else {
            // u beyond the first 32 children: the 33rd if it exists, else the last child with p > 0
            const uint32_t pos = __ballot_sync(AZ_FULL, lane < k && p0 > 0.0);
            int last = 31 - __clz(pos);
            chosen_move = (k > 32 && __shfl_sync(AZ_FULL, p1, 0) > 0.0) ? __shfl_sync(AZ_FULL, m1, 0)
                                                                       : __shfl_sync(AZ_FULL, m0, last);
        }
    }
    if (lane == 0) t.rng_ctr[g] = event + 1u;
    return chosen_move;
}

template <int GAME>
__global__ void __launch_bounds__(AZ_WARPS_PER_CTA * 32) k_tree_choose(az_tree t, const float* __restrict__ temp,
                                                                     uint8_t* __restrict__ action,
                                                                     float* __restrict__ pi) {
    using R = az_rules<GAME>;
    const int g = blockIdx.x * AZ_WARPS_PER_CTA + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= t.G || !t.active[g]) return;
    const az_dims d = R::dims(t.n, t.height);
    const int chosen = az_choose_warp<GAME>(t, d, g, lane, temp ? temp[g] : 0.f,
                                            pi ? pi + (size_t)g * R::action_size(d) : nullptr);
    if (lane == 0) action[g] = (uint8_t)chosen;
}

// ------------------------------------------------------------------------------------------------ advance

// MCT.change_root (mcts.py:118-125) for every active game + the move on the root board.  The kept subtree is
// copied breadth-first into the other half of the game's pool (32 nodes per round: each lane looks at one copied
// node, an exclusive warp scan of their child counts assigns the destination blocks, then the child blocks are
// copied with coalesced 32-byte-per-lane loads/stores), so the root is always node 0 and a game's tree stays
// dense no matter how long the game lasts.
// Body of az_tree_advance for game g by its warp; returns false (status bit set, game untouched) for an illegal move,
// else the new root position and the size of the re-rooted tree.
template <int GAME>
__device__ __forceinline__ bool az_advance_warp(const az_tree& t, const az_dims& d, int g, int lane, int a,
                                                uint64_t* new_own, uint64_t* new_opp, int* new_player,
                                                uint32_t* new_nodes) {
    using R = az_rules<GAME>;
    uint64_t own = t.root_own[g], opp = t.root_opp[g];
    const bool ok = R::is_legal(az_warp_rules<GAME>::legal(own, opp, d, lane, nullptr), a);
    if (!ok) {
        if (lane == 0) t.status[g] |= AZ_STATUS_ILLEGAL;
        return false;
    }
    const int h = t.half[g];
    const az_node* src = t.nodes + ((size_t)g * 2 + h) * (size_t)t.capacity;
    az_node* dst = t.nodes + ((size_t)g * 2 + (1 - h)) * (size_t)t.capacity;
    NodeRec root = az_load_node(&src[0]);
    int found = -1;
    if (root.flags() & AZ_NODE_EXPANDED) {
        for (int j = lane; j < root.n_children(); j += 32) {
            NodeRec c = az_load_node(&src[root.first_child + j]);
            if (c.move() == a) found = j;
        }
        const uint32_t who = __ballot_sync(AZ_FULL, found >= 0);
        found = who ? __shfl_sync(AZ_FULL, found, __ffs(who) - 1) : -1;
    }
    uint32_t n_dst = 1u;
    if (found < 0) {
        if (lane == 0) az_store_node(&dst[0], 0.0, 0.f, 0u, 0u, az_make_meta(0, a, 0u));
    } else {
        NodeRec nr = az_load_node(&src[root.first_child + found]);
        if (lane == 0) az_store_node(&dst[0], nr.Q, nr.P, nr.N, nr.first_child, nr.meta);
        __syncwarp();
        uint32_t i = 0u;
        while (i < n_dst) {
            const uint32_t cnt = min(32u, n_dst - i);
            NodeRec r;
            r.first_child = 0u; r.meta = 0u;
            if ((uint32_t)lane < cnt) r = az_load_node(&dst[i + lane]);
            const bool has = (uint32_t)lane < cnt && r.first_child != 0u;
            const uint32_t kk = has ? (uint32_t)r.n_children() : 0u;
            uint32_t incl = kk;
#pragma unroll
            for (int off = 1; off < 32; off <<= 1) {
                uint32_t o = __shfl_up_sync(AZ_FULL, incl, off);
                if (lane >= off) incl += o;
            }
            const uint32_t total = __shfl_sync(AZ_FULL, incl, 31);
            const uint32_t my_base = n_dst + incl - kk;
            if (has) reinterpret_cast<uint32_t*>(&dst[i + lane])[4] = my_base;
            uint32_t mask = __ballot_sync(AZ_FULL, has);
            while (mask) {
                const int l = __ffs(mask) - 1;
                mask &= mask - 1;
                const uint32_t sfc = __shfl_sync(AZ_FULL, r.first_child, l);
                const uint32_t kc = __shfl_sync(AZ_FULL, kk, l);
                const uint32_t db = __shfl_sync(AZ_FULL, my_base, l);
                for (uint32_t j = lane; j < kc; j += 32) {
                    const uint4* s = reinterpret_cast<const uint4*>(&src[sfc + j]);
                    uint4* dq = reinterpret_cast<uint4*>(&dst[db + j]);
                    uint4 x = s[0], y = s[1];
                    dq[0] = x;
                    dq[1] = y;
                }
            }
            n_dst += total;
            i += cnt;
            __syncwarp();
        }
    }
    az_warp_rules<GAME>::apply(own, opp, a, d, lane);
    const int player = -(int)t.root_player[g];
    __syncwarp();
    if (lane == 0) {
        t.root_own[g] = own;
        t.root_opp[g] = opp;
        t.root_player[g] = (int8_t)player;
        t.half[g] = (uint8_t)(1 - h);
        t.n_alloc[g] = n_dst;
        t.leaf_kind[g] = AZ_LEAF_IDLE;
    }
    *new_own = own; *new_opp = opp; *new_player = player; *new_nodes = n_dst;
    return true;
}

template <int GAME>
__global__ void __launch_bounds__(AZ_WARPS_PER_CTA * 32) k_tree_advance(az_tree t, const uint8_t* __restrict__ action) {
    const int g = blockIdx.x * AZ_WARPS_PER_CTA + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= t.G || !t.active[g]) return;
    const int a = action[g];
    if (a == 255) return;
    const az_dims d = az_rules<GAME>::dims(t.n, t.height);
    uint64_t own, opp;
    int player;
    uint32_t n_nodes;
    az_advance_warp<GAME>(t, d, g, lane, a, &own, &opp, &player, &n_nodes);
}

// One ply of batched self-play for every active game, by the game's warp (SelfPlayEngine.play): record the position
// in the history (trainer.py:241-246), choose the move and the policy target (mcts.py:95-116, players.py:184-189),
// play it and re-root (mcts.py:118-125), test the new position (othello.py:212-229) and retire finished games.
// Replaces az_tree_choose + az_tree_advance + az_*_terminal_score + a dozen elementwise launches per ply.
//   slot_game  game id of every slot (NULL = identity): the history / winner arrays are indexed by game id
//   h_*        history row of this ply, [n_games]; h_pi [n_games][A]
//   counters   int64[8] of this ply, zeroed by the caller: [0] games still live after the ply, [1] largest tree (nodes)
//              among them, [2] simulations of the searches just consumed, [3] moves played, [4] largest tree before
//              the move (pool high-water mark)
template <int GAME>
__global__ void __launch_bounds__(AZ_WARPS_PER_CTA * 32) k_tree_move(az_tree t, float temp, const int32_t* __restrict__ slot_game,
                                                                   uint64_t* __restrict__ h_own, uint64_t* __restrict__ h_opp,
                                                                   int8_t* __restrict__ h_player, uint8_t* __restrict__ h_live,
                                                                   float* __restrict__ h_pi, int8_t* __restrict__ winner,
                                                                   unsigned long long* __restrict__ counters) {
    using R = az_rules<GAME>;
    const int g = blockIdx.x * AZ_WARPS_PER_CTA + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= t.G) return;
    const bool live = t.active[g] != 0;
    if (!live) {
        if (!slot_game && lane == 0) h_live[g] = 0;
        return;
    }
    const int gid = slot_game ? slot_game[g] : g;
    const az_dims d = R::dims(t.n, t.height);
    const int A = R::action_size(d);
    if (lane == 0) {
        h_own[gid] = t.root_own[g];
        h_opp[gid] = t.root_opp[g];
        h_player[gid] = t.root_player[g];
        h_live[gid] = 1;
        atomicAdd(&counters[2], (unsigned long long)t.sims_done[g]);
        atomicAdd(&counters[3], 1ULL);
        atomicMax(&counters[4], (unsigned long long)t.n_alloc[g]);
    }
    const int a = az_choose_warp<GAME>(t, d, g, lane, temp, h_pi + (size_t)gid * A);
    __syncwarp();
    uint64_t own, opp;
    int player;
    uint32_t n_nodes;
    if (!az_advance_warp<GAME>(t, d, g, lane, a, &own, &opp, &player, &n_nodes)) return;
    bool terminal;
    az_warp_rules<GAME>::legal(own, opp, d, lane, &terminal);
    if (lane == 0) {
        if (terminal) {
            winner[gid] = (int8_t)R::winner(own, opp, player, d);
            t.active[g] = 0;
        } else {
            atomicAdd(&counters[0], 1ULL);
            atomicMax(&counters[1], (unsigned long long)n_nodes);
        }
    }
}

// ------------------------------------------------------------------------------------------------ hash evaluator

__device__ __forceinline__ uint64_t az_mix64(uint64_t z) {
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL;
    z ^= z >> 27; z *= 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return z;
}

// One warp per position: lane a computes probs[a], a + 32, ...; see tests/hash_eval.py for the definition.
__global__ void __launch_bounds__(AZ_WARPS_PER_CTA * 32) k_eval_hash(const uint64_t* __restrict__ own,
                                                                   const uint64_t* __restrict__ opp, int A, int64_t G,
                                                                   uint64_t salt, float* __restrict__ probs,
                                                                   float* __restrict__ value) {
    const int64_t g = (int64_t)blockIdx.x * AZ_WARPS_PER_CTA + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= G) return;
    const uint64_t h = az_mix64(own[g] * 0x9E3779B97F4A7C15ULL + az_mix64(opp[g] + 0xD1B54A32D192ED03ULL) + salt);
    for (int a = lane; a < A; a += 32) {
        uint64_t q = ((az_mix64(h + (uint64_t)a * 0x632BE59BD9B4E019ULL) >> 40) & 31ULL) + 1ULL;
        probs[g * A + a] = (float)q / 4096.0f;
    }
    if (lane == 0) value[g] = (float)((int)((h >> 20) & 0x1FFULL) - 256) / 256.0f;
}

// ------------------------------------------------------------------------------------------------ C ABI

static int az_check_tree(const az_tree* t) {
    if (!t || t->G <= 0 || t->capacity < 64 || t->max_depth < 2) return AZ_ERR_ARG;
    if (t->game == AZ_GAME_OTHELLO) {
        if (!az_valid_n(t->n)) return AZ_ERR_ARG;
    } else if (t->game == AZ_GAME_CONNECT4) {
        if (t->n < 4 || t->n > 8 || t->height < 4 || t->height > 8) return AZ_ERR_ARG;
    } else if (t->game == AZ_GAME_TICTACTOE) {
        if (t->n != 3 || t->height != 3) return AZ_ERR_ARG;
    } else {
        return AZ_ERR_ARG;
    }
    if (!t->nodes || !t->half || !t->n_alloc || !t->root_own || !t->root_opp || !t->root_player || !t->active ||
        !t->sims_done || !t->rng_ctr || !t->status || !t->leaf_node || !t->leaf_kind || !t->leaf_own ||
        !t->leaf_opp || !t->leaf_player || !t->path || !t->path_len)
        return AZ_ERR_ARG;
    if (t->mode != AZ_MODE_ROLLOUT && t->mode != AZ_MODE_NEURAL) return AZ_ERR_ARG;
    if (t->mode == AZ_MODE_NEURAL && (!t->probs || !t->value)) return AZ_ERR_ARG;
    if (t->mode == AZ_MODE_ROLLOUT && !t->value) return AZ_ERR_ARG;
    if (t->n_leaves < 0 || t->n_leaves > 8 || (t->n_leaves > 1 && t->mode != AZ_MODE_NEURAL)) return AZ_ERR_ARG;
    if (t->logits) {                                            // raw network rows: A policy logits + the value logit
        const int A = t->game == AZ_GAME_OTHELLO ? t->n * t->n + 1 : (t->game == AZ_GAME_CONNECT4 ? t->n : 9);
        if (t->mode != AZ_MODE_NEURAL || t->logits_ld < A + 1) return AZ_ERR_ARG;
    }
    return AZ_OK;
}

static inline int az_warp_grid(int64_t G) { return (int)((G + AZ_WARPS_PER_CTA - 1) / AZ_WARPS_PER_CTA); }

// Launches kernel<GAME><<<grid, block, 0, stream>>>(args...) for the game of the tree.
#define AZ_LAUNCH_GAME(game, kernel, grid, block, stream, ...)                                     \
    do {                                                                                           \
        if ((game) == AZ_GAME_OTHELLO) kernel<AZ_GAME_OTHELLO><<<grid, block, 0, stream>>>(__VA_ARGS__); \
        else if ((game) == AZ_GAME_CONNECT4) kernel<AZ_GAME_CONNECT4><<<grid, block, 0, stream>>>(__VA_ARGS__); \
        else kernel<AZ_GAME_TICTACTOE><<<grid, block, 0, stream>>>(__VA_ARGS__);                    \
    } while (0)

extern "C" int az_tree_reset(const az_tree* t, const uint8_t* mask, void* stream) {
    int rc = az_check_tree(t);
    if (rc) return rc;
    k_tree_reset<<<(t->G + 255) / 256, 256, 0, (cudaStream_t)stream>>>(*t, mask);
    AZ_CHECK_LAUNCH("k_tree_reset");
    return AZ_OK;
}

extern "C" int az_tree_step(const az_tree* t, int n_sim, int do_finish, int do_select, void* stream) {
    int rc = az_check_tree(t);
    if (rc) return rc;
    if (n_sim < 0) return AZ_ERR_ARG;
    // Warps (= games) per CTA of the parity step kernel.  All games of a batch must be resident at once (one warp each, a
    // long dependent chain per warp), and 28 warps fill an SM's register file either way; what the block size decides is
    // how many CTAs have to be launched behind the network kernel that held every SM: at 4096 games one CTA of 28 warps
    // per SM (147 CTAs) against 1024 CTAs of 4 warps is 0.5 us per simulation step (52.6 against 53.1 us; 8 warps per CTA
    // would leave room for 3 CTAs = 24 warps per SM and split the batch in two waves: 57.8 us).  Small batches keep small
    // CTAs so that they spread over the SMs.  AZ_B200_TREE_WARPS overrides.
    static const int forced_warps = []() {
        const char* v = getenv("AZ_B200_TREE_WARPS");
        const int w = v ? atoi(v) : 0;
        return w >= 1 && w <= AZ_WARPS_PER_CTA * AZ_TREE_MIN_CTAS ? w : 0;
    }();
    int step_warps = forced_warps;
    if (!step_warps) {
        static int n_sm = 0;
        if (!n_sm) {
            int dev = 0;
            if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n_sm, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n_sm < 1)
                n_sm = 148;
        }
        const int64_t per_sm = (t->G + n_sm - 1) / n_sm;
        step_warps = per_sm <= AZ_WARPS_PER_CTA ? AZ_WARPS_PER_CTA
                   : (per_sm <= 2 * AZ_WARPS_PER_CTA * AZ_TREE_MIN_CTAS / 4 ? 2 * AZ_WARPS_PER_CTA * AZ_TREE_MIN_CTAS / 4
                                                                          : AZ_WARPS_PER_CTA * AZ_TREE_MIN_CTAS);
    }
    const bool multi = t->n_leaves >= 1 && t->mode == AZ_MODE_NEURAL;
    const int wpc = multi ? AZ_WARPS_PER_CTA : step_warps;
    const dim3 grid((unsigned)((t->G + wpc - 1) / wpc)), block(wpc * 32);
    cudaError_t e;
#define AZ_STEP(GAME, MODE) az_launch_pdl(k_tree_step<GAME, MODE>, grid, block, 0, (cudaStream_t)stream, *t, n_sim, do_finish, do_select)
    const bool neural = t->mode == AZ_MODE_NEURAL;
    if (t->n_leaves >= 1 && neural) {          // throughput mode (n_leaves = 0: the parity kernel)
#define AZ_STEP_MULTI(GAME) az_launch_pdl(k_tree_step_multi<GAME>, grid, block, 0, (cudaStream_t)stream, *t, n_sim, do_finish, do_select)
        e = t->game == AZ_GAME_OTHELLO ? AZ_STEP_MULTI(AZ_GAME_OTHELLO)
            : (t->game == AZ_GAME_CONNECT4 ? AZ_STEP_MULTI(AZ_GAME_CONNECT4) : AZ_STEP_MULTI(AZ_GAME_TICTACTOE));
#undef AZ_STEP_MULTI
        if (e != cudaSuccess) return az_set_cuda_error(e, "k_tree_step_multi");
        AZ_CHECK_LAUNCH("k_tree_step_multi");
        return AZ_OK;
    }
    if (t->game == AZ_GAME_OTHELLO)
        e = neural ? AZ_STEP(AZ_GAME_OTHELLO, AZ_MODE_NEURAL) : AZ_STEP(AZ_GAME_OTHELLO, AZ_MODE_ROLLOUT);
    else if (t->game == AZ_GAME_CONNECT4)
        e = neural ? AZ_STEP(AZ_GAME_CONNECT4, AZ_MODE_NEURAL) : AZ_STEP(AZ_GAME_CONNECT4, AZ_MODE_ROLLOUT);
    else
        e = neural ? AZ_STEP(AZ_GAME_TICTACTOE, AZ_MODE_NEURAL) : AZ_STEP(AZ_GAME_TICTACTOE, AZ_MODE_ROLLOUT);
#undef AZ_STEP
    if (e != cudaSuccess) return az_set_cuda_error(e, "k_tree_step");
    AZ_CHECK_LAUNCH("k_tree_step");
    return AZ_OK;
}

extern "C" int az_tree_rollout(const az_tree* t, void* stream) {
    int rc = az_check_tree(t);
    if (rc) return rc;
    AZ_LAUNCH_GAME(t->game, k_tree_rollout, (t->G + 127) / 128, 128, (cudaStream_t)stream, *t);
    AZ_CHECK_LAUNCH("k_tree_rollout");
    return AZ_OK;
}

extern "C" int az_tree_root_stats(const az_tree* t, int32_t* N, double* Q, float* P, uint32_t* root_N, void* stream) {
    int rc = az_check_tree(t);
    if (rc) return rc;
    k_tree_root_stats<<<az_warp_grid(t->G), AZ_WARPS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(*t, N, Q, P, root_N);
    AZ_CHECK_LAUNCH("k_tree_root_stats");
    return AZ_OK;
}

extern "C" int az_tree_choose(const az_tree* t, const float* temp, uint8_t* action, float* pi, void* stream) {
    int rc = az_check_tree(t);
    if (rc) return rc;
    if (!action) return AZ_ERR_ARG;
    AZ_LAUNCH_GAME(t->game, k_tree_choose, az_warp_grid(t->G), AZ_WARPS_PER_CTA * 32, (cudaStream_t)stream, *t, temp,
                   action, pi);
    AZ_CHECK_LAUNCH("k_tree_choose");
    return AZ_OK;
}

extern "C" int az_tree_advance(const az_tree* t, const uint8_t* action, void* stream) {
    int rc = az_check_tree(t);
    if (rc) return rc;
    if (!action) return AZ_ERR_ARG;
    AZ_LAUNCH_GAME(t->game, k_tree_advance, az_warp_grid(t->G), AZ_WARPS_PER_CTA * 32, (cudaStream_t)stream, *t, action);
    AZ_CHECK_LAUNCH("k_tree_advance");
    return AZ_OK;
}

extern "C" int az_tree_move(const az_tree* t, float temp, const int32_t* slot_game, uint64_t* h_own, uint64_t* h_opp,
                            int8_t* h_player, uint8_t* h_live, float* h_pi, int8_t* winner, int64_t* counters,
                            void* stream) {
    int rc = az_check_tree(t);
    if (rc) return rc;
    if (!h_own || !h_opp || !h_player || !h_live || !h_pi || !winner || !counters || temp < 0.f) return AZ_ERR_ARG;
    AZ_LAUNCH_GAME(t->game, k_tree_move, az_warp_grid(t->G), AZ_WARPS_PER_CTA * 32, (cudaStream_t)stream, *t, temp,
                   slot_game, h_own, h_opp, h_player, h_live, h_pi, winner, (unsigned long long*)counters);
    AZ_CHECK_LAUNCH("k_tree_move");
    return AZ_OK;
}

extern "C" int az_eval_hash(const uint64_t* own, const uint64_t* opp, int action_size, int64_t G, uint64_t salt,
                            float* probs, float* value, void* stream) {
    if (action_size < 1 || action_size > AZ_CANVAS_ACTIONS || G < 0 || (G > 0 && (!own || !opp || !probs || !value)))
        return AZ_ERR_ARG;
    if (G == 0) return AZ_OK;
    k_eval_hash<<<az_warp_grid(G), AZ_WARPS_PER_CTA * 32, 0, (cudaStream_t)stream>>>(own, opp, action_size, G, salt, probs,
                                                                                    value);
    AZ_CHECK_LAUNCH("k_eval_hash");
    return AZ_OK;
}
