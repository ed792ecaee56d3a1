// TEST-ONLY: compiles the HOST side of the device headers (othello_bits.cuh / philox.cuh are __host__ __device__)
// so the bit logic can be checked against the oracle on a machine without a GPU.  Never loaded by the product.
#include <stdint.h>
#include "../alphazero_b200/csrc/othello_bits.cuh"
#include "../alphazero_b200/csrc/philox.cuh"

extern "C" {
void hb_legal(const uint64_t* own, const uint64_t* opp, int n, int64_t G, uint64_t* legal) {
    uint64_t bm = az_board_mask(n);
    for (int64_t g = 0; g < G; ++g) legal[g] = az_legal_mask(own[g], opp[g], bm);
}
void hb_flips(const uint64_t* own, const uint64_t* opp, const uint8_t* action, int64_t G, uint64_t* flips) {
    for (int64_t g = 0; g < G; ++g) {
        uint64_t mv = action[g] < 64 ? 1ULL << action[g] : 0ULL;
        uint64_t f = mv ? az_flips(mv, own[g], opp[g]) : 0ULL;
        uint64_t f2 = 0;
        for (int d = 0; d < 8; ++d) f2 |= mv ? az_flips_dir_dyn(mv, own[g], opp[g], d) : 0ULL;
        flips[g] = (f == f2) ? f : ~0ULL;  // the per-direction (lane-split) form must agree with the fused one
    }
}
void hb_apply(uint64_t* own, uint64_t* opp, const uint8_t* action, int64_t G) {
    for (int64_t g = 0; g < G; ++g) az_apply(own[g], opp[g], action[g]);
}
void hb_terminal(const uint64_t* own, const uint64_t* opp, int n, int64_t G, uint8_t* term) {
    uint64_t bm = az_board_mask(n);
    for (int64_t g = 0; g < G; ++g) term[g] = az_is_terminal(own[g], opp[g], bm);
}
int hb_nth(uint64_t m, int k) { return az_nth_set_bit(m, k); }
void hb_philox(const uint32_t* c, const uint32_t* k, uint32_t* o) {
    az_u32x4 r = az_philox4x32_10(c[0], c[1], c[2], c[3], k[0], k[1]);
    o[0] = r.x; o[1] = r.y; o[2] = r.z; o[3] = r.w;
}
}

// ---- Connect4 (connect4_bits.cuh / rules.cuh host side)
#include "../include/az_b200.h"
#include "../alphazero_b200/csrc/rules.cuh"

extern "C" {
void hb_c4_legal(const uint64_t* own, const uint64_t* opp, int width, int64_t G, uint64_t* legal) {
    for (int64_t g = 0; g < G; ++g) legal[g] = az_c4_legal_cols(own[g], opp[g], width);
}
void hb_c4_apply(uint64_t* own, uint64_t* opp, const uint8_t* action, int width, int height, int64_t G) {
    az_dims d = az_rules<AZ_GAME_CONNECT4>::dims(width, height);
    for (int64_t g = 0; g < G; ++g) az_rules<AZ_GAME_CONNECT4>::apply(own[g], opp[g], action[g], d);
}
void hb_c4_terminal(const uint64_t* own, const uint64_t* opp, const int8_t* player, int width, int height, int64_t G,
                    uint8_t* term, int8_t* winner) {
    az_dims d = az_rules<AZ_GAME_CONNECT4>::dims(width, height);
    for (int64_t g = 0; g < G; ++g) {
        term[g] = az_rules<AZ_GAME_CONNECT4>::is_terminal(own[g], opp[g], d);
        winner[g] = (int8_t)az_rules<AZ_GAME_CONNECT4>::winner(own[g], opp[g], player[g], d);
    }
}
}
