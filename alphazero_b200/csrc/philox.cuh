// Philox4x32-10 counter-based RNG (Salmon et al., "Parallel random numbers: as easy as 1, 2, 3", SC'11).
// The reference draws from NumPy's global MT19937 (alphazero/utils.py:34, mcts.py:165,238, othello.py:194,
// players.py:189); stream-level reproduction on a GPU is out of scope, so every random choice
// on the device is keyed by (seed, game id, event counter) which makes results independent of grid shape and of
// the number of GPUs.  The oracle (oracle/az_oracle.c) implements the same function for bit-exact playout parity.
#pragma once
#include <stdint.h>

#ifdef __CUDACC__
#define AZ_PHILOX_HD __host__ __device__ __forceinline__
#else
#define AZ_PHILOX_HD static inline
#endif

struct az_u32x4 { uint32_t x, y, z, w; };

AZ_PHILOX_HD void az_mulhilo32(uint32_t a, uint32_t b, uint32_t& hi, uint32_t& lo) {
    uint64_t p = (uint64_t)a * (uint64_t)b;
    hi = (uint32_t)(p >> 32);
    lo = (uint32_t)p;
}

AZ_PHILOX_HD az_u32x4 az_philox4x32_10(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0, lo0, hi1, lo1;
        az_mulhilo32(0xD2511F53u, c0, hi0, lo0);
        az_mulhilo32(0xCD9E8D57u, c2, hi1, lo1);
        uint32_t n0 = hi1 ^ c1 ^ k0;
        uint32_t n2 = hi0 ^ c3 ^ k1;
        c0 = n0; c1 = lo1; c2 = n2; c3 = lo0;
        k0 += 0x9E3779B9u;
        k1 += 0xBB67AE85u;
    }
    az_u32x4 o; o.x = c0; o.y = c1; o.z = c2; o.w = c3;
    return o;
}

// Uniform integer in [0, n) from one 32-bit draw by multiply-high (exact integer arithmetic, same on CPU and GPU).
AZ_PHILOX_HD uint32_t az_bounded(uint32_t r, uint32_t n) { return (uint32_t)(((uint64_t)r * (uint64_t)n) >> 32); }

// Uniform double in (0, 1) from two 32-bit draws (53 random bits, never 0).
AZ_PHILOX_HD double az_uniform53(uint32_t a, uint32_t b) {
    uint64_t v = (((uint64_t)a << 21) ^ (uint64_t)(b >> 11)) & ((1ULL << 53) - 1ULL);
    return ((double)v + 0.5) * (1.0 / 9007199254740992.0);
}
