// PolicyValueNetwork.predict's output transform (base.py:350-355: exp(log_softmax) of the policy logits; othello.py:382 /
// connect4.py:411: tanh of the value logit) on one warp, logits held lane-wise.  Shared by k_net_heads (which
// materialises probs / value) and k_tree_step (which, given az_tree.logits, consumes the logits row directly), so
// that both produce the same bits.
#pragma once
#include <cuda_bf16.h>
#include <stdint.h>

// x[i] = logit of action lane + 32*i (anything for actions >= A); returns p[i] = softmax over the A <= 96 actions
// (0 for actions >= A).  Every lane of the warp must call it.
__device__ __forceinline__ void az_heads_softmax(const float (&x)[3], int lane, int A, float (&p)[3]) {
    float v[3];
    float mx = -INFINITY;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        v[i] = (lane + 32 * i) < A ? x[i] : -INFINITY;
        mx = fmaxf(mx, v[i]);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, off));
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        p[i] = (lane + 32 * i) < A ? __expf(__fsub_rn(v[i], mx)) : 0.f;
        s = __fadd_rn(s, p[i]);
    }
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) s = __fadd_rn(s, __shfl_xor_sync(0xffffffffu, s, off));
    const float inv = __fdiv_rn(1.f, s);
#pragma unroll
    for (int i = 0; i < 3; ++i) p[i] = __fmul_rn(p[i], inv);
}

__device__ __forceinline__ float az_heads_value(__nv_bfloat16 logit) { return tanhf(__bfloat162float(logit)); }
