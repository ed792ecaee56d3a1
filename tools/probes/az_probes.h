/* az_probes.h -- measurement-only micro-benchmarks (roofline denominators, design probes).  NOT part of the product
 * ABI: built into tools/probes/libaz_probes.so by tools/probes/build.py, never loaded by alphazero_b200. */
#ifndef AZ_PROBES_H
#define AZ_PROBES_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* Integer-ALU issue-rate micro-benchmark (the roofline denominator of the board kernels):
 * launches 148*8 CTAs of 256 threads, each thread running `iters` x 32 instructions of one kind in 8 independent
 * chains.  kind 0 LOP3, 1 SHF, 2 IADD3, 3 shift-mask-or (the movegen step, 2 instructions).  out: >= 148*8*256
 * uint32 (sink).  *ops_per_thread (host, optional) = instructions each thread executed in the timed loop. */
int az_peak_int32(int kind, int iters, uint32_t* out, int64_t out_len, int64_t* ops_per_thread, void* stream);

/* tcgen05.mma issue-rate probe (the evidence behind the trunk's N = 96 design): every SM issues
 * `reps` x `k_steps` MMAs of shape m x n x 16 (bf16 from shared memory, K-major no-swizzle, fp32 in TMEM), the
 * k_steps of a rep accumulating into one of `regions` (1 or 2) accumulators; cycles[sm] (device, 148 int64) = SM
 * clocks from the first issue to the completion of the last.
 * flags: 1 = every MMA accumulates (no zero-initialising MMA), 2 = tcgen05.commit after every chain, 4 = three warps
 * keep reading 96 accumulator columns with tcgen05.ld meanwhile. */
int az_peak_mma(int m, int n, int k_steps, int reps, int regions, int flags, int64_t* cycles, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* AZ_PROBES_H */
