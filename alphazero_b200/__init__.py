"""alphazero_b200 -- B200-native (sm_100a) batched MCTS self-play engine behind t0m1ab/alphazero's API.

Only the reference's hot path lives here: Othello bitboard kernels, SoA MCTS tree kernels, leaf encode/scatter
around the reference's own PyTorch policy-value network, and the host-side mirror of the reference interface
(Board / Player / PolicyValueNetwork / MCT / MCTSPlayer / AlphaZeroPlayer / Arena / self-play).
There is no CPU fallback: every compute entry point goes through libaz_b200.so and raises if it is missing.
"""
__version__ = "0.1.0"
