"""NumPy walk-through of k_net_fcstack's data movement (csrc/fcstack_kernels.cu), for the CPU test-suite.

It is not an oracle of the network (that is the fp32 PyTorch module); it restates the KERNEL's address arithmetic --
which bytes each CTA of a cluster copies into which stage, how the tensor core's K-major no-swizzle descriptors read a
stage, where the epilogues store the next layer's operand blocks and the logits -- so that the host-side packing
(fused_net.pack_fcstack / pack_rows, the trunk's packed feature layout) can be checked against it without a GPU.
"""
import numpy as np


def _bf16_round(x: np.ndarray) -> np.ndarray:
    """float32 -> nearest-even bfloat16 -> float32."""
    u = np.ascontiguousarray(x, dtype=np.float32).view(np.uint32).astype(np.uint64)
    r = ((u + 0x7FFF + ((u >> 16) & 1)) >> 16) << 16
    return r.astype(np.uint32).view(np.float32)


def _bf16_to_f32(bits: np.ndarray) -> np.ndarray:
    return (bits.astype(np.uint32) << 16).view(np.float32)


def _f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    return (_bf16_round(x).view(np.uint32) >> 16).astype(np.uint16)


def _read_operand(stage: np.ndarray, rows: int, ksteps: int, lbo: int) -> np.ndarray:
    """Decode what `ksteps` tcgen05.mma of K = 16 read from a stage (uint8) through the descriptor
    umma_desc(start + ks * 2 * lbo, lbo, 128): element (row i, k) of step ks at ks*2*lbo + (k//8)*lbo + i*16 + (k%8)*2."""
    out = np.zeros((rows, ksteps * 16), dtype=np.float32)
    u16 = stage.view(np.uint16)
    for ks in range(ksteps):
        for k in range(16):
            base = ks * 2 * lbo + (k // 8) * lbo + (k % 8) * 2
            idx = (base + np.arange(rows) * 16) // 2
            out[:, ks * 16 + k] = _bf16_to_f32(u16[idx])
    return out


def emulate(x_packed: np.ndarray, blob: np.ndarray, plan: dict, G: int, ld: int, RT: int, NS: int) -> np.ndarray:
    """x_packed, blob: uint8 arrays; returns logits float32 [G][ld] (bf16-rounded values, untouched entries NaN)."""
    K, Nc, kch = plan["K"], plan["Nc"], plan["kch"]
    hpc, n_out = plan["hpc"], plan["n_out"]
    tiles = (G + 127) // 128
    clusters = (tiles + RT - 1) // RT
    T = clusters * RT
    assert x_packed.size >= T * 128 * K[0] * 2
    h = [x_packed, np.zeros(T * 128 * K[1] * 2, np.uint8), np.zeros(T * 128 * K[2] * 2, np.uint8)]
    logits = np.full((G, ld), np.nan, dtype=np.float32)
    for cl in range(clusters):
        for L in range(3):
            nch = K[L] // kch[L]
            a_bytes, b_bytes = kch[L] * 256, kch[L] * Nc[L] * 2
            a_part, b_part = a_bytes // NS, b_bytes // RT
            bps = nch // NS
            acc = {(r, c): np.zeros((128, Nc[L]), np.float32) for r in range(RT) for c in range(NS)}
            for q in range(nch):
                kc = q          # ascending K: all blocks of column slice 0, then slice 1, ...
                # every CTA's stage after all issuers' (multicast) parts have landed
                stage_a = {r: np.zeros(a_bytes, np.uint8) for r in range(RT)}
                stage_b = {c: np.zeros(b_bytes, np.uint8) for c in range(NS)}
                for r in range(RT):
                    m = cl * RT + r
                    for c in range(NS):       # issuer (r, c): part c of the A block to row group r, part r of B to slice c
                        src = m * nch * a_bytes + kc * a_bytes + c * a_part
                        stage_a[r][c * a_part:(c + 1) * a_part] = h[L][src:src + a_part]
                        srcb = plan["w_off"][L] + c * nch * b_bytes + kc * b_bytes + r * b_part
                        stage_b[c][r * b_part:(r + 1) * b_part] = blob[srcb:srcb + b_part]
                for r in range(RT):
                    A = _read_operand(stage_a[r], 128, kch[L] // 16, 2048)
                    for c in range(NS):
                        B = _read_operand(stage_b[c], Nc[L], kch[L] // 16, 16 * Nc[L])
                        acc[(r, c)] += A @ B.T
            for r in range(RT):
                m = cl * RT + r
                for c in range(NS):
                    bias = blob[plan["b_off"][L]:plan["b_off"][L] + NS * Nc[L] * 4].view(np.float32)[c * Nc[L]:(c + 1) * Nc[L]]
                    y = acc[(r, c)] + bias[None, :]
                    if L < 2:
                        bw = kch[L + 1]
                        nblk, nchn = Nc[L] // bw, K[L + 1] // bw
                        y = np.maximum(y, 0.0)
                        bits = _f32_to_bf16_bits(y)
                        dst16 = h[L + 1].view(np.uint16)
                        for j in range(nblk):
                            kcn = c * nblk + j
                            for pl in range(bw // 8):
                                for row in range(128):
                                    off = (m * nchn * bw * 256 + kcn * bw * 256 + pl * 2048 + row * 16) // 2
                                    dst16[off:off + 8] = bits[row, j * bw + pl * 8:j * bw + pl * 8 + 8]
                    else:
                        for col in range(min(hpc, Nc[2])):
                            hidx = c * hpc + col
                            if hidx < n_out:
                                rows = np.arange(128) + m * 128
                                ok = rows < G
                                logits[rows[ok], hidx] = _bf16_round(y[ok, col])
    return logits
