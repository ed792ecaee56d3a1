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


def _kc2(q: int, c: int, bpp: int) -> int:
    """f2_kc of csrc/fcstack2_kernels.cu: the K chunk of fc2's q-th chunk in CTA c."""
    per = 4 * bpp
    h, rem = divmod(q, per)
    cs, j = ((rem // bpp) + c) & 3, rem % bpp
    return cs * 2 * bpp + h * bpp + j


def emulate2(x_packed: np.ndarray, blob: np.ndarray, plan: dict, G: int, ld: int) -> np.ndarray:
    """The same walk-through for k_net_fcstack2 (variant 2 of the plan): fc1 in two passes, h1 through global blocks
    fetched stage by stage in the order f2_kc names, h2 in shared memory, the heads split over K and their four partial
    sums added by the CTA that owns the row quarter."""
    assert plan["variant"] in (2, 3)
    pair = plan["variant"] == 3       # k_net_fcstack3: every B block comes in two column halves, one per CTA of a pair
    K1, N1, N2 = plan["K"]
    PW, Nc2, n3 = plan["Nc"]
    n_out = plan["n_out"]
    nkx, nch2, bpp = K1 // 64, N1 // 64, PW // 64
    cps = 1 if nkx & 1 else 2
    cps2 = 2 if PW == 128 else 1
    n1p, n2 = nkx // cps, nch2 // cps2
    a1, b1 = cps * 16384, cps * 64 * PW * 2
    a2, b2, hb_bytes = cps2 * 16384, 64 * Nc2 * 2, Nc2 * n3 * 2
    tiles = (G + 127) // 128

    def b_operand(base, group, blocks_per_group, block, nbytes, cols, ksteps):
        """B of a stage: `nbytes` at block `block` of column group `group`; for a pair the two CTAs stage the halves
        2 group + r, r = 0, 1 (cols / 2 columns, nbytes / 2 bytes each) and the tensor cores read [half 0; half 1]."""
        if not pair:
            src = base + (group * blocks_per_group + block) * nbytes
            return _read_operand(blob[src:src + nbytes], cols, ksteps, 16 * cols)
        halves = []
        for r in range(2):
            src = base + ((group * 2 + r) * blocks_per_group + block) * (nbytes // 2)
            halves.append(_read_operand(blob[src:src + nbytes // 2], cols // 2, ksteps, 16 * (cols // 2)))
        return np.vstack(halves)

    h1 = np.zeros(tiles * 128 * N1 * 2, np.uint8)
    logits = np.full((G, ld), np.nan, dtype=np.float32)
    bias1 = blob[plan["b_off"][0]:plan["b_off"][0] + N1 * 4].view(np.float32)
    bias2 = blob[plan["b_off"][1]:plan["b_off"][1] + N2 * 4].view(np.float32)
    bias3 = blob[plan["b_off"][2]:plan["b_off"][2] + n3 * 4].view(np.float32)
    for m in range(tiles):
        X = x_packed[m * nkx * 16384:(m + 1) * nkx * 16384]
        for c in range(4):                                       # fc1: both passes of every CTA of the row tile
            for h in range(2):
                acc = np.zeros((128, PW), np.float32)
                for q in range(n1p):
                    A = _read_operand(X[q * a1:(q + 1) * a1], 128, cps * 4, 2048)
                    B = b_operand(plan["w_off"][0], c * 2 + h, n1p, q, b1, PW, cps * 4)
                    acc += A @ B.T
                bits = _f32_to_bf16_bits(np.maximum(acc + bias1[c * 2 * PW + h * PW:c * 2 * PW + (h + 1) * PW][None, :], 0.0))
                dst16 = h1.view(np.uint16)
                for g in range(PW // 32):                        # warp (quarter, g): 32 columns = half a block
                    j, hb = g >> 1, g & 1
                    kc = c * 2 * bpp + h * bpp + j
                    for pl in range(4):
                        for row in range(128):
                            off = ((m * nch2 + kc) * 16384 + row * 16 + (hb * 4 + pl) * 2048) // 2
                            dst16[off:off + 8] = bits[row, g * 32 + pl * 8:g * 32 + pl * 8 + 8]
        partial = np.zeros((4, 128, n3), np.float32)
        for c in range(4):                                       # fc2 and this CTA's K slice of the heads
            acc = np.zeros((128, Nc2), np.float32)
            seen = []
            for s in range(n2):
                kc = _kc2(s * cps2, c, bpp)
                seen += list(range(kc, kc + cps2))
                a0 = (m * nch2 + kc) * 16384
                A = _read_operand(h1[a0:a0 + a2], 128, cps2 * 4, 2048)
                # blocks of b2 bytes per chunk; a stage takes cps2 consecutive ones
                B = b_operand(plan["w_off"][1], c, nch2 // cps2, kc // cps2, cps2 * b2, Nc2, cps2 * 4)
                acc += A @ B.T
            assert sorted(seen) == list(range(nch2))             # every K chunk exactly once
            bits = _f32_to_bf16_bits(np.maximum(acc + bias2[c * Nc2:(c + 1) * Nc2][None, :], 0.0))
            h2s = np.zeros(Nc2 * 256, np.uint8)
            d16 = h2s.view(np.uint16)
            for g in range(Nc2 // 32):
                for pl in range(4):
                    for row in range(128):
                        off = ((g * 4 + pl) * 2048 + row * 16) // 2
                        d16[off:off + 8] = bits[row, g * 32 + pl * 8:g * 32 + pl * 8 + 8]
            A = _read_operand(h2s, 128, Nc2 // 16, 2048)
            B = b_operand(plan["w_off"][2], c, 1, 0, hb_bytes, n3, Nc2 // 16)
            partial[c] = A @ B.T
        for owner in range(4):                                   # rows 32 owner .. : own registers + bias, then the peers in sender order
            rows = slice(owner * 32, owner * 32 + 32)
            total = partial[owner][rows] + bias3[None, :]
            for si in range(3):
                total = total + partial[(owner + si + 1) & 3][rows]
            for rr in range(32):
                r = m * 128 + owner * 32 + rr
                if r < G:
                    logits[r, :n_out] = _bf16_round(total[rr, :n_out])
    return logits
