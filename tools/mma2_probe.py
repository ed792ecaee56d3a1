"""tcgen05.mma.cta_group::2 semantics probe (tools/probes/mma2_probe.cu): which half of B each CTA of the pair supplies
and which rows of D each CTA's TMEM holds.

    python tools/mma2_probe.py
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from tools.probes import build

L = build.lib()
torch.manual_seed(0)
for N in (128, 256, 64):
    A = torch.randn(256, 64).to(torch.bfloat16).cuda()
    B = torch.randn(N, 64).to(torch.bfloat16).cuda()
    want = (A.double() @ B.double().t()).float()
    for swap in (0, 1):
        D = torch.full((2, 128, N), float("nan"), device="cuda")
        rc = L.az_probe_mma2(A.data_ptr(), B.data_ptr(), D.data_ptr(), N, swap, 0, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        got = D.reshape(256, N)
        err_direct = (got - want).abs().max().item()
        # hypotheses: columns of the two halves swapped; rows of the two CTAs swapped
        cs = torch.cat([want[:, N // 2:], want[:, :N // 2]], 1)
        rs = torch.cat([want[128:], want[:128]], 0)
        print(f"N={N} swap_b={swap} rc={rc}: |D - A B^T| max {err_direct:.4f} | vs column-halves swapped {(got - cs).abs().max().item():.4f}"
              f" | vs row-halves swapped {(got - rs).abs().max().item():.4f} | nan {int(torch.isnan(got).sum())}")

# A operand whose descriptor starts `shift` rows (16 bytes each) off the data: D[i] = A[i + shift] . B^T per CTA, rows
# outside the CTA's 128 read the zero margin
N = 96
A = torch.randn(256, 64).to(torch.bfloat16).cuda()
B = torch.randn(N, 64).to(torch.bfloat16).cuda()
for shift in (0, 1, -1, 3, 8, -8):
    D = torch.full((2, 128, N), float("nan"), device="cuda")
    rc = L.az_probe_mma2(A.data_ptr(), B.data_ptr(), D.data_ptr(), N, 0, shift, torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    As = torch.zeros(2, 128, 64, dtype=torch.float64, device="cuda")
    Ad = A.double().reshape(2, 128, 64)
    if shift >= 0:
        As[:, :128 - shift] = Ad[:, shift:]
    else:
        As[:, -shift:] = Ad[:, :128 + shift]
    want = (As @ B.double().t()).float()
    print(f"A start shifted by {shift:+d} rows: rc={rc} max |D - shifted(A) B^T| = {(D - want).abs().max().item():.5f}")
