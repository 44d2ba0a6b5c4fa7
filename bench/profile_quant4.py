#!/usr/bin/env python
"""Tiny driver for `ncu --set full`: NF4 quantise + expand of one Llama-2-7B up_proj weight
(11008 x 4096, bf16).  Run it plain first, then under ncu with -k regex:q4_dequantize_kernel."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_lion_pytorch_b200.ops import quant4  # noqa: E402

torch.manual_seed(0)
w = (torch.randn(11008, 4096, device="cuda") * 0.02).to(torch.bfloat16)
packed, absmax = quant4.quantize_4bit(w)
out = torch.empty(w.numel(), device="cuda", dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
for i in range(3):
    flush.fill_(i)
    quant4.dequantize_4bit(packed, absmax, w.shape, torch.bfloat16, out=out)
torch.cuda.synchronize()
err = (out.view_as(w).float() - w.float()).abs().max().item()
print(f"profile driver done, max abs error {err:.5f}")
