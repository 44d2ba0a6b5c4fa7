#!/usr/bin/env python
"""Tiny driver for `ncu --set full`: one GPT-2-small micro-batch forward/backward (norm, colsum and
cross-entropy kernels at their real shapes) followed by the fused exchange kernel with 8 emulated
ranks on a GPT-2-sized bucket.  Run it plain first, then under ncu with -k regex:<kernels>."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_lion import Lion  # noqa: E402
from distributed_lion_pytorch_b200.models import build_model  # noqa: E402
from distributed_lion_pytorch_b200.ops import lion_ops  # noqa: E402
from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange  # noqa: E402

torch.manual_seed(0)
dev = torch.device("cuda")
model = build_model("gpt2", dtype=torch.bfloat16, device=dev, pretrained=False)
opt = Lion(model.parameters(), lr=1e-4, weight_decay=0.1)
ids = torch.randint(0, 50257, (20, 1024), device=dev)
for _ in range(2):
    model(ids, labels=ids).loss.backward()
    opt.step()
    opt.zero_grad()
torch.cuda.synchronize()
del model, opt
W, dtype = 8, torch.bfloat16
n = (124_440_000 // W) // lion_ops.fused_tile(dtype) * lion_ops.fused_tile(dtype)
ex = EmulatedExchange(W, n, dtype, dev)
P = torch.randn(W, n, device=dev).to(dtype)
G = torch.randn(W, n, device=dev).to(dtype)
M = torch.zeros(W, n, device=dev, dtype=dtype)
for _ in range(2):
    ex.step(P, G, M, 1e-4, 0.1, 0.9, 0.99)
torch.cuda.synchronize()
ex.check_status()
print("profile driver done")
