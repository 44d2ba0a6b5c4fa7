#!/usr/bin/env python
"""Driver for `ncu`: three launches of the tcgen05 dgrad + GELU-backward kernel at the GPT-2 shape — capture the last."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_lion_pytorch_b200.ops import mlp  # noqa: E402

m, n, k = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (20480, 3072, 768)))
g = torch.Generator(device="cuda").manual_seed(0)
dy = torch.randn(m, k, device="cuda", generator=g).to(torch.bfloat16)
w = (torch.randn(n, k, device="cuda", generator=g) * k ** -0.5).to(torch.bfloat16)
pre = torch.randn(m, n, device="cuda", generator=g).to(torch.bfloat16)
for _ in range(3):
    dpre, db = mlp.dgrad_dgelu(dy, w, pre)
torch.cuda.synchronize()
print("ok", float(db.float().abs().mean()))
