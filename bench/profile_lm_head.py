#!/usr/bin/env python
"""Driver for `ncu`: three launches of the tcgen05 LM-head kernel at the GPT-2 micro-batch shape
([20480, 768] x [50257, 768], bf16) — capture the last one."""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_lion_pytorch_b200.ops import lm_head  # noqa: E402

rows, vocab, hidden = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (20480, 50257, 768)))
g = torch.Generator(device="cuda").manual_seed(0)
h = torch.randn(rows, hidden, device="cuda", generator=g).to(torch.bfloat16)
w = (torch.randn(vocab, hidden, device="cuda", generator=g) * hidden ** -0.5).to(torch.bfloat16)
labels = torch.randint(0, vocab, (rows,), device="cuda", generator=g)
for _ in range(3):
    lp, _ = lm_head.lm_head_log_probs(h, w, labels, want_lse=False)
torch.cuda.synchronize()
print("ok", float(lp.mean()))
