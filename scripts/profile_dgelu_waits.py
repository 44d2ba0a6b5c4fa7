#!/usr/bin/env python
"""Where the tcgen05 dgrad + GELU-backward kernel waits: per-wait-site cycle counters (the kernel's optional profile
buffer), mean over CTAs, as a share of one epilogue warp's whole loop.  Diagnostics for DESIGN / tuning."""
import json
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_lion_pytorch_b200.ops import _cabi, mlp  # noqa: E402

m, n, k = (int(x) for x in (sys.argv[1:4] if len(sys.argv) > 3 else (20480, 3072, 768)))
g = torch.Generator(device="cuda").manual_seed(0)
dy = torch.randn(m, k, device="cuda", generator=g).to(torch.bfloat16)
w = (torch.randn(n, k, device="cuda", generator=g) * k ** -0.5).to(torch.bfloat16)
pre = torch.randn(m, n, device="cuda", generator=g).to(torch.bfloat16)
for _ in range(3):
    mlp.dgrad_dgelu(dy, w, pre, want_bias_grad=False)
prof = torch.zeros(148, 16, dtype=torch.int64, device="cuda")
_cabi.lib().dlion_dgrad_dgelu_set_profile(prof.data_ptr())
mlp.dgrad_dgelu(dy, w, pre, want_bias_grad=False)
torch.cuda.synchronize()
_cabi.lib().dlion_dgrad_dgelu_set_profile(None)
p = prof.double().mean(0).tolist()
names = ["producer: empty stage", "mma: acc_empty", "mma: operands (full)", "store: box_done", "store: wait_group.read",
         "epilogue: acc_full", "epilogue: aux_full"]
total = p[8]
out = {"shape": [m, n, k], "epilogue_loop_cycles": total, "waits": {nm: {"cycles": round(p[i]), "share": round(p[i] / total, 3)}
                                                               for i, nm in enumerate(names)}}
print(json.dumps(out, indent=1))
