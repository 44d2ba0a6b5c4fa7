"""Which Python call sites launch the small ATen kernels (fill, copy, cat, add) of a GPT-2 micro-batch?
torch.profiler with stacks over one steady-state micro-batch; prints, per (kernel family, innermost repo frame),
the launch count and summed device time.  Usage: python scripts/attribute_aten_kernels.py [out.json]"""
import collections
import json
import re
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
from distributed_lion import Lion  # noqa: E402
from distributed_lion_pytorch_b200.models import build_model  # noqa: E402

dev = torch.device("cuda", 0)
torch.manual_seed(0)
model = build_model("gpt2", dtype=torch.bfloat16, device=dev, pretrained=False)
opt = Lion(model.parameters(), lr=1e-4, weight_decay=0.1, zero_grads_in_step=True)
ids = torch.randint(0, 50257, (20, 1024), device=dev)


def micro():
    loss = model(ids, labels=ids).loss
    (loss / 8).backward()


for _ in range(2):
    micro()
    opt.clip_grad_norm_(1.0)
    opt.step()
    opt.zero_grad()
micro()
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

with profile(activities=[ProfilerActivity.CPU, ProfilerActivity.CUDA], with_stack=True) as prof:
    micro()
    torch.cuda.synchronize()

fam = [("fill", r"FillFunctor"), ("copy", r"copy_kernel|direct_copy|CatArray|cat"), ("add/mul", r"CUDAFunctor_add|MulFunctor|BinaryFunctor|AddFunctor"),
       ("reduce", r"reduce_kernel"), ("index/embedding", r"index|embedding|gather|scatter"), ("other-aten", r"at::native|native::")]
agg = collections.defaultdict(lambda: [0, 0.0])
total = collections.defaultdict(lambda: [0, 0.0])
events = prof.events()
cpu_by_id = {}
for e in events:
    if e.device_type == torch.autograd.DeviceType.CPU:
        for k in e.kernels:
            cpu_by_id.setdefault(id(k), e)
for e in events:
    if e.device_type != torch.autograd.DeviceType.CPU or not e.kernels:
        continue
    stack = [s for s in (e.stack or []) if "distributed_lion" in s or "/repo/" in s]
    site = stack[0].strip() if stack else "(no repo frame: autograd engine / library)"
    site = re.sub(r"^.*/repo/", "", site)
    for k in e.kernels:
        name = k.name
        f = next((n for n, pat in fam if re.search(pat, name)), None)
        if f is None:
            continue
        agg[(f, e.name, site)][0] += 1
        agg[(f, e.name, site)][1] += k.duration
        total[f][0] += 1
        total[f][1] += k.duration
rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
print("families:", {k: (v[0], round(v[1], 1)) for k, v in total.items()})
for (f, op, site), (c, us) in rows[:45]:
    print(f"{us:9.1f} us x{c:4d}  [{f}] {op}  <- {site}")
if len(sys.argv) > 1:
    Path(sys.argv[1]).write_text(json.dumps([{"family": f, "op": op, "site": site, "count": c, "us": us}
                                             for (f, op, site), (c, us) in rows], indent=1))
