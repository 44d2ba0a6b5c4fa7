#!/usr/bin/env python
"""Driver for `ncu --set full` over every hand-written kernel at the shape the training path launches it with
(GPT-2 small micro-batch 20 x 1024, Llama-2-7B micro-batch 4 x 1024).  Each kernel is launched twice (the first launch
warms caches and the instruction cache; summaries use the second).  Run it plain first, then under ncu:

    python bench/profile_all_kernels.py &&
    ncu --set full --clock-control none --import-source on -k regex:dlion -o gpurun_out/prof_all python bench/profile_all_kernels.py
"""
import ctypes as C
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from distributed_lion_pytorch_b200.ops import _cabi, bucket_ops, lion_ops  # noqa: E402
from distributed_lion_pytorch_b200.ops import linear as lin  # noqa: E402
from distributed_lion_pytorch_b200.ops.activation import gelu_tanh, swiglu  # noqa: E402
from distributed_lion_pytorch_b200.ops.cross_entropy import linear_cross_entropy  # noqa: E402
from distributed_lion_pytorch_b200.ops.embedding import embed  # noqa: E402
from distributed_lion_pytorch_b200.ops.norm import add_layer_norm, add_rms_norm  # noqa: E402
from distributed_lion_pytorch_b200.ops.qkv import split_qkv  # noqa: E402
from distributed_lion_pytorch_b200.ops.rope import apply_rope, rope_tables  # noqa: E402
from distributed_lion_pytorch_b200.parallel.symm import EmulatedExchange  # noqa: E402

torch.manual_seed(0)
dev = torch.device("cuda")
bf = torch.bfloat16
L = _cabi.lib()


def twice(fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()


def rnd(*shape, dtype=bf, grad=False):
    t = torch.randn(*shape, device=dev).to(dtype)
    return t.requires_grad_() if grad else t


# ---- optimizer side, GPT-2-sized bucket (124 M bf16 elements)
n = 124_444_672 // 8192 * 8192
p, g, m = rnd(n), rnd(n), rnd(n) * 0.1
ss = torch.empty(1, dtype=torch.float32, device=dev)
twice(lambda: bucket_ops.sumsq(g, ss))
twice(lambda: lion_ops.local_step(p, g, m, 1e-4, 0.1, 0.9, 0.99, clip=(ss, 1.0), zero_g=False))
words = torch.empty(n // 8, dtype=torch.uint8, device=dev)
twice(lambda: lion_ops.sign_pack(m, g, 0.9, words, clip=(ss, 1.0)))
gathered = words.view(1, -1).repeat(8, 1).contiguous()
twice(lambda: lion_ops.vote_update(p, g, m, gathered, 8, None, 1e-4, 0.1, 0.99, clip=(ss, 1.0)))
twice(lambda: bucket_ops.clip_scale_(g, ss, 1.0))
srcs = [rnd(768, 3072), rnd(3072, 768), rnd(50257, 768), rnd(768), rnd(2304)]
dsts = [torch.empty_like(s) for s in srcs]
twice(lambda: bucket_ops.multi_tensor_gather(srcs, dsts))
del gathered, words, srcs, dsts
# the fused exchange kernel, 8 ranks emulated in one cooperative launch, clip + zero-fill folded in
W = 8
nf = (n // W) // lion_ops.fused_tile(bf) * lion_ops.fused_tile(bf)
ex = EmulatedExchange(W, nf, bf, dev)
P, G, M = rnd(W, nf), rnd(W, nf), torch.zeros(W, nf, device=dev, dtype=bf)
ssf = bucket_ops.sumsq(G.view(-1))
twice(lambda: ex.step(P, G.copy_(P), M, 1e-4, 0.1, 0.9, 0.99, clip=(ssf, 1.0), zero_g=True))
ex.check_status()
del p, g, m, P, G, M, ex

# ---- GPT-2 model-side kernels, rows = 20 * 1024
rows = 20 * 1024
ids = torch.randint(0, 50257, (20, 1024), device=dev)
wte = torch.nn.Parameter(rnd(50257, 768) * 0.02)
wpe = torch.nn.Parameter(rnd(1024, 768) * 0.02)
wte.grad, wpe.grad = torch.zeros_like(wte), torch.zeros_like(wpe)


def emb():
    h = embed(ids, wte, wpe)
    h.backward(torch.ones_like(h))


twice(emb)
lw, lb = torch.nn.Parameter(torch.ones(768, device=dev, dtype=bf)), torch.nn.Parameter(torch.zeros(768, device=dev, dtype=bf))
cb = torch.nn.Parameter(torch.zeros(768, device=dev, dtype=bf))
lw.grad, lb.grad, cb.grad = torch.zeros_like(lw), torch.zeros_like(lb), torch.zeros_like(cb)
x, res = rnd(rows, 768, grad=True), rnd(rows, 768, grad=True)


def ln():
    y, s = add_layer_norm(x, res, lw, lb, 1e-5, 0.1, True, x_bias=cb)
    (y.sum() + s.sum()).backward()


twice(ln)
pre, dy = rnd(rows, 3072), rnd(rows, 3072)
dpre = torch.empty_like(pre)
outb = torch.zeros(3072, device=dev, dtype=bf)
part = torch.empty(int(L.dlion_colsum_parts(rows, 3072)), 3072, dtype=torch.float32, device=dev)
twice(lambda: L.dlion_gelu_tanh_bwd_colsum(pre.data_ptr(), dy.data_ptr(), dpre.data_ptr(), rows, 3072, part.data_ptr(),
                                           lin._tickets(dev).data_ptr(), outb.data_ptr(), 1, _cabi.dtype_code(bf),
                                           _cabi.stream_ptr(dev)))
twice(lambda: gelu_tanh(pre))
twice(lambda: L.dlion_gelu_tanh_bwd(pre.data_ptr(), dy.data_ptr(), dpre.data_ptr(), pre.numel(), _cabi.dtype_code(bf),
                                    _cabi.stream_ptr(dev)))
twice(lambda: lin.column_sum(dy, bf, accumulate_into=outb))
qkv = rnd(20, 1024, 2304, grad=True)
qb = torch.nn.Parameter(torch.zeros(2304, device=dev, dtype=bf))


def qk():
    q, k, v = split_qkv(qkv, 12, qb)
    torch.autograd.backward([q, k, v], [torch.ones_like(q), torch.ones_like(k), torch.ones_like(v)])


twice(qk)
h = rnd(4096, 768, grad=True)
lab = torch.randint(0, 50257, (4096,), device=dev)
twice(lambda: linear_cross_entropy(h, wte, lab).backward())
del pre, dy, dpre, qkv, x, res, h

# ---- Llama-2-7B model-side kernels, rows = 4 * 1024, hidden 4096, vocab 32000
rows = 4096
rw = torch.nn.Parameter(torch.ones(4096, device=dev, dtype=bf))
rw.grad = torch.zeros_like(rw)
x, res = rnd(rows, 4096, grad=True), rnd(rows, 4096, grad=True)


def rms():
    y, s = add_rms_norm(x, res, rw, 1e-5)
    (y.sum() + s.sum()).backward()


twice(rms)
gate, up = rnd(rows, 11008, grad=True), rnd(rows, 11008, grad=True)
twice(lambda: swiglu(gate, up).sum().backward())
cos, sin = rope_tables(1024, 128, 10000.0, dev)
q, k = rnd(4, 1024, 32 * 128, grad=True), rnd(4, 1024, 32 * 128, grad=True)


def rp():
    a, b = apply_rope(q.clone(), k.clone(), cos, sin, 128)        # in place on fresh projection-like outputs
    (a.sum() + b.sum()).backward()


twice(rp)
hw = torch.nn.Parameter(rnd(32000, 4096) * 0.02)
h = rnd(4096, 4096, grad=True)
lab = torch.randint(0, 32000, (4096,), device=dev)
twice(lambda: linear_cross_entropy(h, hw, lab).backward())
from distributed_lion_pytorch_b200.ops import quant4  # noqa: E402

w = rnd(11008, 4096)
packed, absmax = quant4.quantize_4bit(w, quant_type="nf4")
twice(lambda: quant4.quantize_4bit(w, quant_type="nf4"))
twice(lambda: quant4.dequantize_4bit(packed, absmax, w.shape, bf, quant_type="nf4"))
print("profile driver done")
