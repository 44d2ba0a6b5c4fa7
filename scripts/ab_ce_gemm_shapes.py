"""A/B for the chunked cross-entropy GEMMs at the GPT-2 shape: strided [c, V] views of a [c, Vp] block against
fully padded operands (V = 50257, Vp = 50304).  Prints CUDA-event times (us, median of 20)."""
import json
import torch

torch.manual_seed(0)
dev = "cuda"
c, H, V, Vp = 4096, 768, 50257, 50304
h = torch.randn(c, H, device=dev, dtype=torch.bfloat16)
w = torch.randn(V, H, device=dev, dtype=torch.bfloat16) * 0.02
wp = torch.zeros(Vp, H, device=dev, dtype=torch.bfloat16)
wp[:V] = w
buf = torch.empty(c, Vp, device=dev, dtype=torch.bfloat16)
view = buf[:, :V]
dh = torch.empty(c, H, device=dev, dtype=torch.bfloat16)
dw = torch.zeros(V, H, device=dev, dtype=torch.bfloat16)
dwp = torch.zeros(Vp, H, device=dev, dtype=torch.bfloat16)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def t(fn, n=20):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.fill_(1)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    return sorted(ts)[len(ts) // 2]


res = {
    "fwd_view": t(lambda: torch.mm(h, w.t(), out=view)),
    "fwd_padded": t(lambda: torch.mm(h, wp.t(), out=buf)),
    "dh_view": t(lambda: torch.mm(view, w, out=dh)),
    "dh_padded": t(lambda: torch.mm(buf, wp, out=dh)),
    "dw_view": t(lambda: dw.addmm_(view.t(), h)),
    "dw_padded": t(lambda: dwp.addmm_(buf.t(), h)),
    "pad_copy": t(lambda: wp[:V].copy_(w)),
}
print(json.dumps({k: round(v, 1) for k, v in res.items()}))
