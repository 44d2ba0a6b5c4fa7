"""GPU tier: csrc/quant4.cu (NF4 / FP4 block codec) against the plain-torch oracle in ops/quant4.py,
and Linear4bit / LoRA on a 4-bit base against the same math done with expanded fp32 weights."""
import pytest
import torch

from distributed_lion_pytorch_b200.ops import quant4

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("quant_type", ["nf4", "fp4"])
@pytest.mark.parametrize("shape,blocksize", [((256, 192), 64), ((4096, 1024), 64), ((11008, 512), 128), ((3, 64), 64)])
def test_quantize_kernel_matches_oracle_bit_for_bit(dtype, quant_type, shape, blocksize):
    from distributed_lion_pytorch_b200.ops import _cabi
    torch.manual_seed(hash((shape, blocksize)) % 1000)
    w = (torch.randn(shape, device="cuda") * 0.02).to(dtype)
    w.view(-1)[:blocksize] = 0                                   # an all-zero block
    before = _cabi.LAUNCHES
    packed, absmax = quant4.quantize_4bit(w, blocksize, quant_type)
    assert _cabi.LAUNCHES == before + 1                          # the CUDA kernel ran, not the torch path
    p_ref, a_ref = quant4.quantize_4bit_reference(w.cpu(), blocksize, quant_type)
    assert torch.equal(absmax.cpu(), a_ref)
    assert torch.equal(packed.cpu(), p_ref)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("quant_type", ["nf4", "fp4"])
@pytest.mark.parametrize("n,blocksize", [(64, 64), (8192 * 33 + 64, 64), (4096 * 4096, 64), (1 << 20, 256)])
def test_dequantize_kernel_matches_oracle_bit_for_bit(dtype, quant_type, n, blocksize):
    from distributed_lion_pytorch_b200.ops import _cabi
    g = torch.Generator(device="cuda").manual_seed(n % 9973)
    packed = torch.randint(0, 256, (n // 2,), dtype=torch.uint8, device="cuda", generator=g)
    absmax = torch.rand(n // blocksize, device="cuda", generator=g) * 0.1
    before = _cabi.LAUNCHES
    out = quant4.dequantize_4bit(packed, absmax, (n,), dtype, blocksize, quant_type)
    assert _cabi.LAUNCHES == before + 1
    ref = quant4.dequantize_4bit_reference(packed, absmax, (n,), dtype, blocksize, quant_type)
    assert torch.equal(out, ref)
    # the shared-workspace form writes the same values into a caller buffer and leaves the tail alone
    buf = torch.full((n + 128,), 7.0, dtype=dtype, device="cuda")
    view = quant4.dequantize_4bit(packed, absmax, (n,), dtype, blocksize, quant_type, out=buf)
    assert view.data_ptr() == buf.data_ptr() and torch.equal(view, ref)
    assert bool((buf[n:] == 7.0).all())


def test_codec_argument_checks_on_gpu():
    w = torch.zeros(128, device="cuda")
    with pytest.raises(ValueError):
        quant4.quantize_4bit(w, 96)
    packed, absmax = quant4.quantize_4bit(w, 64)
    with pytest.raises(ValueError):
        quant4.dequantize_4bit(packed, absmax.half(), (128,), torch.float32)
    with pytest.raises(ValueError):
        quant4.dequantize_4bit(packed, absmax, (128,), torch.float32, out=torch.zeros(64, device="cuda"))


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_linear4bit_matches_expanded_weight_math(dtype):
    from distributed_lion_pytorch_b200.models import Linear4bit
    torch.manual_seed(3)
    lin = torch.nn.Linear(512, 256, bias=True, device="cuda", dtype=dtype)
    q = Linear4bit.from_linear(lin)
    other = Linear4bit.from_linear(torch.nn.Linear(1024, 512, device="cuda", dtype=dtype))   # shares the workspace
    wq = q.dequantize().clone()
    x = torch.randn(4, 33, 512, device="cuda", dtype=dtype, requires_grad=True)
    y = q(x)
    _ = other(torch.randn(2, 1024, device="cuda", dtype=dtype))     # overwrites the workspace between fwd and bwd
    gy = torch.randn_like(y)
    y.backward(gy)
    x2 = x.detach().clone().requires_grad_()
    y2 = torch.nn.functional.linear(x2, wq, lin.bias)
    y2.backward(gy)
    assert torch.equal(y, y2) and torch.equal(x.grad, x2.grad)


def test_llama_lora_on_4bit_base_trains_on_gpu():
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200.models import (Linear4bit, LoraConfig, apply_lora, build_model, merge_and_unload,
                                                      quantize_base_)
    torch.manual_seed(0)
    m = build_model("llama-tiny", dtype=torch.bfloat16, device="cuda")
    quantize_base_(m)
    apply_lora(m, LoraConfig(r=8, lora_alpha=16, lora_dropout=0.0))
    opt = Lion([p for p in m.parameters() if p.requires_grad], lr=3e-3)
    ids = torch.randint(0, m.config.vocab_size, (4, 64), device="cuda")
    losses = []
    for _ in range(8):
        opt.zero_grad()
        loss = m(ids, labels=ids).loss
        loss.backward()
        opt.step()
        losses.append(float(loss))
    assert losses[-1] < losses[0]
    m.eval()
    with torch.no_grad():
        tuned = m(ids).logits.float()
    merge_and_unload(m)
    assert not any(isinstance(x, Linear4bit) for x in m.modules())
    with torch.no_grad():
        assert torch.allclose(m(ids).logits.float(), tuned, atol=0.1, rtol=0.05)
