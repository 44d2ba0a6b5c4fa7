"""4-bit base weights (NF4 / FP4) on CPU: codec properties, Linear4bit autograd, LoRA on a 4-bit
base, merge back to 16-bit, and the sft_llama2 ``--load_in_4bit`` path end to end.
reference: sft_llama2.py:141-154 (BitsAndBytesConfig nf4), dpo_llama2.py:137,151 (load_in_4bit)."""
import pytest
import torch

from distributed_lion_pytorch_b200.ops import quant4


@pytest.mark.parametrize("quant_type", ["nf4", "fp4"])
@pytest.mark.parametrize("blocksize", [64, 128])
def test_codec_layout_and_error_bound(quant_type, blocksize):
    torch.manual_seed(1)
    w = torch.randn(48, 256) * 0.02
    packed, absmax = quant4.quantize_4bit(w, blocksize, quant_type)
    assert packed.dtype == torch.uint8 and packed.numel() == w.numel() // 2
    assert absmax.dtype == torch.float32 and absmax.numel() == w.numel() // blocksize
    assert torch.equal(absmax, w.reshape(-1, blocksize).abs().amax(1))
    back = quant4.dequantize_4bit(packed, absmax, w.shape, torch.float32, blocksize, quant_type)
    code = torch.tensor(quant4.code_book(quant_type))
    # every value lands on the nearest code-book entry: the error is at most half the widest gap, scaled
    gap = code.sort().values.diff().max() / 2
    err = (back - w).reshape(-1, blocksize).abs()
    assert (err <= gap * absmax[:, None] * (1 + 1e-6)).all()
    # the element holding the block maximum is reproduced exactly (codes contain +1 and -1)
    blocks, bb = w.reshape(-1, blocksize), back.reshape(-1, blocksize)
    arg = blocks.abs().argmax(1, keepdim=True)
    assert torch.equal(blocks.gather(1, arg), bb.gather(1, arg))
    # re-quantising the expanded weight is a fixed point
    p2, a2 = quant4.quantize_4bit(back, blocksize, quant_type)
    assert torch.equal(a2, absmax)
    assert torch.equal(quant4.dequantize_4bit(p2, a2, w.shape, torch.float32, blocksize, quant_type), back)


def test_codec_nibble_order_and_zero_blocks():
    code = quant4.code_book("nf4")
    w = torch.zeros(128)
    w[0], w[1] = 1.0, -1.0            # block 0: element 0 -> code 15, element 1 -> code 0
    packed, absmax = quant4.quantize_4bit(w, 64, "nf4")
    assert packed[0].item() == (15 << 4) | 0            # element 2i in the high nibble
    assert packed[1].item() == (7 << 4) | 7             # zeros map to the 0.0 entry (index 7)
    assert absmax.tolist() == [1.0, 0.0]
    back = quant4.dequantize_4bit(packed, absmax, w.shape, torch.float32)
    assert torch.equal(back, w) and code[7] == 0.0
    with pytest.raises(ValueError):
        quant4.quantize_4bit(torch.zeros(100), 64)
    with pytest.raises(ValueError):
        quant4.quantize_4bit(torch.zeros(128), 64, "int4")


def test_linear4bit_forward_backward_and_module_casts():
    from distributed_lion_pytorch_b200.models import Linear4bit
    torch.manual_seed(2)
    lin = torch.nn.Linear(64, 32, bias=True)
    q = Linear4bit.from_linear(lin)
    wq = q.dequantize()
    assert wq.shape == (32, 64) and wq.dtype == torch.float32
    x = torch.randn(5, 3, 64, requires_grad=True)
    y = q(x)
    y.pow(2).sum().backward()
    x2 = x.detach().clone().requires_grad_()
    y2 = torch.nn.functional.linear(x2, wq, lin.bias)
    y2.pow(2).sum().backward()
    assert torch.allclose(y, y2, atol=1e-6) and torch.allclose(x.grad, x2.grad, atol=1e-5)
    assert not any(p.requires_grad for p in q.parameters())
    q.to(torch.bfloat16)                                   # module casts must not touch the fp32 block scales
    assert q.absmax.dtype == torch.float32 and q.weight_packed.dtype == torch.uint8
    back = q.to_linear()
    assert type(back) is torch.nn.Linear and torch.allclose(back.weight.float(), wq, atol=1e-2)


def test_lora_on_4bit_base_trains_and_merges_to_16bit(tmp_path):
    from distributed_lion import Lion
    from distributed_lion_pytorch_b200.models import (Linear4bit, LoraConfig, apply_lora, build_model, merge_and_unload,
                                                      quantize_base_, quantized_bytes)
    torch.manual_seed(0)
    m = build_model("llama-tiny").eval()
    full = sum(p.numel() * p.element_size() for n, p in m.named_parameters() if "proj" in n)
    ids = torch.randint(0, 512, (2, 16))
    with torch.no_grad():
        dense = m(ids).logits.clone()
    quantize_base_(m)
    assert isinstance(m.model.layers[0].mlp.down_proj, Linear4bit) and type(m.lm_head) is torch.nn.Linear
    assert quantized_bytes(m) < full / 6                   # fp32 -> 4.5 bits per weight
    with torch.no_grad():
        base = m(ids).logits.clone()
    assert (base - dense).abs().max() < 0.25 * dense.abs().max()     # same function up to quantisation noise
    apply_lora(m, LoraConfig(r=4, lora_alpha=8, lora_dropout=0.0))
    trainable = [n for n, p in m.named_parameters() if p.requires_grad]
    assert trainable and all("lora_" in n for n in trainable)
    opt = Lion([p for p in m.parameters() if p.requires_grad], lr=1e-2)
    m.train()
    first = None
    for _ in range(4):
        opt.zero_grad()
        loss = m(ids, labels=ids).loss
        loss.backward()
        opt.step()
        first = first if first is not None else loss.item()
    assert loss.item() < first
    m.eval()
    with torch.no_grad():
        tuned = m(ids).logits.clone()
    merge_and_unload(m)
    assert not any(isinstance(x, Linear4bit) for x in m.modules())
    with torch.no_grad():
        assert torch.allclose(m(ids).logits, tuned, atol=1e-4)
    m.save_pretrained(str(tmp_path / "merged"))            # plain HF layout again
    again = build_model(str(tmp_path / "merged"))
    with torch.no_grad():
        assert torch.allclose(again(ids).logits, tuned, atol=1e-4)


def test_sft_entry_point_with_4bit_base(tmp_path):
    import sft_llama2
    out = tmp_path / "sft4"
    metrics = sft_llama2.main([
        "--model_name", "llama-tiny", "--synthetic_data", "--seq_length", "32", "--max_steps", "3",
        "--per_device_train_batch_size", "2", "--learning_rate", "1e-3", "--lion", "--async_grad",
        "--load_in_4bit", "--logging_steps", "1", "--save_strategy", "no", "--output_dir", str(out),
        "--model_dtype", "float32"])
    assert metrics["train_loss"] > 0
    assert (out / "final_merged_checkpoint" / "model.safetensors").exists()
    with pytest.raises(ValueError):
        sft_llama2.main(["--model_name", "llama-tiny", "--synthetic_data", "--max_steps", "1", "--load_in_4bit",
                         "--use_lora", "False", "--output_dir", str(out)])


def test_dpo_entry_point_with_4bit_policy_and_reference(tmp_path):
    """reference: dpo_llama2.py:133-152 loads BOTH the policy and the frozen reference model in 4 bit."""
    import dpo_llama2
    out = tmp_path / "dpo4"
    dpo_llama2.main([
        "--model_name_or_path", "llama-tiny", "--synthetic_data", "--synthetic_samples", "16", "--max_prompt_length", "8",
        "--max_length", "24", "--max_steps", "2", "--per_device_train_batch_size", "2", "--learning_rate", "1e-3",
        "--lion", "--async_grad", "--load_in_4bit", "--logging_steps", "1", "--eval_steps", "100", "--save_steps", "100",
        "--model_dtype", "float32", "--output_dir", str(out)])
    assert (out / "final_checkpoint").is_dir()
