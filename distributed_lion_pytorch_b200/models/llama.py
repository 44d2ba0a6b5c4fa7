"""Llama-2 family with HF parameter names (``model.layers.N.self_attn.q_proj.weight`` …).

reference: ``AutoModelForCausalLM.from_pretrained("meta-llama/Llama-2-7b-hf")`` in
sft_llama2.py:147 and dpo_llama2.py:133/147 (policy and frozen reference model).  The block
is built around the fused kernels: residual-add + RMSNorm (N3), in-place RoPE on the q/k
projection outputs (N4), chunked linear-cross-entropy / log-probs (N5).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.checkpoint import checkpoint

from ..ops.activation import swiglu
from ..ops.embedding import embed
from ..ops.cross_entropy import linear_cross_entropy, linear_log_probs
from ..ops.linear import fused_linear
from ..ops.norm import add_rms_norm
from ..ops.rope import apply_rope, rope_tables
from .common import CausalLMOutput, ConfigBase, PreTrainedMixin, shift_labels


@dataclass
class LlamaConfig(ConfigBase):
    vocab_size: int = 32000
    hidden_size: int = 4096
    intermediate_size: int = 11008
    num_hidden_layers: int = 32
    num_attention_heads: int = 32
    num_key_value_heads: int = 32
    max_position_embeddings: int = 4096
    rms_norm_eps: float = 1e-5
    rope_theta: float = 10000.0
    initializer_range: float = 0.02
    tie_word_embeddings: bool = False
    hidden_act: str = "silu"
    bos_token_id: int = 1
    eos_token_id: int = 2
    dtype_name: str = "float32"
    model_type = "llama"
    architectures = ("LlamaForCausalLM",)


LLAMA_PRESETS = {
    "llama-2-7b": dict(),
    "llama-2-13b": dict(hidden_size=5120, intermediate_size=13824, num_hidden_layers=40, num_attention_heads=40,
                        num_key_value_heads=40),
    "llama-1b": dict(hidden_size=2048, intermediate_size=5632, num_hidden_layers=16, num_attention_heads=16,
                     num_key_value_heads=16),
    "llama-tiny": dict(vocab_size=512, hidden_size=64, intermediate_size=176, num_hidden_layers=2,
                       num_attention_heads=4, num_key_value_heads=2, max_position_embeddings=128),
}


class Linear(nn.Linear):
    """nn.Linear (HF parameter names, LoRA-wrappable) whose backward accumulates the weight
    gradient inside the GEMM when a gradient buffer already exists (ops/linear.py)."""

    def forward(self, x):
        return fused_linear(x, self.weight, self.bias, in_out=False)


class RMSNormParams(nn.Module):
    def __init__(self, n: int, eps: float):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(n))
        self.eps = eps

    def forward(self, x, residual=None):
        return add_rms_norm(x, residual, self.weight, self.eps)


class LlamaAttention(nn.Module):
    def __init__(self, cfg: LlamaConfig):
        super().__init__()
        H = cfg.hidden_size
        self.n_head, self.n_kv = cfg.num_attention_heads, cfg.num_key_value_heads
        self.head_dim = H // self.n_head
        self.q_proj = Linear(H, self.n_head * self.head_dim, bias=False)
        self.k_proj = Linear(H, self.n_kv * self.head_dim, bias=False)
        self.v_proj = Linear(H, self.n_kv * self.head_dim, bias=False)
        self.o_proj = Linear(self.n_head * self.head_dim, H, bias=False)

    def forward(self, x, cos, sin, attn_mask=None):
        B, S, _ = x.shape
        q, k = apply_rope(self.q_proj(x), self.k_proj(x), cos, sin, self.head_dim)
        v = self.v_proj(x)
        q = q.view(B, S, self.n_head, self.head_dim).transpose(1, 2)
        k = k.view(B, S, self.n_kv, self.head_dim).transpose(1, 2)
        v = v.view(B, S, self.n_kv, self.head_dim).transpose(1, 2)
        y = F.scaled_dot_product_attention(q, k, v, attn_mask=attn_mask, is_causal=attn_mask is None,
                                           enable_gqa=self.n_kv != self.n_head)
        return self.o_proj(y.transpose(1, 2).reshape(B, S, -1))


class LlamaMLP(nn.Module):
    def __init__(self, cfg: LlamaConfig):
        super().__init__()
        self.gate_proj = Linear(cfg.hidden_size, cfg.intermediate_size, bias=False)
        self.up_proj = Linear(cfg.hidden_size, cfg.intermediate_size, bias=False)
        self.down_proj = Linear(cfg.intermediate_size, cfg.hidden_size, bias=False)

    def forward(self, x):
        return self.down_proj(swiglu(self.gate_proj(x), self.up_proj(x)))


class LlamaDecoderLayer(nn.Module):
    def __init__(self, cfg: LlamaConfig):
        super().__init__()
        self.self_attn = LlamaAttention(cfg)
        self.mlp = LlamaMLP(cfg)
        self.input_layernorm = RMSNormParams(cfg.hidden_size, cfg.rms_norm_eps)
        self.post_attention_layernorm = RMSNormParams(cfg.hidden_size, cfg.rms_norm_eps)

    def forward(self, hidden, residual, cos, sin, attn_mask=None):
        x, residual = self.input_layernorm(hidden, residual)
        a = self.self_attn(x, cos, sin, attn_mask)
        x, residual = self.post_attention_layernorm(a, residual)
        return self.mlp(x), residual


class LlamaModel(nn.Module):
    def __init__(self, cfg: LlamaConfig):
        super().__init__()
        self.embed_tokens = nn.Embedding(cfg.vocab_size, cfg.hidden_size)
        self.layers = nn.ModuleList(LlamaDecoderLayer(cfg) for _ in range(cfg.num_hidden_layers))
        self.norm = RMSNormParams(cfg.hidden_size, cfg.rms_norm_eps)


class LlamaForCausalLM(nn.Module, PreTrainedMixin):
    config_class = LlamaConfig
    tied_keys = ("lm_head.weight",)

    def __init__(self, config: LlamaConfig):
        super().__init__()
        self.config = config
        self.model = LlamaModel(config)
        self.lm_head = nn.Linear(config.hidden_size, config.vocab_size, bias=False)
        self.gradient_checkpointing = False
        self._rope = None
        self.reset_parameters()
        self.tie_weights()

    def tie_weights(self):
        if self.config.tie_word_embeddings:
            self.lm_head.weight = self.model.embed_tokens.weight

    @torch.no_grad()
    def reset_parameters(self):
        std = self.config.initializer_range
        for name, p in self.named_parameters():
            if name.endswith("norm.weight") or name.endswith("layernorm.weight"):
                p.fill_(1.0)
            else:
                p.normal_(0.0, std)

    def _tables(self, S: int, device):
        if self._rope is None or self._rope[0].shape[0] < S or self._rope[0].device != device:
            hd = self.config.hidden_size // self.config.num_attention_heads
            n = max(S, min(self.config.max_position_embeddings, 4096))
            self._rope = rope_tables(n, hd, self.config.rope_theta, device)
        return self._rope

    def hidden_states(self, input_ids, attention_mask=None):
        B, S = input_ids.shape
        m = self.model
        h = embed(input_ids, m.embed_tokens.weight)
        cos, sin = self._tables(S, h.device)
        mask = None
        if attention_mask is not None and not bool(attention_mask.all()):
            causal = torch.ones(S, S, dtype=torch.bool, device=h.device).tril()
            mask = causal[None, None] & attention_mask[:, None, None, :].bool()
        hidden, residual = h, None
        for layer in m.layers:
            if self.gradient_checkpointing and self.training:
                hidden, residual = checkpoint(layer, hidden, residual, cos, sin, mask, use_reentrant=False)
            else:
                hidden, residual = layer(hidden, residual, cos, sin, mask)
        out, _ = m.norm(hidden, residual)
        return out

    def forward(self, input_ids, labels=None, attention_mask=None, return_logits: Optional[bool] = None, **_):
        h = self.hidden_states(input_ids, attention_mask)
        loss = logits = None
        if labels is not None:
            loss = linear_cross_entropy(h.view(-1, h.shape[-1]), self.lm_head.weight, shift_labels(labels).view(-1))
        if return_logits or (labels is None and return_logits is None):
            logits = h @ self.lm_head.weight.t()
        return CausalLMOutput(loss=loss, logits=logits)

    def token_log_probs(self, input_ids, labels, attention_mask=None):
        h = self.hidden_states(input_ids, attention_mask)
        lp = linear_log_probs(h.view(-1, h.shape[-1]), self.lm_head.weight, shift_labels(labels).view(-1))
        return lp.view(labels.shape)

    @classmethod
    def from_preset(cls, name: str, **overrides):
        if name not in LLAMA_PRESETS:
            raise KeyError(f"unknown Llama preset {name}; have {sorted(LLAMA_PRESETS)}")
        return cls(LlamaConfig(**{**LLAMA_PRESETS[name], **overrides}))
