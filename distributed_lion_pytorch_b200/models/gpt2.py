"""GPT-2 with HF parameter names (``transformer.h.N.attn.c_attn.weight`` …, Conv1D layout
``[in, out]``), so checkpoints interchange with ``transformers.GPT2LMHeadModel``.

reference: the reference trains ``AutoModelForCausalLM.from_config(GPT2Config())``
(run_clm.py:425-450) through HF eager code.  Here the block is laid out around the fused
kernels: residual-add + LayerNorm is one kernel (N3), the loss is the chunked
linear-cross-entropy (N5: the [B, S, 50257] logits are never built), GEMMs go to cuBLAS and
attention to SDPA/flash, as the north star prescribes.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn
import torch.nn.functional as F
from torch.utils.checkpoint import checkpoint

from ..ops.activation import gelu_tanh
from ..ops.cross_entropy import linear_cross_entropy, linear_log_probs
from ..ops.embedding import embed
from ..ops.linear import fused_linear
from ..ops.lt_gemm import linear_gelu
from ..ops.mlp import gelu_mlp
from ..ops.norm import add_layer_norm
from ..ops.qkv import split_qkv
from .common import CausalLMOutput, ConfigBase, PreTrainedMixin, shift_labels


@dataclass
class GPT2Config(ConfigBase):
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    n_inner: Optional[int] = None
    activation_function: str = "gelu_new"
    resid_pdrop: float = 0.1
    embd_pdrop: float = 0.1
    attn_pdrop: float = 0.1
    layer_norm_epsilon: float = 1e-5
    initializer_range: float = 0.02
    tie_word_embeddings: bool = True
    bos_token_id: int = 50256
    eos_token_id: int = 50256
    dtype_name: str = "float32"
    model_type = "gpt2"
    architectures = ("GPT2LMHeadModel",)

    @property
    def hidden_size(self):
        return self.n_embd


GPT2_PRESETS = {
    "gpt2": dict(),
    "gpt2-medium": dict(n_embd=1024, n_layer=24, n_head=16),
    "gpt2-large": dict(n_embd=1280, n_layer=36, n_head=20),
    "gpt2-xl": dict(n_embd=1600, n_layer=48, n_head=25),
    "gpt2-tiny": dict(n_embd=64, n_layer=2, n_head=4, vocab_size=512, n_positions=128),
}


class Conv1D(nn.Module):
    """HF's GPT-2 linear: ``y = x @ W + b`` with ``W`` stored ``[in, out]``."""

    def __init__(self, nf: int, nx: int):
        super().__init__()
        self.nf = nf
        self.weight = nn.Parameter(torch.empty(nx, nf))
        self.bias = nn.Parameter(torch.zeros(nf))

    def forward(self, x):
        return fused_linear(x, self.weight, self.bias, in_out=True)


class LayerNormParams(nn.Module):
    """Holds ``weight``/``bias`` under the HF names; the math runs in ``ops.norm``."""

    def __init__(self, n: int, eps: float):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(n))
        self.bias = nn.Parameter(torch.zeros(n))
        self.eps = eps

    def forward(self, x, residual=None, dropout_p: float = 0.0, x_bias=None):
        """LayerNorm(dropout(x) + residual); the dropout of the branch output is fused in.  ``x_bias``: the bias of
        the linear layer that produced ``x`` — its gradient falls out of this norm's backward pass."""
        return add_layer_norm(x, residual, self.weight, self.bias, self.eps, dropout_p, self.training, x_bias)


class GPT2Attention(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.n_head, self.head_dim = cfg.n_head, cfg.n_embd // cfg.n_head
        self.c_attn = Conv1D(3 * cfg.n_embd, cfg.n_embd)
        self.c_proj = Conv1D(cfg.n_embd, cfg.n_embd)
        self.attn_pdrop = cfg.attn_pdrop

    def forward(self, x, attn_mask=None):
        B, S, E = x.shape
        # views in forward; in backward ONE kernel packs dq, dk, dv into d(qkv) and reduces c_attn's bias gradient
        q, k, v = split_qkv(self.c_attn(x), self.n_head, self.c_attn.bias)
        y = F.scaled_dot_product_attention(q, k, v, attn_mask=attn_mask, is_causal=attn_mask is None,
                                           dropout_p=self.attn_pdrop if self.training else 0.0)
        return self.c_proj(y.transpose(1, 2).reshape(B, S, E))


class GPT2MLP(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        inner = cfg.n_inner or 4 * cfg.n_embd
        self.c_fc = Conv1D(inner, cfg.n_embd)
        self.c_proj = Conv1D(cfg.n_embd, inner)
        self.act = cfg.activation_function

    def forward(self, x):
        if self.act in ("gelu_new", "gelu_pytorch_tanh"):
            # one autograd node: GELU in the c_fc GEMM's epilogue forward; backward = tcgen05 dgrad GEMM with the GELU
            # backward and c_fc's bias gradient in ITS epilogue (ops/mlp.py, csrc/gemm_dgelu_sm100.cu)
            y = gelu_mlp(x, self.c_fc.weight, self.c_fc.bias, self.c_proj.weight, self.c_proj.bias)
            if y is not None:
                return y
            h = linear_gelu(x, self.c_fc.weight, self.c_fc.bias)        # GEMM with the GELU in its epilogue
            if h is None:
                h = gelu_tanh(self.c_fc(x))
        else:
            h = F.gelu(self.c_fc(x))
        return self.c_proj(h)


class GPT2Block(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.ln_1 = LayerNormParams(cfg.n_embd, cfg.layer_norm_epsilon)
        self.attn = GPT2Attention(cfg)
        self.ln_2 = LayerNormParams(cfg.n_embd, cfg.layer_norm_epsilon)
        self.mlp = GPT2MLP(cfg)
        self.resid_pdrop = cfg.resid_pdrop

    def forward(self, hidden, residual, attn_mask=None, in_dropout: float = 0.0, in_bias=None):
        """(branch output, residual stream) in → same out.  The residual add AND the dropout of the
        incoming branch output live inside the norm kernel; this block's own MLP output is returned
        un-dropped and is dropped by whoever normalises it next (the next block's ln_1, or ln_f).
        ``in_bias``: the bias of the layer that produced ``hidden`` (the previous block's ``mlp.c_proj``): the
        norm backward reduces its gradient on the way."""
        x, residual = self.ln_1(hidden, residual, in_dropout, in_bias)
        a = self.attn(x, attn_mask)
        x, residual = self.ln_2(a, residual, self.resid_pdrop, self.attn.c_proj.bias)
        return self.mlp(x), residual


class GPT2Model(nn.Module):
    def __init__(self, cfg: GPT2Config):
        super().__init__()
        self.wte = nn.Embedding(cfg.vocab_size, cfg.n_embd)
        self.wpe = nn.Embedding(cfg.n_positions, cfg.n_embd)
        self.h = nn.ModuleList(GPT2Block(cfg) for _ in range(cfg.n_layer))
        self.ln_f = LayerNormParams(cfg.n_embd, cfg.layer_norm_epsilon)
        self.embd_pdrop = cfg.embd_pdrop


class GPT2LMHeadModel(nn.Module, PreTrainedMixin):
    config_class = GPT2Config
    tied_keys = ("lm_head.weight",)

    def __init__(self, config: GPT2Config):
        super().__init__()
        self.config = config
        self.transformer = GPT2Model(config)
        self.lm_head = nn.Linear(config.n_embd, config.vocab_size, bias=False)
        self.gradient_checkpointing = False
        self.reset_parameters()
        self.tie_weights()

    def tie_weights(self):
        if self.config.tie_word_embeddings:
            self.lm_head.weight = self.transformer.wte.weight

    @torch.no_grad()
    def reset_parameters(self):
        std = self.config.initializer_range
        for name, p in self.named_parameters():
            if name.endswith("bias"):
                p.zero_()
            elif ".ln_" in name or name.endswith("ln_f.weight"):
                p.fill_(1.0)
            elif name.endswith("c_proj.weight"):
                p.normal_(0.0, std / math.sqrt(2 * self.config.n_layer))
            else:
                p.normal_(0.0, std)

    def hidden_states(self, input_ids, attention_mask=None):
        B, S = input_ids.shape
        t = self.transformer
        h = embed(input_ids, t.wte.weight, t.wpe.weight)      # wte[ids] + wpe[0..S-1], one pass
        mask = None
        if attention_mask is not None and not bool(attention_mask.all()):
            causal = torch.ones(S, S, dtype=torch.bool, device=h.device).tril()
            mask = causal[None, None] & attention_mask[:, None, None, :].bool()
        hidden, residual = h, None
        p_in = t.embd_pdrop                      # layer 0 drops the embedding sum, later layers the previous MLP output
        b_in = None                              # bias of the linear that produced `hidden` (none for the embeddings)
        for blk in t.h:
            if self.gradient_checkpointing and self.training:
                hidden, residual = checkpoint(blk, hidden, residual, mask, p_in, b_in, use_reentrant=False)
            else:
                hidden, residual = blk(hidden, residual, mask, p_in, b_in)
            p_in = blk.resid_pdrop
            b_in = blk.mlp.c_proj.bias
        out, _ = t.ln_f(hidden, residual, p_in, b_in)
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
        """fp32 [B, S] log p(labels[t+1] | ≤t); 0 where the target is ignored (DPO building block)."""
        h = self.hidden_states(input_ids, attention_mask)
        lp = linear_log_probs(h.view(-1, h.shape[-1]), self.lm_head.weight, shift_labels(labels).view(-1))
        return lp.view(labels.shape)

    @classmethod
    def from_preset(cls, name: str, **overrides):
        if name not in GPT2_PRESETS:
            raise KeyError(f"unknown GPT-2 preset {name}; have {sorted(GPT2_PRESETS)}")
        return cls(GPT2Config(**{**GPT2_PRESETS[name], **overrides}))
