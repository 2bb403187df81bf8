"""GPT-2 (BASELINE.json config 5: GPT-2 small, 124M, sequence 1024, bf16, synchronous DP).

Not in the reference (it ships only the MNIST models); needed as a named workload.  All projections are Dense
layers in the `[in, out]` layout (GPT-2's own Conv1D layout) running the tcgen05 GEMM, GELU and the bias are in
the GEMM path, the tied LM head runs the same kernel on the `[vocab, d]` embedding table (K-major B operand), the
loss is the fused softmax-CE kernel writing the logit gradient in place, causal attention (forward and backward) is the
tcgen05 kernel of `ops/csrc/attention_sm100.cu` on the packed QKV tensor, LayerNorm + residual adds are fused kernels; the
tied table's gradient is built in place in the gradient bucket (LM-head weight-gradient GEMM, then the token rows added by
`ops/embed.py`); only the forward embedding gather and the position add stay on PyTorch ops.
"""
from __future__ import annotations

import math
import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.nn as nn

from .. import ops
from . import layers as L


@dataclass
class GPT2Config:
    vocab_size: int = 50257
    n_positions: int = 1024
    n_embd: int = 768
    n_layer: int = 12
    n_head: int = 12
    padded_vocab: int = 50304          # multiple of 128: aligned TMA pitches for the LM head / CE kernel

    @staticmethod
    def small() -> "GPT2Config":
        return GPT2Config()

    @staticmethod
    def tiny() -> "GPT2Config":
        return GPT2Config(vocab_size=500, n_positions=128, n_embd=128, n_layer=2, n_head=2, padded_vocab=512)   # head dim 64: the tcgen05 attention kernel


class Block(nn.Module):
    def __init__(self, c: GPT2Config, i: int):
        super().__init__()
        d = c.n_embd
        self.n_head = c.n_head
        self.ln_1 = L.LayerNorm(d, name=f"h{i}/ln_1")
        self.c_attn = L.Dense(d, 3 * d, name=f"h{i}/attn/c_attn")
        self.c_proj = L.Dense(d, d, name=f"h{i}/attn/c_proj")
        self.ln_2 = L.LayerNorm(d, name=f"h{i}/ln_2")
        self.c_fc = L.Dense(d, 4 * d, activation="gelu", name=f"h{i}/mlp/c_fc")
        self.c_proj2 = L.Dense(4 * d, d, name=f"h{i}/mlp/c_proj")
        for lin, std in ((self.c_attn, 0.02), (self.c_fc, 0.02), (self.c_proj, 0.02 / math.sqrt(2 * c.n_layer)),
                         (self.c_proj2, 0.02 / math.sqrt(2 * c.n_layer))):
            nn.init.normal_(lin.kernel, std=std)

    def forward(self, x, h):
        """`x`: residual stream, `h` = ln_1(x) (computed by the caller, fused with the previous residual add).
        Returns the attention branch output and leaves the adds to the fused add+LayerNorm kernels."""
        a = ops.causal_self_attention(self.c_attn(h), self.n_head)
        x, h2 = ops.add_layer_norm(x, self.c_proj(a), self.ln_2.gamma, self.ln_2.beta, self.ln_2.epsilon)
        return x, ops.mlp(h2, self.c_fc.kernel, self.c_fc.bias, self.c_proj2.kernel, self.c_proj2.bias)


class GPT2(nn.Module):
    def __init__(self, config: Optional[GPT2Config] = None):
        super().__init__()
        c = self.config = config or GPT2Config.small()
        self.wte = nn.Parameter(torch.empty(c.padded_vocab, c.n_embd))
        self.wpe = nn.Parameter(torch.empty(c.n_positions, c.n_embd))
        nn.init.normal_(self.wte, std=0.02)
        nn.init.normal_(self.wpe, std=0.02)
        # used by the embedding gather AND the LM head.  Under an engine both gradients go straight into the bucket slice: the
        # head's weight-gradient GEMM writes it, the embedding backward adds its token rows on top (two ready() calls).
        # Without an engine (no sink) plain autograd accumulates the two dense gradients.
        if os.environ.get("DFB_EMBEDDING_SCATTER", "1") == "1":
            self.wte._dfb_grad_parts = 2
        else:                                # A/B: dense embedding gradient, accumulated by autograd, copied by the hook
            self.wte._dfb_shared = True
        self.h = nn.ModuleList([Block(c, i) for i in range(c.n_layer)])
        self.ln_f = L.LayerNorm(c.n_embd, name="ln_f")
        self.input_shape = [c.n_positions]
        self.output_shape = [c.n_positions, c.padded_vocab]

    def first_dense_kernel(self):
        """The first GEMM of the forward pass (block 0 QKV projection): its all-gather is fused into that GEMM."""
        return self.h[0].c_attn.kernel

    def num_params(self) -> int:
        c = self.config
        return sum(p.numel() for p in self.parameters()) - (c.padded_vocab - c.vocab_size) * c.n_embd

    def forward(self, idx: torch.Tensor, targets: Optional[torch.Tensor] = None):
        B, T = idx.shape
        pos = torch.arange(T, device=idx.device)
        x = ops.embedding(idx, self.wte) + self.wpe[pos]
        h = self.h[0].ln_1(x)
        for i, blk in enumerate(self.h):
            x, m = blk(x, h)
            nxt = self.h[i + 1].ln_1 if i + 1 < len(self.h) else self.ln_f
            x, h = ops.add_layer_norm(x, m, nxt.gamma, nxt.beta, nxt.epsilon)      # residual add fused with the next LayerNorm
        x = h
        logits = ops.linear(x, self.wte, None, None, transposed=True)
        if targets is None:
            return logits
        return ops.softmax_cross_entropy(logits, targets, inplace=True, unit_grad=True)


def gpt2_loss(model: GPT2, idx: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
    return model(idx, targets)
