"""tcgen05 causal attention (`ops/csrc/attention_sm100.cu`) against a plain PyTorch fp32 restatement: forward, log-sum-exp and
all three gradients, on the packed `[B,T,3,H,64]` layout, for sequence lengths that are a partial tile (64), ragged (1000)
and full (1024), with an odd number of (batch x head) pairs."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda", 0)


def _reference(qkv32, n_head):
    """fp32 math on the same (bf16-rounded) inputs: softmax(Q K^T / sqrt(D) + causal mask) V."""
    B, T, C3 = qkv32.shape
    D = C3 // (3 * n_head)
    q, k, v = qkv32.view(B, T, 3, n_head, D).permute(2, 0, 3, 1, 4)             # [B,H,T,D]
    s = (q @ k.transpose(-1, -2)) / math.sqrt(D)
    mask = torch.ones(T, T, dtype=torch.bool, device=qkv32.device).tril()
    s = s.masked_fill(~mask, float("-inf"))
    lse = torch.logsumexp(s, dim=-1)                                             # [B,H,T]
    o = torch.softmax(s, dim=-1) @ v                                             # [B,H,T,D]
    return o.transpose(1, 2).reshape(B, T, n_head * D), lse


def _rel(a, b):
    return float((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-6))


# long sequences: 2304 = 18 query tiles (one CTA per work item, 9 key-tile pairs), 4096 = 32 (up to 32 query tiles and 63
# sub-iterations per item: barrier phases, the operand pool and the dQ accumulation run long); 170 x 3 heads x 1024 tokens
# gives more work items (2040) than SMs, so the persistent grid draws tickets
@pytest.mark.parametrize("B,H,T", [(1, 1, 64), (3, 1, 128), (1, 3, 1000), (3, 3, 1024), (1, 5, 300), (1, 2, 2304), (1, 1, 4096),
                                   (170, 3, 1024)])
def test_flash_attention_forward_and_gradients_match_fp32_reference(B, H, T):
    from distriflow_b200.ops.attention import _FlashAttention
    D = 64
    g = torch.Generator(device="cuda").manual_seed(1000 * B + 10 * H + T)
    qkv = (torch.randn(B, T, 3 * H * D, device=dev(), generator=g) * 1.5).to(torch.bfloat16).requires_grad_(True)
    do = torch.randn(B, T, H * D, device=dev(), generator=g).to(torch.bfloat16)
    out = _FlashAttention.apply(qkv, H)
    out.backward(do)
    ref_in = qkv.detach().float().requires_grad_(True)
    ref_out, ref_lse = _reference(ref_in, H)
    ref_out.backward(do.float())
    torch.cuda.synchronize()
    assert torch.isfinite(out).all() and torch.isfinite(qkv.grad).all()
    assert _rel(out, ref_out) <= 2e-2, ("out", _rel(out, ref_out))
    dg, rg = qkv.grad.view(B, T, 3, H, D).float(), ref_in.grad.view(B, T, 3, H, D)
    for w, name in enumerate(("dq", "dk", "dv")):
        assert _rel(dg[:, :, w], rg[:, :, w]) <= 2e-2, (name, _rel(dg[:, :, w], rg[:, :, w]))


def test_flash_attention_lse_and_repeated_calls():
    """The log-sum-exp saved for backward is exact to fp32 rounding, and two calls in a row give identical results (the
    shared dQ accumulator is left at zero by every backward call)."""
    from distriflow_b200.ops import _lib
    from distriflow_b200.ops._lib import check, ptr, stream
    from distriflow_b200.ops.attention import _FlashAttention, _sigs
    L = _sigs()
    B, H, T, D = 2, 3, 333, 64
    qkv = torch.randn(B, T, 3 * H * D, device=dev()).to(torch.bfloat16)
    out = torch.empty(B, T, H * D, device=dev(), dtype=torch.bfloat16)
    lse = torch.empty(B, H, T, device=dev(), dtype=torch.float32)
    check(L.dfb_attn_fwd(ptr(qkv), ptr(out), ptr(lse), B, T, H, D, stream()), "fwd")
    _, ref_lse = _reference(qkv.float(), H)
    assert float((lse - ref_lse).abs().max()) < 2e-3
    grads = []
    for _ in range(2):
        x = qkv.clone().requires_grad_(True)
        _FlashAttention.apply(x, H).float().square().sum().backward()
        grads.append(x.grad.clone())
    assert torch.equal(grads[0][:, :, H * D:], grads[1][:, :, H * D:])              # dK, dV: deterministic
    assert _rel(grads[0], grads[1]) < 1e-2                                          # dQ: fp32 reduce-add order may differ
    assert _lib.LAUNCHES[0] > 0


def test_gpt2_block_uses_the_native_attention_kernel():
    from distriflow_b200.ops import attention
    assert attention.FLASH
    x = torch.randn(2, 128, 3 * 2 * 64, device=dev()).to(torch.bfloat16).requires_grad_(True)
    y = attention.causal_self_attention(x, 2)
    assert type(y.grad_fn).__name__ == "_FlashAttentionBackward"


def test_attention_backward_emits_the_qkv_bias_gradient():
    """`dbias` of dfb_attn_bwd = column sums of the packed dqkv it writes (bias gradient of the QKV projection, K4: bias-grad
    in the producer), and the Dense backward picks it up from the gradient tensor instead of running a column-sum pass."""
    import ctypes
    from distriflow_b200.ops._lib import check, ptr, stream
    from distriflow_b200.ops.attention import _bwd_scratch, _sigs
    from distriflow_b200.ops.dense import take_colsum
    L = _sigs()
    B, H, T, D = 2, 3, 200, 64
    qkv = torch.randn(B, T, 3 * H * D, device=dev()).to(torch.bfloat16)
    do = torch.randn(B, T, H * D, device=dev()).to(torch.bfloat16)
    out = torch.empty(B, T, H * D, device=dev(), dtype=torch.bfloat16)
    lse = torch.empty(B, H, T, device=dev(), dtype=torch.float32)
    check(L.dfb_attn_fwd(ptr(qkv), ptr(out), ptr(lse), B, T, H, D, stream()), "fwd")
    dqkv = torch.empty_like(qkv)
    dbias = torch.full((3 * H * D,), 7.0, device=dev())            # the entry point clears it itself
    scratch = _bwd_scratch(dev(), (B, T, H, D), int(L.dfb_attn_bwd_scratch_bytes(B, T, H, D)))
    for _ in range(2):
        check(L.dfb_attn_bwd(ptr(qkv), ptr(out), ptr(do), ptr(lse), ptr(dqkv), ptr(dbias), ptr(scratch),
                             ctypes.c_size_t(scratch.numel()), B, T, H, D, stream()), "bwd")
    ref = dqkv.float().sum((0, 1))
    assert float((dbias - ref).abs().max()) <= 1e-3 * float(ref.abs().max()) + 1e-3
    # through autograd the sums ride on the gradient tensor that reaches the producer of qkv (the QKV Dense layer's backward)
    from distriflow_b200.ops.attention import _FlashAttention
    seen = []

    class Probe(torch.autograd.Function):
        @staticmethod
        def forward(ctx, t):
            return t.view_as(t)

        @staticmethod
        def backward(ctx, g):
            seen.append(g)
            return g

    x = qkv.clone().requires_grad_(True)
    _FlashAttention.apply(Probe.apply(x), H).backward(do)
    cs = take_colsum(seen[0], 3 * H * D)
    assert cs is not None and float((cs - seen[0].float().sum((0, 1))).abs().max()) <= 1e-3 * float(ref.abs().max()) + 1e-3
