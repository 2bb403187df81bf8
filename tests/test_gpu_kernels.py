"""Numerics of every hand-written kernel against a plain PyTorch fp32 reference of the same op (B200 box)."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def dev():
    return torch.device("cuda", 0)


def rel_err(a, b):
    a, b = a.float(), b.float()
    return float((a - b).abs().max() / b.abs().max().clamp(min=1e-6))


# ----------------------------------------------------------------------------------------- tcgen05 GEMM
def _gemm_case(M, N, K, a_mn, b_mn, out_dtype=torch.bfloat16, bias=False, act=None, block_n=None, splits=None, seed=0):
    from distriflow_b200.ops.dense import _init_sigs, gemm
    _init_sigs()
    g = torch.Generator(device="cuda").manual_seed(seed)
    A = torch.randn((K, M) if a_mn else (M, K), device=dev(), generator=g).bfloat16()
    B = torch.randn((K, N) if b_mn else (N, K), device=dev(), generator=g).bfloat16()
    bvec = torch.randn(N, device=dev(), generator=g) if bias else None
    out = gemm(A, B, a_mn=a_mn, b_mn=b_mn, M=M, N=N, K=K, bias=bvec, act=act, out_dtype=out_dtype, block_n=block_n, splits=splits)
    Af = (A.t() if a_mn else A).float()
    Bf = (B.t() if b_mn else B).float()
    ref = Af @ Bf.t()
    if bias:
        ref = ref + bvec
    if act == "relu":
        ref = F.relu(ref)
    elif act == "gelu":
        ref = F.gelu(ref, approximate="tanh")
    torch.cuda.synchronize()
    return rel_err(out, ref)


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, True), (True, False)])
@pytest.mark.parametrize("block_n", [128, 256, 512])
def test_gemm_tcgen05_layouts(a_mn, b_mn, block_n):
    e = _gemm_case(256, 512, 256, a_mn, b_mn, block_n=block_n, splits=1)       # 512 = 2-CTA (cta_group::2) 256x256 tiles
    assert e < 1e-2, f"rel err {e}"


@pytest.mark.parametrize("shape", [(300, 264, 200), (1000, 776, 1032), (8192, 2304, 768), (768, 3072, 4096)])
def test_gemm_two_cta_tails_epilogues_splitk(shape):
    M, N, K = shape
    for a_mn, b_mn in [(False, True), (False, False), (True, True)]:
        if a_mn and M % 8:
            continue
        assert _gemm_case(M, N, K, a_mn, b_mn, block_n=512, splits=1) < 1e-2, (shape, a_mn, b_mn)
    assert _gemm_case(M, N, K, False, True, block_n=512, splits=1, bias=True, act="gelu") < 1e-2
    assert _gemm_case(M, N, K, False, False, block_n=512, splits=1, bias=True, out_dtype=torch.float32) < 5e-3
    if K >= 1024:
        assert _gemm_case(M, N, K, True, True, block_n=512, splits=4) < 1e-2


@pytest.mark.parametrize("shape", [(300, 264, 200), (128, 16, 64), (1000, 776, 1032), (4096, 768, 768), (2048, 3072, 768)])
def test_gemm_tcgen05_tails(shape):
    M, N, K = shape
    for a_mn, b_mn in [(False, True), (False, False), (True, True)]:
        if a_mn and M % 8:
            continue
        e = _gemm_case(M, N, K, a_mn, b_mn, splits=1)
        assert e < 1e-2, f"{shape} a_mn={a_mn} b_mn={b_mn}: rel err {e}"


def test_gemm_tcgen05_epilogues_and_fp32_out():
    assert _gemm_case(512, 512, 256, False, True, bias=True, act="relu") < 1e-2
    assert _gemm_case(512, 768, 256, False, True, bias=True, act="gelu") < 1e-2
    assert _gemm_case(384, 256, 512, True, True, out_dtype=torch.float32) < 5e-3
    assert _gemm_case(384, 256, 512, False, False, out_dtype=torch.float32, bias=True) < 5e-3


def test_gemm_tcgen05_split_k():
    # wgrad-like: small MN, long K; scratch workspace path
    assert _gemm_case(768, 768, 8192, True, True, splits=4) < 1e-2
    assert _gemm_case(256, 384, 4096, True, True, out_dtype=torch.float32, splits=8) < 5e-3
    assert _gemm_case(768, 3072, 16384, True, True) < 1e-2          # heuristic picks the split


def test_gemm_persistent_many_tiles_and_cta_cap():
    from distriflow_b200.ops import dense as lin
    assert _gemm_case(8192, 2304, 768, False, True) < 1e-2           # > 148 tiles: several per CTA, both TMEM stages
    lin.set_max_ctas(100)
    try:
        assert _gemm_case(4096, 3072, 768, False, True) < 1e-2
    finally:
        lin.set_max_ctas(0)


# ----------------------------------------------------------------------------------------- SIMT GEMM
def test_gemm_simt_reference_shapes():
    from distriflow_b200.ops.dense import _init_sigs, gemm
    _init_sigs()
    x = torch.randn(32, 784, device=dev())
    w = torch.randn(784, 10, device=dev())
    b = torch.randn(10, device=dev())
    y = gemm(x, w, a_mn=False, b_mn=True, M=32, N=10, K=784, bias=b, act="relu")
    assert rel_err(y, F.relu(x @ w + b)) < 1e-5
    dy = torch.randn(32, 10, device=dev())
    dx = gemm(dy, w, a_mn=False, b_mn=False, M=32, N=784, K=10)
    assert rel_err(dx, dy @ w.t()) < 1e-5
    dw = gemm(x, dy, a_mn=True, b_mn=True, M=784, N=10, K=32)
    assert rel_err(dw, x.t() @ dy) < 1e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("act", [None, "relu", "gelu"])
def test_linear_autograd(dtype, act):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    K, N = (784, 10) if dtype == torch.float32 else (256, 384)
    x = torch.randn(4, 24, K, device=dev(), dtype=dtype, requires_grad=True)
    w = (torch.randn(K, N, device=dev()) * 0.05).to(dtype).requires_grad_(True)
    b = torch.randn(N, device=dev()).to(dtype).requires_grad_(True)
    y = ops.linear(x, w, b, act)
    gy = torch.randn_like(y)
    y.backward(gy)
    xr, wr, br = (t.detach().float().requires_grad_(True) for t in (x, w, b))
    yr = xr @ wr + br
    if act == "relu":
        yr = F.relu(yr)
    elif act == "gelu":
        yr = F.gelu(yr, approximate="tanh")
    yr.backward(gy.float())
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    assert rel_err(y, yr) < tol
    assert rel_err(x.grad, xr.grad) < tol and rel_err(w.grad, wr.grad) < tol and rel_err(b.grad, br.grad) < tol
    assert w.grad.dtype == dtype and b.grad.dtype == dtype


def test_linear_transposed_weight():
    from distriflow_b200 import ops
    torch.manual_seed(1)
    x = torch.randn(512, 128, device=dev()).bfloat16().requires_grad_(True)
    w = (torch.randn(640, 128, device=dev()) * 0.05).bfloat16().requires_grad_(True)     # [out, in]
    y = ops.linear(x, w, None, None, transposed=True)
    gy = torch.randn_like(y)
    y.backward(gy)
    xr, wr = x.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    yr = xr @ wr.t()
    yr.backward(gy.float())
    assert rel_err(y, yr) < 2e-2 and rel_err(x.grad, xr.grad) < 2e-2 and rel_err(w.grad, wr.grad) < 2e-2


# ----------------------------------------------------------------------------------------- softmax-CE
@pytest.mark.parametrize("dtype,C", [(torch.bfloat16, 50304), (torch.float32, 10), (torch.float32, 1003), (torch.bfloat16, 512)])
def test_softmax_ce_index(dtype, C):
    from distriflow_b200.ops.softmax_ce import softmax_cross_entropy
    torch.manual_seed(0)
    rows = 64
    logits = (torch.randn(rows, C, device=dev()) * 3).to(dtype).requires_grad_(True)
    labels = torch.randint(0, C, (rows,), device=dev())
    loss = softmax_cross_entropy(logits, labels)
    loss.backward()
    lr = logits.detach().float().requires_grad_(True)
    ref = F.cross_entropy(lr, labels)
    ref.backward()
    assert abs(float(loss) - float(ref)) < 2e-3 * max(1.0, abs(float(ref)))
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert rel_err(logits.grad, lr.grad) < tol + 5e-3


def test_softmax_ce_inplace_unit_grad_and_int32():
    from distriflow_b200.ops.softmax_ce import softmax_cross_entropy
    torch.manual_seed(0)
    base = (torch.randn(32, 2048, device=dev()) * 2).bfloat16()
    labels = torch.randint(0, 2048, (32,), device=dev(), dtype=torch.int32)
    lr = base.float().requires_grad_(True)
    F.cross_entropy(lr, labels.long()).backward()
    x = (base.clone() * 1.0).requires_grad_(True)       # non-leaf so the in-place overwrite is legal to inspect
    x2 = x * 1.0
    loss = softmax_cross_entropy(x2, labels, inplace=True, unit_grad=True)
    loss.backward()
    assert rel_err(x.grad, lr.grad) < 1.5e-2


def test_softmax_ce_dense_and_reference_compat():
    from distriflow_b200.losses import _RefCompatCE
    from distriflow_b200.ops.softmax_ce import reference_compat_ce, softmax_cross_entropy
    torch.manual_seed(0)
    z = torch.randn(32, 10, device=dev(), requires_grad=True)
    y = F.one_hot(torch.randint(0, 10, (32,), device=dev()), 10).float()
    loss = softmax_cross_entropy(z, y)
    loss.backward()
    zr = z.detach().clone().requires_grad_(True)
    ref = (-(y * F.log_softmax(zr, -1)).sum(-1)).mean()
    ref.backward()
    assert abs(float(loss) - float(ref)) < 1e-5 and rel_err(z.grad, zr.grad) < 1e-5
    for variant in ("custom", "true"):
        p = torch.softmax(torch.randn(32, 10, device=dev()), -1).requires_grad_(True)
        l1 = reference_compat_ce(y, p, variant)
        l1.backward()
        pc = p.detach().cpu().requires_grad_(True)
        l2 = _RefCompatCE.apply(pc, y.cpu(), variant)
        l2.backward()
        assert abs(float(l1) - float(l2)) < 1e-5 and rel_err(p.grad.cpu(), pc.grad) < 1e-5


# ----------------------------------------------------------------------------------------- conv / pool / helpers
@pytest.mark.parametrize("cfg", [
    dict(N=2, H=28, W=28, C=1, K=32, k=3, s=1, pad="valid"),          # reference CNN layer 1 (model.json)
    dict(N=2, H=26, W=26, C=32, K=32, k=3, s=1, pad="valid"),         # reference CNN layer 2
    dict(N=2, H=16, W=16, C=16, K=24, k=3, s=2, pad="same"),
    dict(N=1, H=15, W=17, C=3, K=8, k=5, s=2, pad="same"),
    dict(N=2, H=8, W=8, C=16, K=32, k=1, s=1, pad="valid"),
])
def test_conv2d_fp32_fwd_bwd(cfg):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(cfg["N"], cfg["H"], cfg["W"], cfg["C"], device=dev(), requires_grad=True)
    w = (torch.randn(cfg["k"], cfg["k"], cfg["C"], cfg["K"], device=dev()) * 0.2).requires_grad_(True)
    b = torch.randn(cfg["K"], device=dev(), requires_grad=True)
    y = ops.conv2d_nhwc(x, w, b, (cfg["s"], cfg["s"]), cfg["pad"], "relu")
    gy = torch.randn_like(y)
    y.backward(gy)
    xc, wc, bc = (t.detach().cpu().requires_grad_(True) for t in (x, w, b))
    yr = ops.conv2d_nhwc(xc, wc, bc, (cfg["s"], cfg["s"]), cfg["pad"], "relu")
    yr.backward(gy.cpu())
    assert y.shape == yr.shape
    assert rel_err(y.cpu(), yr) < 1e-4
    assert rel_err(x.grad.cpu(), xc.grad) < 1e-4 and rel_err(w.grad.cpu(), wc.grad) < 1e-3 and rel_err(b.grad.cpu(), bc.grad) < 1e-4


def test_conv2d_bf16_simt_path():
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(2, 12, 12, 8, device=dev()).bfloat16().requires_grad_(True)
    w = (torch.randn(3, 3, 8, 24, device=dev()) * 0.2).bfloat16().requires_grad_(True)
    y = ops.conv2d_nhwc(x, w, None, (1, 1), "same", None)
    gy = torch.randn_like(y)
    y.backward(gy)
    xc, wc = x.detach().float().cpu().requires_grad_(True), w.detach().float().cpu().requires_grad_(True)
    yr = ops.conv2d_nhwc(xc, wc, None, (1, 1), "same", None)
    yr.backward(gy.float().cpu())
    assert rel_err(y.cpu(), yr) < 2e-2 and rel_err(x.grad.cpu(), xc.grad) < 2e-2 and rel_err(w.grad.cpu(), wc.grad) < 2e-2


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_maxpool(dtype):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(2, 24, 24, 32, device=dev()).to(dtype).requires_grad_(True)
    y = ops.maxpool2d_nhwc(x, (2, 2))
    gy = torch.randn_like(y)
    y.backward(gy)
    xc = x.detach().float().cpu().requires_grad_(True)
    yr = ops.maxpool2d_nhwc(xc, (2, 2))
    yr.backward(gy.float().cpu())
    assert torch.equal(y.float().cpu(), yr) and rel_err(x.grad.cpu(), xc.grad) < 1e-6
    # overlapping 3x3/2 'same' windows (ResNet stem)
    x2 = torch.randn(1, 9, 9, 8, device=dev()).to(dtype).requires_grad_(True)
    y2 = ops.maxpool2d_nhwc(x2, (3, 3), (2, 2), "same")
    y2.sum().backward()
    x2c = x2.detach().float().cpu().requires_grad_(True)
    y2r = ops.maxpool2d_nhwc(x2c, (3, 3), (2, 2), "same")
    y2r.sum().backward()
    assert torch.equal(y2.float().cpu(), y2r) and rel_err(x2.grad.cpu(), x2c.grad) < 1e-6


def test_colsum_act_accuracy():
    from distriflow_b200.ops import _lib
    from distriflow_b200.ops._lib import check, dt, ptr, stream
    from distriflow_b200.ops.dense import _init_sigs, act_op, colsum
    _init_sigs()
    x = torch.randn(1000, 776, device=dev()).bfloat16()
    assert rel_err(colsum(x), x.float().sum(0)) < 1e-3
    x32 = torch.randn(333, 10, device=dev())
    assert rel_err(colsum(x32), x32.sum(0)) < 1e-5
    a = torch.randn(4099, device=dev()).bfloat16()
    b = torch.randn(4099, device=dev()).bfloat16()
    assert rel_err(act_op(a, None, 0), F.gelu(a.float(), approximate="tanh")) < 1e-2
    br = b.float().requires_grad_(True)
    F.gelu(br, approximate="tanh").backward(a.float())
    assert rel_err(act_op(a, b, 1), br.grad) < 1e-2
    assert torch.equal(act_op(a, b, 2), torch.where(b > 0, a, torch.zeros_like(a)))
    L = _lib.lib()
    L.dfb_accuracy.restype = ctypes.c_int
    L.dfb_accuracy.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int, ctypes.c_void_p, ctypes.c_int, ctypes.c_int,
                               ctypes.c_long, ctypes.c_long, ctypes.c_int, ctypes.c_void_p]
    pred = torch.randn(257, 10, device=dev())
    lab = torch.randint(0, 10, (257,), device=dev())
    slot = torch.zeros(2, device=dev())
    check(L.dfb_accuracy(ptr(pred), ptr(lab), 0, ptr(slot), 257, 10, 10, 0, dt(pred), stream()), "acc")
    assert int(slot[0]) == int((pred.argmax(-1) == lab).sum()) and int(slot[1]) == 257
    onehot = F.one_hot(lab, 10).float()
    slot.zero_()
    check(L.dfb_accuracy(ptr(pred), ptr(onehot), 2, ptr(slot), 257, 10, 10, 10, dt(pred), stream()), "acc")
    assert int(slot[0]) == int((pred.argmax(-1) == lab).sum())


# ----------------------------------------------------------------------------------------- LayerNorm / attention layout
@pytest.mark.parametrize("dtype,D", [(torch.bfloat16, 768), (torch.float32, 128), (torch.bfloat16, 2048), (torch.float32, 36)])
@pytest.mark.parametrize("with_res", [False, True])
def test_add_layer_norm(dtype, D, with_res):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    rows = 333
    x = torch.randn(3, rows // 3, D, device=dev()).to(dtype).requires_grad_(True)
    r = torch.randn(3, rows // 3, D, device=dev()).to(dtype).requires_grad_(True) if with_res else None
    g = (1 + 0.1 * torch.randn(D, device=dev())).to(dtype).requires_grad_(True)
    b = (0.1 * torch.randn(D, device=dev())).to(dtype).requires_grad_(True)
    if with_res:
        from distriflow_b200.ops.dense import take_colsum
        branch_grads = []

        class Probe(torch.autograd.Function):        # sees the gradient object exactly as the branch's producer would
            @staticmethod
            def forward(ctx, t):
                return t.view_as(t)

            @staticmethod
            def backward(ctx, gg):
                branch_grads.append(gg)
                return gg

        s, y = ops.add_layer_norm(x, Probe.apply(r), g, b, 1e-5)
        gs, gy = torch.randn_like(s), torch.randn_like(y)
        torch.autograd.backward([s, y], [gs, gy])
    else:
        y = ops.layer_norm(x, g, b, 1e-5)
        gy = torch.randn_like(y)
        y.backward(gy)
    xr, gr, br = (t.detach().float().requires_grad_(True) for t in (x, g, b))
    if with_res:
        rr = r.detach().float().requires_grad_(True)
        sr = (xr + rr).to(dtype).float() if dtype == torch.bfloat16 else xr + rr
        sr = xr + rr + (sr - (xr + rr)).detach()          # same rounding of the stored residual stream, gradient of a plain add
        yr = F.layer_norm(sr, (D,), gr, br, 1e-5)
        torch.autograd.backward([sr, yr], [gs.float(), gy.float()])
        assert rel_err(s, sr) < 1e-2 and rel_err(r.grad, rr.grad) < 2e-2
        # the branch gradient carries its own column sums (the bias gradient of the layer that produced the branch)
        cs = take_colsum(branch_grads[0], D)
        assert cs is not None and rel_err(cs, r.grad.float().reshape(-1, D).sum(0)) < (1e-5 if dtype == torch.float32 else 2e-3)
    else:
        yr = F.layer_norm(xr, (D,), gr, br, 1e-5)
        yr.backward(gy.float())
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    assert rel_err(y, yr) < tol and rel_err(x.grad, xr.grad) < tol
    assert rel_err(g.grad, gr.grad) < tol and rel_err(b.grad, br.grad) < tol


@pytest.mark.parametrize("with_res", [False, True])
def test_layer_norm_backward_staged_equals_register_kernel(with_res):
    """The bulk-copy staged backward (ring of rows per warp, many ring wraps per warp at this size) does the arithmetic
    of the register-only kernel in the same order: every output is bitwise equal."""
    from distriflow_b200 import ops
    from distriflow_b200.ops.norm import _sigs
    L = _sigs()
    torch.manual_seed(1)
    rows, D = 20011, 768
    x0 = torch.randn(rows, D, device=dev()).bfloat16()
    r0 = torch.randn(rows, D, device=dev()).bfloat16()
    g0 = (1 + 0.1 * torch.randn(D, device=dev())).bfloat16()
    b0 = (0.1 * torch.randn(D, device=dev())).bfloat16()
    gs, gy = torch.randn(rows, D, device=dev()).bfloat16(), torch.randn(rows, D, device=dev()).bfloat16()
    res = []
    try:
        for staged in (1, 0):
            L.dfb_layernorm_bwd_staged(staged)
            x, r, g, b = (t.clone().requires_grad_(True) for t in (x0, r0, g0, b0))
            if with_res:
                s, y = ops.add_layer_norm(x, r, g, b, 1e-5)
                torch.autograd.backward([s, y], [gs, gy])
            else:
                ops.layer_norm(x, g, b, 1e-5).backward(gy)
            res.append((x.grad.clone(), g.grad.clone(), b.grad.clone()))
    finally:
        L.dfb_layernorm_bwd_staged(0)
    for a, c in zip(*res):
        assert torch.equal(a, c)
    xr, gr, br = (t.float().requires_grad_(True) for t in (x0, g0, b0))
    F.layer_norm(xr, (D,), gr, br, 1e-5).backward(gy.float())
    if not with_res:
        assert rel_err(res[0][0], xr.grad) < 2e-2 and rel_err(res[0][1], gr.grad) < 2e-2


def test_attention_layout_kernels_match_torch():
    from distriflow_b200.ops.attention import causal_self_attention
    torch.manual_seed(0)
    B, T, H, D = 2, 128, 4, 64
    qkv = torch.randn(B, T, 3 * H * D, device=dev()).bfloat16().requires_grad_(True)
    out = causal_self_attention(qkv, H)
    go = torch.randn_like(out)
    out.backward(go)
    ref_in = qkv.detach().clone().requires_grad_(True)
    q, k, v = ref_in.view(B, T, 3, H, D).permute(2, 0, 3, 1, 4)
    ref = F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(B, T, H * D)
    ref.backward(go)
    assert rel_err(out, ref) < 2e-2 and rel_err(qkv.grad, ref_in.grad) < 2e-2


# ----------------------------------------------------------------------------------------- tcgen05 implicit-GEMM conv
@pytest.mark.parametrize("cfg", [
    dict(N=4, H=16, W=16, C=64, K=64, k=3, s=1, pad="same"),           # ResNet-20 last stage / ResNet-50 stage 1 3x3
    dict(N=2, H=28, W=28, C=128, K=128, k=3, s=1, pad="same"),         # partial pixel tiles (tw=28, th=4 -> 112 rows)
    dict(N=3, H=14, W=14, C=256, K=256, k=3, s=1, pad="same"),
    dict(N=4, H=7, W=7, C=512, K=512, k=3, s=1, pad="same"),           # two images per tile
    dict(N=2, H=56, W=56, C=64, K=256, k=1, s=1, pad="same"),          # 1x1 expand
    dict(N=2, H=56, W=56, C=256, K=64, k=1, s=1, pad="same"),          # 1x1 reduce
    dict(N=2, H=28, W=28, C=128, K=128, k=3, s=2, pad="same"),         # strided 3x3 (element strides in the tensor map)
    dict(N=2, H=28, W=28, C=256, K=512, k=1, s=2, pad="same"),         # strided 1x1 projection shortcut
    dict(N=2, H=12, W=12, C=64, K=128, k=3, s=1, pad="valid"),
])
def test_conv2d_tcgen05_vs_fp32_reference(cfg):
    from distriflow_b200 import ops
    from distriflow_b200.ops import conv as convmod
    torch.manual_seed(0)
    x = torch.randn(cfg["N"], cfg["H"], cfg["W"], cfg["C"], device=dev()).bfloat16().requires_grad_(True)
    w = (torch.randn(cfg["k"], cfg["k"], cfg["C"], cfg["K"], device=dev()) * (1.0 / (cfg["k"] * cfg["C"] ** 0.5))).bfloat16().requires_grad_(True)
    b = torch.randn(cfg["K"], device=dev()).bfloat16().requires_grad_(True)
    geom_ok = convmod._sigs().dfb_conv2d_tc_supported(0, cfg["N"], cfg["H"], cfg["W"], cfg["C"], -(-cfg["H"] // cfg["s"]),
                                                      -(-cfg["W"] // cfg["s"]), cfg["K"], cfg["k"], cfg["k"], cfg["s"], cfg["s"])
    assert geom_ok or cfg["pad"] == "valid"
    y = ops.conv2d_nhwc(x, w, b, (cfg["s"], cfg["s"]), cfg["pad"], "relu")
    gy = torch.randn_like(y)
    y.backward(gy)
    xr, wr, br = (t.detach().float().cpu().requires_grad_(True) for t in (x, w, b))
    yr = ops.conv2d_nhwc(xr, wr, br, (cfg["s"], cfg["s"]), cfg["pad"], "relu")
    yr.backward(gy.float().cpu())
    assert y.shape == yr.shape
    assert rel_err(y.cpu(), yr) < 2e-2, "forward"
    assert rel_err(x.grad.cpu(), xr.grad) < 2e-2, "dgrad"
    assert rel_err(w.grad.cpu(), wr.grad) < 2e-2, "wgrad"
    assert rel_err(b.grad.cpu(), br.grad) < 2e-2


def test_resnet20_and_resnet50_block_step_on_gpu():
    from distriflow_b200.models.resnet import Bottleneck, resnet20, resnet_loss
    from distriflow_b200.parallel.engine import DataParallelEngine
    torch.manual_seed(0)
    m = resnet20().to(dev())
    eng = DataParallelEngine(m, resnet_loss, {"optimizer": "momentum", "learningRate": 0.05}, device=dev())
    x = torch.randn(32, 32, 32, 3, device=dev())
    y = torch.randint(0, 10, (32,), device=dev())
    l0 = float(eng.step(x, y))
    for _ in range(10):
        l1 = float(eng.step(x, y))
    assert l1 == l1 and l1 < l0, (l0, l1)
    blk = Bottleneck(256, 64, 1, "b").to(dev()).to(torch.bfloat16)        # ResNet-50 stage-1 block: all convs on tcgen05
    xb = torch.randn(4, 56, 56, 256, device=dev()).bfloat16().requires_grad_(True)
    out = blk(xb)
    out.float().square().mean().backward()
    assert torch.isfinite(out.float()).all() and torch.isfinite(xb.grad.float()).all()


def test_fused_mlp_gelu_in_epilogues():
    from distriflow_b200 import ops
    torch.manual_seed(0)
    D, H = 256, 1024
    x = torch.randn(3, 200, D, device=dev()).bfloat16().requires_grad_(True)
    w1 = (torch.randn(D, H, device=dev()) * 0.05).bfloat16().requires_grad_(True)
    b1 = (torch.randn(H, device=dev()) * 0.1).bfloat16().requires_grad_(True)
    w2 = (torch.randn(H, D, device=dev()) * 0.05).bfloat16().requires_grad_(True)
    b2 = (torch.randn(D, device=dev()) * 0.1).bfloat16().requires_grad_(True)
    from distriflow_b200.ops import dense
    dense.FUSE_GELU = True
    try:
        y = ops.mlp(x, w1, b1, w2, b2)
        gy = torch.randn_like(y)
        y.backward(gy)
    finally:
        dense.FUSE_GELU = False
    ref = [t.detach().float().requires_grad_(True) for t in (x, w1, b1, w2, b2)]
    yr = F.gelu(ref[0] @ ref[1] + ref[2], approximate="tanh") @ ref[3] + ref[4]
    yr.backward(gy.float())
    assert rel_err(y, yr) < 2e-2
    for got, r in zip((x, w1, b1, w2, b2), ref):
        assert rel_err(got.grad, r.grad) < 3e-2


@pytest.mark.parametrize("M,N,K", [(4096, 3072, 768), (2000, 2560, 320)])
def test_dgelu_gemm_emits_the_bias_gradient_from_its_epilogue(M, N, K):
    """K4: the gelu'-epilogue dgrad GEMM also produces the column sums of what it stores (fc1's bias gradient)."""
    from distriflow_b200.ops import dense
    torch.manual_seed(0)
    dy = torch.randn(M, K, device=dev()).bfloat16()
    w2 = (torch.randn(N, K, device=dev()) * 0.05).bfloat16()
    pre = torch.randn(M, N, device=dev()).bfloat16()
    assert dense.pick_tile(M, N, K) == (512, 1)
    out, cs = dense.gemm_dgelu_colsum(dy, w2, a_mn=False, b_mn=False, M=M, N=N, K=K, aux=pre)
    plain = dense.gemm_aux(dy, w2, a_mn=False, b_mn=False, M=M, N=N, K=K, aux=pre, aux_mode=2)
    assert cs is not None and torch.equal(out, plain)
    assert rel_err(cs, out.float().sum(0)) < 1e-5
    x = pre.float()
    u = 0.7978845608028654 * (x + 0.044715 * x ** 3)
    t = torch.tanh(u)
    dgelu = 0.5 * (1 + t) + 0.5 * x * (1 - t * t) * 0.7978845608028654 * (1 + 3 * 0.044715 * x * x)
    assert rel_err(out, (dy.float() @ w2.float().t()) * dgelu) < 2e-2


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_embedding_backward_scatters_into_a_filled_gradient_slice(dtype, monkeypatch):
    """The token embedding's gradient is added IN PLACE on top of what the slice already holds (the tied LM head's weight
    gradient), with repeated tokens; and a parameter without a sink takes the plain autograd path."""
    from distriflow_b200 import ops
    from distriflow_b200.ops import embed
    monkeypatch.setattr(embed, "DETERMINISTIC", False)
    torch.manual_seed(0)
    V, D, n = 300, 64, 4096                                   # every row is hit ~14 times
    table = torch.nn.Parameter(torch.randn(V, D, device=dev()).to(dtype))
    idx = torch.randint(0, V, (8, n // 8), device=dev())
    dy = torch.randn(8, n // 8, D, device=dev()).to(dtype)
    head = torch.randn(V, D, device=dev()).to(dtype)
    sink = head.clone().reshape(-1)
    calls = []
    table._dfb_grad_sink, table._dfb_grad_ready, table._dfb_sink_filled = sink, (lambda: calls.append(1)), True
    ops.embedding(idx, table).backward(dy)
    ref = head.float().index_add(0, idx.reshape(-1), dy.reshape(-1, D).float())
    assert calls == [1] and table.grad is None and table._dfb_sink_filled is False
    assert rel_err(sink.view(V, D), ref) < (2e-2 if dtype == torch.bfloat16 else 1e-5)
    sink.fill_(7.0)                                           # nobody wrote the slice first: it is cleared, not added to
    ops.embedding(idx, table).backward(dy)
    ref0 = torch.zeros(V, D, device=dev()).index_add(0, idx.reshape(-1), dy.reshape(-1, D).float())
    assert rel_err(sink.view(V, D), ref0) < (2e-2 if dtype == torch.bfloat16 else 1e-5)
    plain = torch.nn.Parameter(table.detach().clone())
    ops.embedding(idx, plain).backward(dy)
    assert rel_err(plain.grad, ref0) < (2e-2 if dtype == torch.bfloat16 else 1e-5)


def test_thin_channel_conv_stem_im2col_path():
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(4, 64, 64, 3, device=dev()).bfloat16()
    w = (torch.randn(7, 7, 3, 64, device=dev()) * 0.05).bfloat16().requires_grad_(True)
    y = ops.conv2d_nhwc(x, w, None, (2, 2), "same", None)
    gy = torch.randn_like(y)
    y.backward(gy)
    wr = w.detach().float().cpu().requires_grad_(True)
    yr = ops.conv2d_nhwc(x.float().cpu(), wr, None, (2, 2), "same", None)
    yr.backward(gy.float().cpu())
    assert y.shape == yr.shape and rel_err(y.cpu(), yr) < 2e-2 and rel_err(w.grad.cpu(), wr.grad) < 2e-2


@pytest.mark.parametrize("cfg", [dict(C=16, K=16, k=3, s=1), dict(C=16, K=32, k=3, s=2), dict(C=32, K=32, k=3, s=1),
                                 dict(C=32, K=64, k=1, s=2), dict(C=64, K=32, k=3, s=1)])
def test_small_channel_conv_patch_gemm_path(cfg):
    """ResNet-20 channel counts (16 / 32): patch matrix + tcgen05 GEMM for forward, filter and (stride 1) data gradient."""
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(8, 32, 32, cfg["C"], device=dev()).bfloat16().requires_grad_(True)
    w = (torch.randn(cfg["k"], cfg["k"], cfg["C"], cfg["K"], device=dev()) * 0.1).bfloat16().requires_grad_(True)
    y = ops.conv2d_nhwc(x, w, None, (cfg["s"], cfg["s"]), "same", None)
    gy = torch.randn_like(y)
    y.backward(gy)
    xr, wr = x.detach().float().cpu().requires_grad_(True), w.detach().float().cpu().requires_grad_(True)
    yr = ops.conv2d_nhwc(xr, wr, None, (cfg["s"], cfg["s"]), "same", None)
    yr.backward(gy.float().cpu())
    assert y.shape == yr.shape and rel_err(y.cpu(), yr) < 2e-2
    assert rel_err(x.grad.cpu(), xr.grad) < 2e-2 and rel_err(w.grad.cpu(), wr.grad) < 2e-2


# ----------------------------------------------------------------------------------------- fused BatchNorm (+ residual + ReLU)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("shape", [(8, 14, 14, 64), (4, 7, 7, 2048), (16, 8, 8, 16), (3, 5, 5, 24)])
@pytest.mark.parametrize("with_res,relu", [(False, False), (False, True), (True, True), (True, False)])
def test_batch_norm_nhwc_fused(dtype, shape, with_res, relu):
    """`csrc/batchnorm.cu` against PyTorch's batch_norm (+ add + relu) evaluated in fp32 on the same (rounded) inputs."""
    from distriflow_b200 import ops
    g = torch.Generator(device="cuda").manual_seed(3)
    C = shape[-1]
    x = (torch.randn(shape, device=dev(), generator=g) * 2 + 0.5).to(dtype).requires_grad_(True)
    res = torch.randn(shape, device=dev(), generator=g).to(dtype).requires_grad_(True) if with_res else None
    gamma = (torch.rand(C, device=dev(), generator=g) + 0.5).to(dtype).requires_grad_(True)
    beta = (torch.randn(C, device=dev(), generator=g) * 0.1).to(dtype).requires_grad_(True)
    rm, rv = torch.zeros(C, device=dev()), torch.ones(C, device=dev())
    act = "relu" if relu else None
    y = ops.batch_norm_nhwc(x, gamma, beta, rm, rv, True, 0.1, 1e-5, act, res)
    dy = torch.randn(shape, device=dev(), generator=g).to(dtype)
    grads = torch.autograd.grad(y, [t for t in (x, res, gamma, beta) if t is not None], dy)
    # fp32 reference on the same inputs
    xf, gf, bf = (t.detach().float().requires_grad_(True) for t in (x, gamma, beta))
    rf = res.detach().float().requires_grad_(True) if with_res else None
    rm2, rv2 = torch.zeros(C, device=dev()), torch.ones(C, device=dev())
    yr = F.batch_norm(xf.permute(0, 3, 1, 2), rm2, rv2, gf, bf, True, 0.1, 1e-5).permute(0, 2, 3, 1)
    if with_res:
        yr = yr + rf
    if relu:
        yr = F.relu(yr)
    if relu and dtype == torch.bfloat16:      # the kernel's ReLU mask is y > 0 on the bf16 output: use the same mask
        pass
    gref = torch.autograd.grad(yr, [t for t in (xf, rf, gf, bf) if t is not None], dy.float())
    tol = 2e-2 if dtype == torch.bfloat16 else 2e-4
    assert rel_err(y, yr) < tol
    for a, b in zip(grads, gref):
        assert rel_err(a, b) < (4e-2 if dtype == torch.bfloat16 else 1e-3), (a.shape, rel_err(a, b))
    assert rel_err(rm, rm2) < 1e-3 and rel_err(rv, rv2) < 1e-3
    # inference mode uses the running statistics
    with torch.no_grad():
        ye = ops.batch_norm_nhwc(x.detach(), gamma.detach(), beta.detach(), rm, rv, False, 0.1, 1e-5, act, res.detach() if with_res else None)
        yer = F.batch_norm(xf.detach().permute(0, 3, 1, 2), rm2, rv2, gf.detach(), bf.detach(), False, 0.1, 1e-5).permute(0, 2, 3, 1)
        if with_res:
            yer = yer + rf.detach()
        if relu:
            yer = F.relu(yer)
    assert rel_err(ye, yer) < tol
