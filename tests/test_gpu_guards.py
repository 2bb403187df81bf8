"""Out-of-bounds writes, found without a sanitizer (compute-sanitizer is closed on this pool).

Every output / scratch tensor the op wrappers allocate while the block runs is carved out of a larger byte buffer with a
4 KiB guard zone of a known pattern on both sides; after forward + backward of each op at shapes that do NOT divide the
kernels' tiles (partial TMA boxes, masked rows, tail vectors) the zones must be untouched.  A kernel that stores one row
or one vector too many lands in a guard zone instead of in a neighbouring allocation where nothing would notice."""
import pytest
import torch

pytestmark = pytest.mark.gpu

GUARD = 4096
PATTERN = 0xA5


class GuardedAllocations:
    """Routes torch.empty / empty_like / zeros / zeros_like on CUDA through guarded buffers while active."""

    def __init__(self):
        self.bufs = []

    def _carve(self, shape, dtype, device):
        n = 1
        for d in shape:
            n *= int(d)
        nbytes = n * torch.empty((), dtype=dtype).element_size()
        body = (nbytes + 255) // 256 * 256
        raw = self._orig_empty(GUARD + body + GUARD, dtype=torch.uint8, device=device)
        raw[:GUARD].fill_(PATTERN)
        raw[GUARD + nbytes:].fill_(PATTERN)
        self.bufs.append((raw, nbytes))
        return raw[GUARD:GUARD + nbytes].view(dtype).view(*shape)

    @staticmethod
    def _shape(args):
        if len(args) == 1 and isinstance(args[0], (tuple, list, torch.Size)):
            return tuple(args[0])
        return tuple(args)

    def __enter__(self):
        self._orig_empty, self._orig_zeros = torch.empty, torch.zeros
        self._orig_empty_like, self._orig_zeros_like = torch.empty_like, torch.zeros_like

        def is_cuda(device):
            return device is not None and torch.device(device).type == "cuda"

        def empty(*args, dtype=None, device=None, **kw):
            if not is_cuda(device) or kw.get("pin_memory"):
                return self._orig_empty(*args, dtype=dtype, device=device, **kw)
            return self._carve(self._shape(args), dtype or torch.get_default_dtype(), device)

        def zeros(*args, dtype=None, device=None, **kw):
            if not is_cuda(device):
                return self._orig_zeros(*args, dtype=dtype, device=device, **kw)
            return self._carve(self._shape(args), dtype or torch.get_default_dtype(), device).zero_()

        def empty_like(t, dtype=None, **kw):
            if not t.is_cuda or kw:
                return self._orig_empty_like(t, dtype=dtype, **kw)
            return self._carve(tuple(t.shape), dtype or t.dtype, t.device)

        def zeros_like(t, dtype=None, **kw):
            if not t.is_cuda or kw:
                return self._orig_zeros_like(t, dtype=dtype, **kw)
            return self._carve(tuple(t.shape), dtype or t.dtype, t.device).zero_()

        torch.empty, torch.zeros, torch.empty_like, torch.zeros_like = empty, zeros, empty_like, zeros_like
        return self

    def __exit__(self, *exc):
        torch.empty, torch.zeros = self._orig_empty, self._orig_zeros
        torch.empty_like, torch.zeros_like = self._orig_empty_like, self._orig_zeros_like
        return False

    def assert_intact(self, what: str):
        torch.cuda.synchronize()
        assert self.bufs, f"{what}: no allocation was routed through the guards"
        for i, (raw, nbytes) in enumerate(self.bufs):
            lo, hi = raw[:GUARD], raw[GUARD + nbytes:]
            assert bool((lo == PATTERN).all()), f"{what}: allocation {i} ({nbytes} B) was written BEFORE its start"
            assert bool((hi == PATTERN).all()), f"{what}: allocation {i} ({nbytes} B) was written PAST its end"


def dev():
    return torch.device("cuda", 0)


def _run(what, fn):
    with GuardedAllocations() as g:
        fn()
        g.assert_intact(what)
    return len(g.bufs)


@pytest.mark.parametrize("M,K,N", [(333, 72, 200), (1000, 768, 2304), (4096, 320, 136), (130, 4104, 520)])
def test_dense_forward_backward_stay_inside_their_buffers(M, K, N):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(M, K, device=dev()).bfloat16().requires_grad_(True)
    w = (torch.randn(K, N, device=dev()) * 0.05).bfloat16().requires_grad_(True)
    b = torch.randn(N, device=dev()).bfloat16().requires_grad_(True)

    def fn():
        y = ops.linear(x, w, b, "relu")
        y.backward(torch.ones_like(y))
    assert _run("dense", fn) >= 3


def test_fused_mlp_stays_inside_its_buffers():
    from distriflow_b200 import ops
    torch.manual_seed(0)
    M, D, H = 2000, 320, 2560                                   # 2-CTA kernels with partial 256-row tiles, colsum epilogue
    x = torch.randn(M, D, device=dev()).bfloat16().requires_grad_(True)
    w1 = (torch.randn(D, H, device=dev()) * 0.05).bfloat16().requires_grad_(True)
    b1 = torch.randn(H, device=dev()).bfloat16().requires_grad_(True)
    w2 = (torch.randn(H, D, device=dev()) * 0.05).bfloat16().requires_grad_(True)
    b2 = torch.randn(D, device=dev()).bfloat16().requires_grad_(True)

    def fn():
        y = ops.mlp(x, w1, b1, w2, b2)
        y.backward(torch.ones_like(y))
    assert _run("mlp", fn) >= 5


@pytest.mark.parametrize("B,T,H", [(3, 1000, 5), (1, 64, 1), (2, 130, 3)])
def test_attention_stays_inside_its_buffers(B, T, H):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    qkv = torch.randn(B, T, 3 * H * 64, device=dev()).bfloat16().requires_grad_(True)

    def fn():
        out = ops.causal_self_attention(qkv, H)
        out.backward(torch.ones_like(out))
    assert _run("attention", fn) >= 3


@pytest.mark.parametrize("rows,D", [(333, 768), (1001, 40), (77, 2048)])
def test_layer_norm_stays_inside_its_buffers(rows, D):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(rows, D, device=dev()).bfloat16().requires_grad_(True)
    r = torch.randn(rows, D, device=dev()).bfloat16().requires_grad_(True)
    g = torch.randn(D, device=dev()).bfloat16().requires_grad_(True)
    b = torch.randn(D, device=dev()).bfloat16().requires_grad_(True)

    def fn():
        s, y = ops.add_layer_norm(x, r, g, b, 1e-5)
        torch.autograd.backward([s, y], [torch.ones_like(s), torch.ones_like(y)])
    assert _run("layer_norm", fn) >= 4


@pytest.mark.parametrize("rows,V", [(257, 1000), (64, 50304), (1000, 10)])
def test_softmax_cross_entropy_stays_inside_its_buffers(rows, V):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    logits = torch.randn(rows, V, device=dev()).bfloat16().requires_grad_(True)
    labels = torch.randint(0, V, (rows,), device=dev())

    def fn():
        ops.softmax_cross_entropy(logits, labels).backward()
    assert _run("softmax_ce", fn) >= 1


@pytest.mark.parametrize("cfg", [dict(N=3, H=14, W=14, C=64, K=128, k=3, s=1), dict(N=2, H=28, W=28, C=128, K=128, k=3, s=2),
                                 dict(N=5, H=9, W=9, C=16, K=32, k=3, s=1), dict(N=2, H=33, W=33, C=3, K=64, k=7, s=2)])
def test_conv_and_batch_norm_stay_inside_their_buffers(cfg):
    from distriflow_b200 import ops
    torch.manual_seed(0)
    x = torch.randn(cfg["N"], cfg["H"], cfg["W"], cfg["C"], device=dev()).bfloat16().requires_grad_(True)
    w = (torch.randn(cfg["k"], cfg["k"], cfg["C"], cfg["K"], device=dev()) * 0.05).bfloat16().requires_grad_(True)
    gam = torch.ones(cfg["K"], device=dev()).bfloat16().requires_grad_(True)
    bet = torch.zeros(cfg["K"], device=dev()).bfloat16().requires_grad_(True)
    rm, rv = torch.zeros(cfg["K"], device=dev()), torch.ones(cfg["K"], device=dev())

    def fn():
        y = ops.conv2d_nhwc(x, w, None, (cfg["s"], cfg["s"]), "same", None)
        z = ops.batch_norm_nhwc(y, gam, bet, rm, rv, True, 0.9, 1e-5, "relu")
        p = ops.maxpool2d_nhwc(z, (2, 2)) if z.shape[1] >= 2 else z
        p.backward(torch.ones_like(p))
    assert _run("conv", fn) >= 3
