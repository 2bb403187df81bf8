"""Probe: does the library SDPA take the packed qkv [B,S,3,H,D] through strided views without copies, and what
layouts come back?  Also prints the device's multicast / VMM handle attributes (for the NVLS path)."""
import ctypes
import torch
import torch.nn.functional as F

dev = torch.device("cuda")
B, S, H, D = 32, 1024, 12, 64
qkv = torch.randn(B, S, 3, H, D, device=dev, dtype=torch.bfloat16, requires_grad=True)


def timed(fn, n=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(True), torch.cuda.Event(True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


q, k, v = (qkv[:, :, i].permute(0, 2, 1, 3) for i in range(3))
o = F.scaled_dot_product_attention(q, k, v, is_causal=True)
print("strided in: q.stride", q.stride(), "o.stride", o.stride(), "o.shape", tuple(o.shape))
do = torch.randn(B, S, H, D, device=dev, dtype=torch.bfloat16).permute(0, 2, 1, 3)
gq, gk, gv = torch.autograd.grad(o, (q, k, v), do, retain_graph=True)
print("grads strides", gq.stride(), gk.stride(), gv.stride())

qc, kc, vc = (t.detach().contiguous().requires_grad_() for t in (q, k, v))
doc = do.contiguous()


def fwd_s():
    return F.scaled_dot_product_attention(q, k, v, is_causal=True)


def fwd_c():
    return F.scaled_dot_product_attention(qc, kc, vc, is_causal=True)


oc = fwd_c()
print("fwd strided %.3f ms   contiguous %.3f ms" % (timed(fwd_s), timed(fwd_c)))
print("bwd strided %.3f ms   contiguous %.3f ms" % (
    timed(lambda: torch.autograd.grad(o, (q, k, v), do, retain_graph=True)),
    timed(lambda: torch.autograd.grad(oc, (qc, kc, vc), doc, retain_graph=True))))
with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CUDA]) as p:
    o2 = fwd_s()
    torch.autograd.grad(o2, (q, k, v), do)
    torch.cuda.synchronize()
for e in sorted(p.key_averages(), key=lambda e: -e.device_time_total)[:8]:
    print("  %-90s %3d  %.1f us" % (e.key[:90], e.count, e.device_time_total))

cu = ctypes.CDLL("libcuda.so.1")
for name, attr in (("MULTICAST_SUPPORTED", 132), ("HANDLE_TYPE_POSIX_FD", 103), ("HANDLE_TYPE_FABRIC", 128),
                   ("VMM_SUPPORTED", 102), ("GPU_DIRECT_RDMA_WITH_CUDA_VMM", 116)):
    v_ = ctypes.c_int(-1)
    rc = cu.cuDeviceGetAttribute(ctypes.byref(v_), attr, 0)
    print("attr", name, "rc", rc, "value", v_.value)
