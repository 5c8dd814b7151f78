"""Time the fused tcgen05 attention (fwd, fwd+bwd) on the GPT-2 medium shape against torch SDPA (flash), bf16."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.nn.functional as F

from distributeddataparallel_b200.ops import native_tf

b, heads, t = (int(v) for v in (sys.argv[1:4] if len(sys.argv) > 3 else (8, 16, 1024)))
c = heads * 64
qkv = torch.randn(b, t, 3 * c, device="cuda").to(torch.bfloat16).requires_grad_(True)
dout = torch.randn(b, t, c, device="cuda").to(torch.bfloat16)


def timeit(fn, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def ours_fwd():
    with torch.no_grad():
        return native_tf.causal_self_attention(qkv, heads)


def ours_fb():
    qkv.grad = None
    native_tf.causal_self_attention(qkv, heads).backward(dout)


def sdpa(x):
    q, k, v = (z.reshape(b, t, heads, 64).transpose(1, 2) for z in x.split(c, dim=2))
    return F.scaled_dot_product_attention(q, k, v, is_causal=True).transpose(1, 2).reshape(b, t, c)


def ref_fwd():
    with torch.no_grad():
        return sdpa(qkv)


def ref_fb():
    qkv.grad = None
    sdpa(qkv).backward(dout)


flops_f = 4.0 * b * heads * t * t * 64 / 2
for fused in (True, False):
    native_tf.set_fused_attention(fused)
    tf_, tfb = timeit(ours_fwd), timeit(ours_fb)
    print(f"ours fused={int(fused)}: fwd {tf_:8.1f} us ({flops_f / tf_ / 1e6:6.1f} TF)  fwd+bwd {tfb:8.1f} us ({3.5 * flops_f / tfb / 1e6:6.1f} TF)")
tf_, tfb = timeit(ref_fwd), timeit(ref_fb)
print(f"torch SDPA     : fwd {tf_:8.1f} us ({flops_f / tf_ / 1e6:6.1f} TF)  fwd+bwd {tfb:8.1f} us ({3.5 * flops_f / tfb / 1e6:6.1f} TF)")
