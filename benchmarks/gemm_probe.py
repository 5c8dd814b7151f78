"""Time the tcgen05 GEMM entry point on the GPT-2 medium shapes against torch.matmul (cuBLAS), bf16.

    python benchmarks/gemm_probe.py [--one M,N,K,mode] [--iters 20]

mode 0: a[M,K] b[N,K]^T   1: a[M,K] b[K,N]   2: a[K,M]^T b[K,N]     (all row-major, bf16 output)
Each timed iteration works on a fresh operand set out of a pool larger than L2 (126 MB).
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from distributeddataparallel_b200.ops import native_tf

SHAPES = [  # (name, M, N, K, mode)
    ("qkv fwd", 8192, 3072, 1024, 0), ("proj fwd", 8192, 1024, 1024, 0), ("fc fwd", 8192, 4096, 1024, 0),
    ("fc2 fwd", 8192, 1024, 4096, 0), ("lm head", 8192, 50257, 1024, 0),
    ("qkv dgrad", 8192, 1024, 3072, 1), ("fc dgrad", 8192, 1024, 4096, 1), ("fc2 dgrad", 8192, 4096, 1024, 1),
    ("qkv wgrad", 3072, 1024, 8192, 2), ("proj wgrad", 1024, 1024, 8192, 2), ("fc wgrad", 4096, 1024, 8192, 2),
    ("fc2 wgrad", 1024, 4096, 8192, 2), ("square 8k", 8192, 8192, 8192, 0),
]


def operands(m, n, k, mode, copies):
    mk = lambda *s: [torch.randn(*s, device="cuda", dtype=torch.bfloat16) * 0.1 for _ in range(copies)]
    if mode == 0:
        return mk(m, k), mk(n, k)
    if mode == 1:
        return mk(m, k), mk(k, n)
    return mk(k, m), mk(k, n)


def ref(a, b, mode):
    if mode == 0:
        return a @ b.t()
    if mode == 1:
        return a @ b
    return a.t() @ b


def time_fn(fn, copies, iters):
    for i in range(3):
        fn(i % copies)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fn(i % copies)
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3      # us


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--one", default=None)
    ap.add_argument("--iters", type=int, default=20)
    ap.add_argument("--no-ref", action="store_true")
    ap.add_argument("--epilogues", action="store_true", help="also time bias / residual / GELU epilogue variants")
    args = ap.parse_args()
    shapes = SHAPES
    if args.one:
        m, n, k, mode = (int(v) for v in args.one.split(","))
        shapes = [("one", m, n, k, mode)]
    for name, m, n, k, mode in shapes:
        per = 2 * (m * k + n * k + m * n)
        copies = max(2, min(8, (160 << 20) // per + 1))
        A, B = operands(m, n, k, mode, copies)
        lda = A[0].shape[1]
        ldb = B[0].shape[1]
        out = [torch.empty(m, n, device="cuda", dtype=torch.bfloat16) for _ in range(copies)]
        ours = lambda i: native_tf.gemm_ex(A[i], B[i], out[i], mode, m, n, k, lda, ldb, n)
        t_ours = time_fn(ours, copies, args.iters)
        flops = 2.0 * m * n * k
        line = f"{name:11s} M={m:5d} N={n:5d} K={k:5d} mode={mode}  ours {t_ours:8.1f} us {flops / t_ours / 1e6:7.1f} TF"
        if not args.no_ref:
            t_ref = time_fn(lambda i: ref(A[i], B[i], mode), copies, args.iters)
            err = (out[0].float() - ref(A[0], B[0], mode).float()).abs().max().item()
            line += f" | cuBLAS {t_ref:8.1f} us {flops / t_ref / 1e6:7.1f} TF | ratio {t_ref / t_ours:5.2f} maxerr {err:.3g}"
        print(line, flush=True)
        if args.epilogues and mode != 2:
            bias = torch.randn(n, device="cuda", dtype=torch.bfloat16)
            aux = [torch.randn(m, n, device="cuda", dtype=torch.bfloat16) for _ in range(copies)]
            t_b = time_fn(lambda i: native_tf.gemm_ex(A[i], B[i], out[i], mode, m, n, k, lda, ldb, n, bias=bias), copies, args.iters)
            row = f"            +bias {t_b:7.1f} us"
            for epi, nm in ((3, "+bias+residual"), (1, "+bias+gelu"), (2, "*gelu'(aux)")):
                kw = dict(aux_out=aux[0]) if epi == 1 else dict(aux=aux[0])
                if native_tf.gemm_epi(A[0], B[0], out[0], mode, m, n, k, lda, ldb, n, epi, bias=bias if epi != 2 else None, **kw):
                    fn = lambda i: native_tf.gemm_epi(A[i], B[i], out[i], mode, m, n, k, lda, ldb, n, epi,
                                                      bias=bias if epi != 2 else None,
                                                      **(dict(aux_out=aux[i]) if epi == 1 else dict(aux=aux[i])))
                    row += f" | {nm} {time_fn(fn, copies, args.iters):7.1f} us"
            print(row, flush=True)
        del A, B, out


if __name__ == "__main__":
    main()
