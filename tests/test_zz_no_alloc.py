"""-m gpu: the allocation-free entry point ``ddimb_denoise2d_static`` (scratch in a file-scope device array of the library, no
counterpart in the reference) against the caller-workspace entry point ``ddimb_denoise2d``, through the C ABI, on two calls in a
row with different shapes (the shared array keeps the previous call's contents).  Collected last on purpose.

Bar: both entry points launch the same kernels on the same inputs; the only difference is where the scratch lives, and the column
sums use fp32 atomics whose order may differ between launches -> max-abs <= 2e-5 (the tolerance of tests/test_gpu_denoise.py)."""
import pytest
import torch

pytestmark = pytest.mark.gpu

TOL = 2e-5


def _both(x, iters=5):
    from ddim_audio_b200 import _C
    from ddim_audio_b200._C import call, ptr

    B, Cc, T, F = x.shape
    need = int(_C.lib.ddimb_denoise2d_workspace(B, Cc, T, F, iters))
    assert 0 < need <= int(_C.lib.ddimb_denoise2d_static_max_bytes())
    ws = torch.empty(need, dtype=torch.uint8, device=x.device)
    ref = torch.empty_like(x)
    call("ddimb_denoise2d", ptr(x, torch.float32), ptr(ref, torch.float32), ptr(ws), need, B, Cc, T, F, iters)
    got = torch.empty_like(x)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    call("ddimb_denoise2d_static", ptr(x, torch.float32), ptr(got, torch.float32), B, Cc, T, F, iters)
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    return got, ref


def test_static_scratch_matches_workspace_entry_point_twice():
    g = torch.Generator(device="cuda").manual_seed(41)
    # two calls in a row: the second has a different (smaller) carve of the same array, over the first call's leftovers
    for shape in [(2, 2, 1030, 256), (1, 2, 257, 128), (2, 2, 1030, 256)]:
        x = torch.randn(*shape, device="cuda", generator=g) * 1.5 + 0.25
        got, ref = _both(x)
        assert torch.isfinite(got).all()
        assert float((got - ref).abs().max()) <= TOL, shape
    x = torch.randn(3, 2, 64, 512, device="cuda", generator=g)
    got, ref = _both(x, iters=0)                       # medians only: selection + one subtraction, bit-exact
    assert torch.equal(got, ref)


def test_static_scratch_rejects_oversized_shapes():
    from ddim_audio_b200 import _C

    x = torch.zeros(8, 2, 8192, 256, device="cuda")
    out = torch.empty_like(x)
    assert int(_C.lib.ddimb_denoise2d_workspace(8, 2, 8192, 256, 5)) > int(_C.lib.ddimb_denoise2d_static_max_bytes())
    with pytest.raises(RuntimeError, match="ddimb_denoise2d_static"):
        _C.call("ddimb_denoise2d_static", _C.ptr(x, torch.float32), _C.ptr(out, torch.float32), 8, 2, 8192, 256, 5)
