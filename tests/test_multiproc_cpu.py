"""CPU tests of the runner's host logic and of the N>1 plumbing (world_size-2 ``gloo``): schedule vs the golden fixture,
seq / select_index construction (runners/diffusion.py:388-394,482-492), shard bounds incl. ragged and empty shards, the
host-side gather of sharded sampling, and the flat-buffer gradient all-reduce of data-parallel training."""
import argparse
import os

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def runner_mod():
    import __graft_entry__ as g
    g.build()
    from ddim_audio_b200.runners import diffusion
    return diffusion


def _args(**kw):
    d = dict(sample_type="generalized", skip_type="uniform", timesteps=1000, sequence=10, eta=0.0, seed=1234)
    d.update(kw)
    return argparse.Namespace(**d)


def test_schedule_matches_reference_golden(runner_mod, golden_dir):
    from ddim_audio_b200.config import audio_config
    r = runner_mod.Diffusion(_args(), audio_config("torch.FloatTensor"), device="cpu")
    want = np.load(os.path.join(golden_dir, "schedule.npz"))["alphas"]
    assert r.num_timesteps == 1000 and r.alphas.dtype == torch.float32
    assert np.array_equal(r.alphas.numpy(), want)              # fp32 cast THEN cumprod: bit-exact with the reference


def test_select_index_and_shards(runner_mod):
    from ddim_audio_b200.config import audio_config
    from ddim_audio_b200.runners.dist import shard_bounds
    r = runner_mod.Diffusion(_args(), audio_config("torch.FloatTensor"), device="cpu")
    assert r.select_index() == {999, 888, 777, 666, 555, 444, 333, 222, 111, 0}
    r.args.sequence = 0
    assert list(r.select_index()) == list(range(1000))
    t = runner_mod.antithetic_timesteps(5, 1000, torch.Generator().manual_seed(0))
    assert t.shape == (5,) and int(t[0] + t[3]) == 999 and int(t[1] + t[4]) == 999
    for n, w in [(512, 8), (64, 4), (7, 4), (2, 4), (0, 2)]:
        b = [shard_bounds(n, r_, w) for r_ in range(w)]
        assert b[0][0] == 0 and b[-1][1] == n and all(b[i][1] == b[i + 1][0] for i in range(w - 1))
        sizes = [hi - lo for lo, hi in b]
        assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, tmp):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    from ddim_audio_b200.runners import dist as rd
    r, _, w = rd.init_from_env("gloo")
    assert (r, w) == (rank, world) == rd.rank_world()
    # --- sharded sampling: every rank draws the same x_T, processes its slice, rank 0 gets the 1-rank result
    g = torch.Generator().manual_seed(7)
    full = torch.randn(5, 2, 4, 8, generator=g)
    lo, hi = rd.shard_bounds(5, rank, world)
    local = [full[lo:hi] * 2.0, full[lo:hi] + 1.0]
    out = rd.gather_host_lists(local)
    if rank == 0:
        assert torch.equal(out[0], full * 2.0) and torch.equal(out[1], full + 1.0)
    else:
        assert out is None
    # --- DP gradients: flat buffer path (one collective) and the packed fallback
    torch.manual_seed(100 + rank)
    flat = torch.zeros(10)
    p1, p2 = torch.nn.Parameter(torch.zeros(2, 3)), torch.nn.Parameter(torch.zeros(4))
    p1.grad, p2.grad = flat[:6].view(2, 3), flat[6:]
    flat.copy_(torch.arange(10.0) * (rank + 1))
    rd.allreduce_gradients([p1, p2], flat)
    assert torch.equal(flat, torch.arange(10.0) * 3) and p1.grad.data_ptr() == flat.data_ptr()
    q = torch.nn.Parameter(torch.zeros(3))
    q.grad = torch.full((3,), float(rank + 1))
    rd.allreduce_gradients([q, torch.nn.Parameter(torch.zeros(1))], flat)
    assert torch.equal(q.grad, torch.full((3,), 3.0))
    dist.barrier()
    open(os.path.join(tmp, f"ok{rank}"), "w").write("ok")
    dist.destroy_process_group()


def test_gloo_world2_shard_gather_and_allreduce(runner_mod, tmp_path):
    port = 29500 + (os.getpid() % 2000)
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    assert os.path.exists(tmp_path / "ok0") and os.path.exists(tmp_path / "ok1")


def test_numa_binding_is_a_noop_without_a_gpu():
    """runners/dist.py::bind_to_gpu_numa must never raise and never shrink the affinity mask when NVML / the GPU is unavailable."""
    import os

    from ddim_audio_b200.runners.dist import bind_to_gpu_numa

    before = os.sched_getaffinity(0)
    ok = bind_to_gpu_numa(0)
    after = os.sched_getaffinity(0)
    assert ok in (True, False)
    assert after == before or (ok and after and after <= before)
