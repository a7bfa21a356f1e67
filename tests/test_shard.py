"""The N > 1 path on CPU: shard plan + the single sum-reduce, world_size 2 over gloo.  The per-rank render
is the CPU oracle here (test infrastructure standing in for the GPU kernels, same sample-range parameters);
on the GPU box test_gpu_parity.test_sample_sharding_is_linear covers the same property through the C ABI."""
import os
import socket

import numpy as np
import pytest

from plutracer_b200.shard import shard_plan, shard_samples


def test_shard_plan_partitions_the_strata():
    for n, world in ((8, 1), (8, 2), (8, 8), (3, 2), (3, 4), (2, 8), (64, 8)):
        seen = np.concatenate([shard_samples(n, world, r) for r in range(world)])
        assert sorted(seen.tolist()) == list(range(n * n))
        counts = [shard_plan(n, world, r)[2] for r in range(world)]
        assert max(counts) - min(counts) <= 1
    # a slice of the strata (bench.py --spp-per-step)
    seen = np.concatenate([shard_samples(8, 4, r, spp=10) for r in range(4)])
    assert sorted(seen.tolist()) == list(range(10))
    assert shard_plan(2, 8, 7) == (7, 8, 0)   # more ranks than samples: empty shard
    with pytest.raises(ValueError):
        shard_plan(8, 2, 2)
    with pytest.raises(ValueError):
        shard_plan(8, 2, 0, spp=65)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, name, w, h, n, tmp):
    import torch
    import torch.distributed as dist
    from plutracer_b200 import GpuContext
    from plutracer_b200.shard import reduce_framebuffer
    from tests.orc import Oracle, golden_scene
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        sc = golden_scene(name, w, h, n)
        begin, stride, count = shard_plan(n, world, rank)
        p = GpuContext.params(w, h, n, seed=7, sample_begin=begin, sample_stride=stride, sample_count=count)
        img, st = Oracle().render(sc, p, trig=0, n_threads=1)
        assert st.samples == w * h * count
        np.save(os.path.join(tmp, f"shard{rank}.npy"), img)
        fb = torch.from_numpy(img.copy())
        reduce_framebuffer(fb, dst=0)
        if rank == 0:
            np.save(os.path.join(tmp, "reduced.npy"), fb.numpy())
        dist.barrier()
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("name", ["cornell-box", "room"])
def test_two_rank_gloo_reduce_equals_single_rank_render(lib, oracle, tmp_path, name):
    import torch.multiprocessing as mp
    from plutracer_b200 import GpuContext
    from tests.orc import golden_scene
    w, h, n, world = 24, 18, 4, 2
    mp.spawn(_rank_main, args=(world, _free_port(), name, w, h, n, str(tmp_path)), nprocs=world, join=True)
    shards = [np.load(tmp_path / f"shard{r}.npy") for r in range(world)]
    reduced = np.load(tmp_path / "reduced.npy")
    # the collective is exactly the float sum of the shards ...
    assert np.array_equal(reduced.view(np.uint32), (shards[0] + shards[1]).view(np.uint32))
    # ... and the shards are a partition of the single-rank render: same samples, other summation order
    full, st = oracle.render(golden_scene(name, w, h, n), GpuContext.params(w, h, n, seed=7), trig=0, n_threads=1)
    assert st.samples == w * h * n * n
    ok = np.isfinite(full)
    assert np.allclose(reduced[ok], full[ok], rtol=2e-5, atol=1e-6)
    assert np.array_equal(np.isfinite(reduced), ok)
