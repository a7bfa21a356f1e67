"""Samples-per-pixel sharding of one render across the GPUs of a box (one process per GPU).

The reference renders every sample of a pixel on one CPU thread (src/renderer.cpp:128-133); samples are
independent, so the path shards by SAMPLE: rank r of N takes the strata s with s = r (mod N) of every
pixel — the per-pixel stratification (sampler.h:58-69) is kept, the scene and BVH are replicated, nothing
is exchanged while rendering — and the N float framebuffers are combined by ONE sum-reduce to rank 0
(NCCL over NVLink on the GPU box, gloo in the CPU tests).  Philox counters carry the global sample index,
so the image is independent of N up to the float summation order of the reduce.
"""
import numpy as np

from .renderer import DEFAULT_SEED, GpuContext


def shard_plan(n_strata, world, rank, spp=None):
    """(sample_begin, sample_stride, sample_count) of `rank`: strata rank, rank + world, ... below `spp`
    (default n_strata**2).  Ranks beyond the sample count get an empty shard (count 0 -> caller skips)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad world/rank")
    total = n_strata * n_strata if spp is None else int(spp)
    if not 0 < total <= n_strata * n_strata:
        raise ValueError("spp must be in 1..n_strata^2")
    count = (total - rank + world - 1) // world if rank < total else 0
    return rank, world, count


def shard_samples(n_strata, world, rank, spp=None):
    """The global sample indices of a shard (for tests and book-keeping)."""
    b, s, c = shard_plan(n_strata, world, rank, spp)
    return np.arange(c, dtype=np.int64) * s + b


def reduce_framebuffer(fb, dst=0, group=None):
    """Sum the per-rank framebuffers (torch tensors, CPU for gloo / CUDA for nccl) into rank `dst`:
    the single collective of the path.  No-op without an initialised process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(fb, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return fb


class sharded_renderer:
    """plu::renderer for one rank of an N-GPU job: render(target) renders this rank's strata on its GPU
    into a device framebuffer, reduces to rank 0 and (rank 0 only) adds the result to the host target —
    `+=` like renderer.cpp:132.  Needs torch for the device buffer and the process group."""

    def __init__(self, S, C_=None, tile_size=(32, 32), smp=None, l=None, device=0, rank=0, world=1, seed=DEFAULT_SEED, flags=0):
        if smp is None:
            raise ValueError("renderer needs a sampler (sample_count gives the strata per pixel)")
        self.scene, self.main_sampler, self.rank, self.world = S, smp, int(rank), int(world)
        self.seed, self.flags, self.device = seed, flags, int(device)
        self.ctx = GpuContext(device)
        self.ctx.upload(S)
        self._fb = self._host = None
        self.last_stats = None

    def render(self, target, spp=None):
        import torch
        W, H = target.size
        n = self.main_sampler.sample_count[0]
        begin, stride, count = shard_plan(n, self.world, self.rank, spp)
        dev = torch.device("cuda", self.device)
        if self._fb is None or self._fb.shape != (H, W, 3):
            self._fb = torch.empty(H, W, 3, dtype=torch.float32, device=dev)
            self._host = torch.empty(H, W, 3, dtype=torch.float32).pin_memory()
        stream = torch.cuda.current_stream(dev)
        # a page-locked target (texture2d(pinned=True)) is copied from and to directly; a pageable one goes through self._host
        direct = torch.from_numpy(target.pixels.reshape(H, W, 3)) if getattr(target, "pinned", False) else None
        if self.rank == 0:   # the caller's pixels enter the sum once
            if direct is not None:
                self._fb.copy_(direct, non_blocking=True)
            else:
                self._host.copy_(torch.from_numpy(target.pixels.reshape(H, W, 3)))
                self._fb.copy_(self._host, non_blocking=True)
        else:
            self._fb.zero_()
        if count > 0:
            p = GpuContext.params(W, H, n, self.seed, self.flags, sample_begin=begin, sample_stride=stride, sample_count=count)
            self.last_stats = self.ctx.render_device(p, self._fb.data_ptr(), stream.cuda_stream)
        reduce_framebuffer(self._fb, dst=0)
        if self.rank == 0:
            (direct if direct is not None else self._host).copy_(self._fb, non_blocking=True)
            stream.synchronize()
            if direct is None:
                target.pixels.reshape(H, W, 3)[...] = self._host.numpy()
        return self.last_stats

    def close(self):
        self.ctx.close()
