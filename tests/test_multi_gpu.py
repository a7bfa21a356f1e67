"""The N-GPU path behind the C ABI (plu_gpu_create_multi: SURVEY.md 8b/8e) and the host wrappers around it.

One process drives all devices: the scene is replicated, device g renders the samples k = g (mod n) of the call's
sample set, ONE ncclReduce combines the float framebuffers on the first device.  The image must be the single-GPU
image up to the float summation order of the reduce; with one device it is the single-GPU path bit for bit.
Cases that need >= 2 GPUs are skipped on a 1-GPU box (run them with `gpurun --gpus 2`)."""
import os
import subprocess
import sys

import numpy as np
import pytest

from tests.conftest import ROOT
from tests.orc import GOLDEN, golden_scene
from tests.test_oracle import same_bits

pytestmark = pytest.mark.gpu


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def _close(a, b):
    ok = np.isfinite(a)
    return np.array_equal(ok, np.isfinite(b)) and np.allclose(a[ok], b[ok], rtol=2e-5, atol=1e-6)


def test_multi_context_with_one_device_is_the_single_gpu_path(gpu):
    from plutracer_b200 import GpuContext
    w, h, n = 64, 48, 4
    sc = golden_scene("cornell-box", w, h, n)
    gpu.upload(sc)
    p = GpuContext.params(w, h, n, seed=11)
    one = np.full((h, w, 3), 0.25, np.float32)
    gpu.render(p, one)
    m = GpuContext([0])
    assert m.n_devices == 1
    m.upload(sc)
    got = np.full((h, w, 3), 0.25, np.float32)
    st = m.render(p, got)
    assert st.n_devices == 1 and st.ms_reduce == 0.0
    assert same_bits(one, got).all()
    m.close()


def test_sample_range_beyond_the_strata_grid_is_rejected(gpu):
    from plutracer_b200 import GpuContext, _ffi
    w, h, n = 16, 16, 4
    gpu.upload(golden_scene("minimal0", w, h, n))
    img = np.zeros((h, w, 3), np.float32)
    gpu.render(GpuContext.params(w, h, n, sample_begin=3, sample_stride=4, sample_count=4), img)   # last index 15: fine
    with pytest.raises(_ffi.PluError) as e:
        gpu.render(GpuContext.params(w, h, n, sample_begin=4, sample_stride=4, sample_count=4), img)   # last index 16
    assert e.value.code == -1


def test_render_device_on_stream_zero_is_ordered_after_the_callers_work(gpu):
    """a NULL stream means stream 0 (the legacy default stream), so a framebuffer the caller initialises
    on torch's default stream is complete before the first accumulate — checked with a large fill in front."""
    import torch
    from plutracer_b200 import GpuContext
    w, h, n = 256, 256, 2
    gpu.upload(golden_scene("cornell-box", w, h, n))
    p = GpuContext.params(w, h, n, seed=3)
    want = np.full((h, w, 3), 2.0, np.float32)
    gpu.render(p, want)
    pad = torch.empty(1 << 28, dtype=torch.float32, device="cuda")   # 1 GiB of fills ahead of the framebuffer init
    fb = torch.empty(h, w, 3, dtype=torch.float32, device="cuda")
    for _ in range(3):
        pad.fill_(1.0)
        fb.fill_(2.0)
        gpu.render_device(p, fb.data_ptr(), None)
        assert same_bits(want, fb.cpu().numpy()).all()


def test_sharded_renderer_world_one_equals_renderer(lib):
    """plutracer_b200.sharded_renderer (the torch.distributed host) with one rank and a NON-ZERO target."""
    from plutracer_b200 import renderer, sharded_renderer, stratified_sampler, texture2d
    w, h, n = 96, 64, 4
    sc = golden_scene("room", w, h, n)
    a, b = texture2d((w, h)), texture2d((w, h))
    a.pixels[...] = 0.5; b.pixels[...] = 0.5
    r = renderer(sc, None, (32, 32), stratified_sampler((w, h), (n, n)), None)
    r.render(a)
    s = sharded_renderer(sc, None, (32, 32), stratified_sampler((w, h), (n, n)), None, rank=0, world=1)
    s.render(b)
    assert same_bits(a.pixels, b.pixels).all()
    r.ctx.close(); s.close()


@pytest.mark.parametrize("name,w,h,n", [("cornell-box", 128, 96, 6), ("room", 160, 90, 4), ("glass0", 64, 48, 3)])
def test_multi_gpu_context_matches_single_gpu(gpu, name, w, h, n):
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    from plutracer_b200 import GpuContext
    sc = golden_scene(name, w, h, n)
    gpu.upload(sc)
    p = GpuContext.params(w, h, n, seed=5)
    one = np.full((h, w, 3), 0.125, np.float32)
    st1 = gpu.render(p, one)
    m = GpuContext("all")
    assert m.n_devices == _n_gpus()
    m.upload(sc)
    got = np.full((h, w, 3), 0.125, np.float32)   # the caller's pixels enter the sum once (device 0 only)
    st = m.render(p, got)
    assert st.n_devices == m.n_devices and st.samples == st1.samples == w * h * n * n
    assert st.rays_extend == st1.rays_extend and st.rays_shadow == st1.rays_shadow   # the same paths, split
    assert _close(one, got)
    # a sample subset (what one rank of an outer split passes) is split again among the devices
    q = GpuContext.params(w, h, n, seed=5, sample_begin=1, sample_stride=2, sample_count=(n * n) // 2)
    a, b = np.zeros((h, w, 3), np.float32), np.zeros((h, w, 3), np.float32)
    gpu.render(q, a); m.render(q, b)
    assert _close(a, b)
    # fewer samples than devices: the idle devices contribute zeros
    q1 = GpuContext.params(w, h, n, seed=5, sample_begin=2, sample_stride=1, sample_count=1)
    a[...] = 0; b[...] = 0
    gpu.render(q1, a); m.render(q1, b)
    assert same_bits(a, b).all()
    # determinism of the multi-GPU image
    again = np.full((h, w, 3), 0.125, np.float32)
    m.render(p, again)
    assert same_bits(got, again).all()
    m.close()


def test_multi_gpu_full_size_cornell_slice(gpu):
    """1024x1024 /smp 64, 16 strata: every device gets several waves; device framebuffer variant."""
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    import torch
    from plutracer_b200 import GpuContext
    w, h, n = 1024, 1024, 64
    sc = golden_scene("cornell-box", w, h, n)
    gpu.upload(sc)
    p = GpuContext.params(w, h, n, seed=9, sample_begin=0, sample_stride=1, sample_count=16)
    one = np.zeros((h, w, 3), np.float32)
    gpu.render(p, one)
    m = GpuContext("all")
    m.upload(sc)
    fb = torch.zeros(h, w, 3, dtype=torch.float32, device="cuda:0")
    st = m.render_device(p, fb.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert st.n_devices == _n_gpus() and st.ms_reduce > 0
    assert _close(one, fb.cpu().numpy())
    m.close()


BIN = os.path.join(ROOT, "oracle", "_ref", "plu_ref_gpu")


def _dropin(args, env_extra, cwd=None):
    env = dict(os.environ); env.update(env_extra)
    res = subprocess.run([BIN] + [str(a) for a in args], capture_output=True, text=True, env=env, cwd=cwd)
    assert res.returncode == 0, res.stderr
    import json
    return json.loads(res.stdout.strip().splitlines()[-1])


def test_dropin_keeps_one_context_and_scales_over_devices(lib, tmp_path):
    """The reference process with plu::gpu_renderer: PLU_DEVICES=0 is bit-identical with the own host path; all
    visible devices give the same image up to summation order; repeated renders reuse the context (the second
    call does not pay CUDA / NCCL set-up again)."""
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/plu_ref_gpu is built only where the reference tree exists (oracle/build_ref.sh)")
    from plutracer_b200 import renderer, stratified_sampler, texture2d
    w, h, n = 128, 96, 4
    scene = os.path.join(GOLDEN, "cornell-box.pluscene")
    o1 = str(tmp_path / "one.bin")
    j1 = _dropin(["render-gpu-flat", scene, w, h, n, o1], {"PLU_DEVICES": "0", "PLU_REF_REPEAT": "3"})
    assert j1["devices"] == 1
    one = np.fromfile(o1, np.float32).reshape(h, w, 3)
    tx = texture2d((w, h))
    r = renderer(golden_scene("cornell-box", w, h, n), None, (32, 32), stratified_sampler((w, h), (n, n)), None)
    r.render(tx)
    assert same_bits(one, tx.pixels).all()
    assert j1["ms"] < j1["ms_first_call"]   # context, wave buffers and framebuffer are reused
    if _n_gpus() >= 2:
        o2 = str(tmp_path / "all.bin")
        env = {k: v for k, v in os.environ.items() if k != "PLU_DEVICES"}
        res = subprocess.run([BIN, "render-gpu-flat", scene, str(w), str(h), str(n), o2], capture_output=True, text=True, env=env)
        assert res.returncode == 0, res.stderr
        import json
        j2 = json.loads(res.stdout.strip().splitlines()[-1])
        assert j2["devices"] == _n_gpus()
        assert _close(one, np.fromfile(o2, np.float32).reshape(h, w, 3))
    r.ctx.close()


def test_dropin_resident_framebuffer_flow(lib, tmp_path):
    """gpu_render_post_write: render -> tone map -> .bmp from the framebuffer that stayed on the device ==
    the three separate drop-in calls (gpu_renderer, gpu_postprocesser, gpu_write_bmp)."""
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/plu_ref_gpu is built only where the reference tree exists (oracle/build_ref.sh)")
    w, h, n = 64, 48, 3
    scene = os.path.join(GOLDEN, "room.pluscene")
    a_img, a_bmp, b_img, b_bmp = (str(tmp_path / f) for f in ("a.bin", "a.bmp", "b.bin", "b.bmp"))
    _dropin(["render-gpu-flat", scene, w, h, n, a_img, a_bmp], {"PLU_DEVICES": "0"})
    _dropin(["render-gpu-flat", scene, w, h, n, b_img, b_bmp], {"PLU_DEVICES": "0", "PLU_REF_FUSED": "1"})
    assert open(a_bmp, "rb").read() == open(b_bmp, "rb").read()
    # plu::gpu::pinned_target around the reference's own texture2d (plu_host_register): same bytes
    c_img, c_bmp = str(tmp_path / "c.bin"), str(tmp_path / "c.bmp")
    _dropin(["render-gpu-flat", scene, w, h, n, c_img, c_bmp], {"PLU_DEVICES": "0", "PLU_REF_PIN": "1", "PLU_REF_REPEAT": "2"})
    assert open(a_bmp, "rb").read() == open(c_bmp, "rb").read() and open(a_img, "rb").read() == open(c_img, "rb").read()
    from plutracer_b200 import GpuContext
    lin = np.fromfile(a_img, np.float32).reshape(h, w, 3)
    ctx = GpuContext(0)
    want = ctx.postprocess(lin.copy())
    assert same_bits(want, np.fromfile(b_img, np.float32).reshape(h, w, 3)).all()
    ctx.close()


_RANK_SCRIPT = r"""
import os, sys, numpy as np, torch, torch.distributed as dist
sys.path.insert(0, sys.argv[1])
from plutracer_b200 import sharded_renderer, stratified_sampler, texture2d
from tests.orc import golden_scene
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
w, h, n = 96, 64, 4
tx = texture2d((w, h)); tx.pixels[...] = 0.5
r = sharded_renderer(golden_scene("room", w, h, n), None, (32, 32), stratified_sampler((w, h), (n, n)), None, device=local, rank=rank, world=world)
r.render(tx)
if rank == 0:
    np.save(sys.argv[2], tx.pixels.reshape(h, w, 3))
dist.barrier(); dist.destroy_process_group()
"""


def test_sharded_renderer_two_ranks_nccl(lib, tmp_path):
    """sharded_renderer (one process per GPU, NCCL reduce through torch.distributed) against renderer.render."""
    if _n_gpus() < 2:
        pytest.skip("needs >= 2 GPUs")
    from plutracer_b200 import renderer, stratified_sampler, texture2d
    script = tmp_path / "rank.py"; script.write_text(_RANK_SCRIPT)
    out = str(tmp_path / "sharded.npy")
    res = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                          "--master-port", "29533", str(script), ROOT, out], capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    w, h, n = 96, 64, 4
    tx = texture2d((w, h)); tx.pixels[...] = 0.5
    r = renderer(golden_scene("room", w, h, n), None, (32, 32), stratified_sampler((w, h), (n, n)), None)
    r.render(tx)
    assert _close(tx.pixels.reshape(h, w, 3), np.load(out))
    r.ctx.close()
