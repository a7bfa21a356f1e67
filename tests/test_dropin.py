"""The drop-in, end to end: oracle/_ref/plu_ref_gpu is the UNMODIFIED reference (its urn loader, scene classes,
live bvh_tree, stratified_sampler and texture2d; oracle/build_ref.sh compiles it from /root/reference in the build
container) with ONE type changed — `plu::gpu_renderer` (integration/plu_gpu_renderer.h) where main.cpp:171 says
`plu::renderer` — linked against libplugpu.so.  Its images must equal, bit for bit, what this repo's own host
path (urn loader / flat scene -> C ABI) renders for the same scene, size and seed; its BMP (gpu_postprocesser +
gpu_write_bmp for main.cpp:183,209) must be the same bytes."""
import os
import subprocess

import numpy as np
import pytest

from tests.conftest import ROOT
from tests.orc import GOLDEN, golden_scene
from tests.test_oracle import same_bits

pytestmark = pytest.mark.gpu
BIN = os.path.join(ROOT, "oracle", "_ref", "plu_ref_gpu")
ONE_GPU = dict(os.environ, PLU_DEVICES="0")   # bit-identity holds on one device; all visible devices: tests/test_multi_gpu.py


def _need_bin():
    if not os.path.exists(BIN):
        pytest.skip("oracle/_ref/plu_ref_gpu is built only where the reference tree exists (oracle/build_ref.sh)")


def _own_render(sc, w, h, n):
    from plutracer_b200 import renderer, stratified_sampler, texture2d
    tx = texture2d((w, h))
    r = renderer(sc, None, (32, 32), stratified_sampler((w, h), (n, n)), None)
    st = r.render(tx)
    assert st.samples == w * h * n * n
    return tx, r.ctx


@pytest.mark.parametrize("name", ["cornell-box", "glass0", "refrac0", "room"])
def test_reference_object_graph_through_gpu_renderer(lib, tmp_path, name):
    """the reference's own surface / material / light / texture objects (rebuilt from the flat fixture, since the
    .urn files of the reference do not travel), flattened by the integration header from its live tree"""
    _need_bin()
    w, h, n = 64, 48, 4
    out = str(tmp_path / "img.bin")
    res = subprocess.run([BIN, "render-gpu-flat", os.path.join(GOLDEN, name + ".pluscene"), str(w), str(h), str(n), out], capture_output=True, text=True, env=ONE_GPU)
    assert res.returncode == 0, res.stderr
    got = np.fromfile(out, np.float32).reshape(h, w, 3)
    tx, _ = _own_render(golden_scene(name, w, h, n), w, h, n)
    assert same_bits(got, tx.pixels).all()
    assert np.nansum(got) > 0


@pytest.mark.parametrize("name", ["atrium", "colonnade", "shard"])
def test_reference_loader_through_gpu_renderer_and_bmp(lib, tmp_path, name):
    """scenes/ of this repo through the REFERENCE's urn loader and scene constructor (scripted objects, lens camera,
    triangle-mesh OBJ), then gpu_postprocesser + gpu_write_bmp"""
    _need_bin()
    from plutracer_b200 import Scene, postprocesser, write_bmp
    w, h, n = 64, 48, 3
    out, bmp = str(tmp_path / "img.bin"), str(tmp_path / "img.bmp")
    scenes = os.path.join(ROOT, "scenes")
    res = subprocess.run([BIN, "render-gpu", os.path.join(scenes, name + ".urn"), str(w), str(h), str(n), out, bmp], capture_output=True, text=True, cwd=scenes, env=ONE_GPU)
    assert res.returncode == 0, res.stderr
    got = np.fromfile(out, np.float32).reshape(h, w, 3)
    cwd = os.getcwd()
    os.chdir(scenes)   # triangle-mesh paths are relative to the working directory (scene.h:138)
    try:
        sc = Scene.load_urn(os.path.join(scenes, name + ".urn")).override(w, h, n)
    finally:
        os.chdir(cwd)
    tx, ctx = _own_render(sc, w, h, n)
    assert same_bits(got, tx.pixels).all()
    postprocesser(ctx).postprocess(tx)
    mine = str(tmp_path / "mine.bmp")
    write_bmp(tx, mine, ctx)
    assert open(bmp, "rb").read() == open(mine, "rb").read()
