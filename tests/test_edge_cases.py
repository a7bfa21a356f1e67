"""Edge cases of the path, CUDA vs oracle bit for bit (per-sample radiance and primary hits): a tree that is a
single leaf (bvh_tree.cpp:8-12: the root IS the surface and sees the caller's hit_record::t), scenes without
lights, sphere area lights, 1x1 images, one stratum, path-length limits, an empty object list."""
import numpy as np
import pytest

HEADER = "[\n\tresolution: [{w} {h}]\n\tantialiasing-samples: {n}\n\tcamera: [\n\t\tposition:\t[0 1 -8]\n\t\ttarget:\t\t[0 1 0]\n\t\tup:\t\t[0 1 0]\n\t]\n"
SCENES = {
    # one surface: the BVH root is a leaf
    "single-sphere": "\tmaterials: [\n\t]\n\tobjects:  [\n\t\tsphere [0 1 0] 1.5 [ diffuse [0.6 0.5 0.4] ]\n\t\tpoint-light [3 5 -4] [60 60 60]\n\t]\n]\n",
    # no light at all: every NEE branch is skipped, radiance is identically zero
    "no-lights": "\tmaterials: [\n\t]\n\tobjects:  [\n\t\tbox [0 -1 0] [4 0.2 4] [ diffuse [0.5 0.5 0.5] ]\n\t\tsphere [0 1 0] 1 [ glass [1 1 1] 1.5 ]\n\t]\n]\n",
    # sphere area light (surface::sample / pdf of a sphere, sphere.h:17-22 incl. the volume-for-area quirk) over mirror + diffuse
    "sphere-light": "\tmaterials: [\n\t\tfloor: [ diffuse [0.7 0.7 0.7] ]\n\t]\n\tobjects:  [\n\t\tdiffuse-area-light [ sphere [0 4 0] 0.75 ] [40 40 35]\n"
                    "\t\tbox [0 -1 0] [5 0.2 5] 'floor\n\t\tsphere [-1.5 0.5 0] 1 [ perfect-reflection [1 1 1] [0.2 0.9 1.1] [3.9 2.4 2.2] ]\n"
                    "\t\tbox [2 0.5 1] [0.8 1.2 0.8] [ diffuse texture [ checkerboard [1 1 1] [0.1 0.1 0.1] 2 ] ]\n\t]\n]\n",
    # only an emitter in view
    "light-only": "\tmaterials: [\n\t]\n\tobjects:  [\n\t\tdiffuse-area-light [ box [0 1 0] [1 1 1] ] [5 4 3]\n\t]\n]\n",
}


def _load(tmp_path, name, w, h, n):
    from plutracer_b200 import Scene
    f = tmp_path / (name + ".urn")
    f.write_text(HEADER.format(w=w, h=h, n=n) + SCENES[name])
    return Scene.load_urn(str(f))


def _bits_equal(a, b):
    return ((a.view(np.uint32) == b.view(np.uint32)) | (np.isnan(a) & np.isnan(b))).all()


@pytest.mark.parametrize("name", sorted(SCENES))
def test_oracle_handles_edge_scenes(lib, oracle, tmp_path, name):
    from plutracer_b200 import GpuContext
    sc = _load(tmp_path, name, 16, 12, 2)
    img, st = oracle.render(sc, GpuContext.params(16, 12, 2, seed=3), trig=0, n_threads=1)
    assert st.samples == 16 * 12 * 4
    if name == "no-lights":
        assert not np.nansum(np.abs(img)) > 0
    if name == "light-only":
        assert np.nanmax(img) > 0


@pytest.mark.gpu
@pytest.mark.parametrize("name", sorted(SCENES))
def test_edge_scene_paths_match_oracle(gpu, oracle, tmp_path, name):
    from plutracer_b200 import GpuContext
    w, h, n = 40, 30, 3
    sc = _load(tmp_path, name, w, h, n)
    gpu.upload(sc)
    a = gpu.trace_primary(w, h)
    b = oracle.trace_primary(sc, w, h, 0)
    assert np.array_equal(a[0], b[0]) and _bits_equal(a[1], b[1]) and _bits_equal(a[2], b[2])
    for mv in (1, 2, 8, 33):
        p = GpuContext.params(w, h, n, seed=5, max_vertices=mv)
        got = gpu.sample_radiance(p, 0, w * h)
        want = oracle.sample_radiance(sc, p, 0, w * h, trig=0)
        assert _bits_equal(got, want), (name, mv)


@pytest.mark.gpu
def test_tiny_images_and_single_stratum(gpu, oracle, tmp_path):
    from plutracer_b200 import GpuContext
    for (w, h, n) in ((1, 1, 1), (1, 7, 2), (5, 1, 1), (3, 2, 5)):
        sc = _load(tmp_path, "sphere-light", w, h, n)
        gpu.upload(sc)
        p = GpuContext.params(w, h, n, seed=9)
        img = np.full((h, w, 3), 0.25, np.float32)
        gpu.render(p, img)
        want, _ = oracle.render(sc, p, out=np.full((h, w, 3), 0.25, np.float32), trig=0, n_threads=1)
        assert _bits_equal(img, want), (w, h, n)


@pytest.mark.gpu
def test_empty_object_list(gpu, tmp_path):
    """The reference cannot build a bvh_tree over zero surfaces; here an empty scene renders black."""
    from plutracer_b200 import GpuContext, Scene, _ffi
    f = tmp_path / "empty.urn"
    f.write_text(HEADER.format(w=8, h=8, n=1) + "\tmaterials: [\n\t]\n\tobjects:  [\n\t]\n]\n")
    sc = Scene.load_urn(str(f))
    assert sc.n_prims == 0
    try:
        gpu.upload(sc)
    except _ffi.PluError:
        return   # rejected loudly is fine too
    img = np.zeros((8, 8, 3), np.float32)
    st = gpu.render(GpuContext.params(8, 8, 1), img)
    assert st.samples == 64 and not img.any()
    ids, t, nrm = gpu.trace_primary(8, 8)
    assert (ids == -1).all()


@pytest.mark.gpu
def test_upload_of_an_unchanged_scene_keeps_the_device_tables(lib, tmp_path):
    """plu_gpu_upload_scene compares the tables it is given with those on the device, byte for byte: the same scene again
    (what the C++ drop-in does before every render) rebuilds nothing (plu_build_stats.reserved == 1); a camera-only change
    keeps the structures and takes the new camera; any other change rebuilds.  Images always equal a fresh context's."""
    from plutracer_b200 import GpuContext, Scene, _ffi
    w, h, n = 48, 36, 3
    sc = _load(tmp_path, "sphere-light", w, h, n)
    p = GpuContext.params(w, h, n, seed=4)

    def fresh(scene):
        c = GpuContext(0)
        try:
            c.upload(scene)
            img = np.zeros((h, w, 3), np.float32)
            c.render(p, img)
            return img
        finally:
            c.close()

    ctx = GpuContext(0)
    try:
        assert ctx.upload(sc).reserved == 0
        a = np.zeros((h, w, 3), np.float32); ctx.render(p, a)
        bs = ctx.upload(sc)
        assert bs.reserved == 1 and bs.ms_total == 0.0 and bs.built_on_device == 1
        b = np.zeros((h, w, 3), np.float32); ctx.render(p, b)
        assert _bits_equal(a, b) and np.nansum(a) > 0
        # camera only: another resolution -> inv_image_size differs, tables are the same
        d = _ffi.SceneDesc.from_buffer_copy(sc.desc)
        d.cam.pos[0] += 0.75
        moved = Scene.from_desc(d, w, h, n)
        assert ctx.upload(moved).reserved == 1
        c_img = np.zeros((h, w, 3), np.float32); ctx.render(p, c_img)
        assert not _bits_equal(a, c_img) and _bits_equal(c_img, fresh(moved))
        # one surface moved: rebuilt
        f = tmp_path / "changed.urn"
        f.write_text(HEADER.format(w=w, h=h, n=n) + SCENES["sphere-light"].replace("sphere [-1.5 0.5 0] 1", "sphere [-1.25 0.5 0] 1"))
        changed = Scene.load_urn(str(f))
        assert ctx.upload(changed).reserved == 0
        d_img = np.zeros((h, w, 3), np.float32); ctx.render(p, d_img)
        assert not _bits_equal(a, d_img) and _bits_equal(d_img, fresh(changed))
        # and back: rebuilt again (only the scene that IS on the device is remembered), same image as at first
        assert ctx.upload(sc).reserved == 0
        e = np.zeros((h, w, 3), np.float32); ctx.render(p, e)
        assert _bits_equal(a, e)
        # the host build is another build: not confused with the device build of the same tables
        assert ctx.upload(sc, build="host").reserved == 0 and ctx.upload(sc, build="host").reserved == 1
    finally:
        ctx.close()


@pytest.mark.gpu
def test_page_locked_targets_change_nothing_but_the_copy_rate(gpu, lib):
    """plu_host_alloc / plu_host_register: a render into a page-locked target (texture2d(pinned=True)), into a registered
    numpy array and into pageable memory gives the same bits, incl. a NON-ZERO target (`+=`); the allocation is zero-filled."""
    import ctypes as C
    from plutracer_b200 import GpuContext, _ffi, renderer, sharded_renderer, stratified_sampler, texture2d
    from tests.orc import golden_scene
    w, h, n = 80, 56, 3
    sc = golden_scene("room", w, h, n)
    r = renderer(sc, None, (32, 32), stratified_sampler((w, h), (n, n)), None)
    a, b = texture2d((w, h)), texture2d((w, h), pinned=True)
    assert b.pinned and not a.pinned and not b.pixels.any()
    a.pixels[...] = 0.25; b.pixels[...] = 0.25
    r.render(a); r.render(b)
    assert _bits_equal(a.pixels, b.pixels) and np.nansum(a.pixels) > 0.25 * a.pixels.size
    c = np.full((h, w, 3), 0.25, np.float32)
    assert lib.plu_host_register(c.ctypes.data, c.nbytes) == _ffi.PLU_OK
    try:
        r.ctx.render(GpuContext.params(w, h, n, seed=r.seed), c)
    finally:
        assert lib.plu_host_unregister(c.ctypes.data) == _ffi.PLU_OK
    assert _bits_equal(a.pixels, c)
    assert lib.plu_host_unregister(c.ctypes.data) == _ffi.PLU_E_CUDA   # not registered any more: refused, and no sticky error is left
    r.render(texture2d((w, h)))
    s = sharded_renderer(sc, None, (32, 32), stratified_sampler((w, h), (n, n)), None, rank=0, world=1)
    d = texture2d((w, h), pinned=True); d.pixels[...] = 0.25
    s.render(d)
    assert _bits_equal(a.pixels, d.pixels)
    r.ctx.close(); s.close()
