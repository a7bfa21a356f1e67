"""the traversal structures built ON THE DEVICE (csrc/bvh_gpu.cu) against the host restatement
of the reference's builder (csrc/host/bvh_build.cpp, itself pinned against the reference's live tree by
tests/test_oracle.py / test_urn_loader.py): every device table byte for byte, on the golden scenes, on synthetic
scenes of sizes that cross every code path (<= 2 surfaces, one warp, > 1024 = element-parallel levels), on
lattices with equal box centres (the host's unstable std::sort decides: host_tie_segments > 0) and with the tie
ranks given by the caller.  Renders through both builds are bit-identical."""
import ctypes as C

import numpy as np
import pytest

from plutracer_b200 import GpuContext, _ffi
from tests.conftest import SCENES
from tests.orc import golden_scene

pytestmark = pytest.mark.gpu

TABLES = {"ref_nodes": _ffi.TABLE_REF_NODES, "prim_meta": _ffi.TABLE_PRIM_META, "fast_nodes": _ffi.TABLE_FAST_NODES,
          "fast_index": _ffi.TABLE_FAST_INDEX, "fast_leaf": _ffi.TABLE_FAST_LEAF}


class _NoRank:
    """The same scene tables with prim_tie_rank = NULL, so that the library derives the leaf order itself."""

    def __init__(self, scene):
        self._keep = scene
        self._d = _ffi.SceneDesc.from_buffer_copy(scene.desc)
        self._d.prim_tie_rank = C.POINTER(C.c_int32)()
        self.desc_ptr = C.pointer(self._d)
        self.n_prims = scene.n_prims


def _tables(ctx, scene, build):
    bs = ctx.upload(scene, build=build)
    return {k: ctx.download_table(v) for k, v in TABLES.items()}, bs


def _assert_same_tables(ctx, scene, what):
    host, hs = _tables(ctx, scene, "host")
    dev, ds = _tables(ctx, scene, "device")
    assert hs.built_on_device == 0 and (ds.built_on_device == 1 or scene.n_prims == 0)
    for k in TABLES:
        assert host[k].size == dev[k].size, (what, k, host[k].size, dev[k].size)
        if not np.array_equal(host[k], dev[k]):
            w = 4
            a, b = host[k].view(np.uint32), dev[k].view(np.uint32)
            bad = np.flatnonzero(a != b)
            raise AssertionError(f"{what}: table {k} differs in {bad.size} of {a.size} words, first at word {bad[0]} "
                                 f"(host {a[bad[0]]:#x}, device {b[bad[0]]:#x}); word size {w}")
    return hs, ds


@pytest.mark.parametrize("name", SCENES)
def test_golden_scenes_same_tables_with_given_and_derived_ranks(gpu, name):
    sc = golden_scene(name)
    _assert_same_tables(gpu, sc, name + " (ranks given)")
    _assert_same_tables(gpu, _NoRank(sc), name + " (ranks derived)")


@pytest.mark.parametrize("n,kind", [(1, "boxes"), (2, "boxes"), (3, "boxes"), (14, "mixed"), (15, "boxes"), (40, "mixed"), (1022, "boxes"),
                                    (1023, "mixed"), (3000, "mixed"), (20000, "boxes"), (262144, "mixed")])
def test_synthetic_scenes_same_tables(gpu, n, kind):
    from tools.gen_synth import make_scene
    sc = make_scene(n, 3, kind, 64, 64, 2)   # n + 2 surfaces, no ranks given
    hs, ds = _assert_same_tables(gpu, sc, f"synthetic {kind} {n}")
    assert ds.n_fast_nodes == hs.n_fast_nodes and ds.levels_ref > 0


def _lattice(n_side, dup=1):
    """boxes on an exact integer lattice (equal centres along every axis), each repeated `dup` times"""
    from plutracer_b200.scene import scene_from_tables
    from tools.gen_synth import camera_for, make_tables
    t = make_tables(1, 1, "boxes")
    g = np.arange(n_side, dtype=np.float32) * 2 - n_side
    c = np.stack(np.meshgrid(g, g, g, indexing="ij"), -1).reshape(-1, 3)
    c = np.repeat(c, dup, axis=0)
    n = c.shape[0]
    data = np.zeros((n, 12), np.float32)
    data[:, 0:3] = c - 0.5
    data[:, 4:7] = c + 0.5
    t["prim_data"] = np.concatenate([data, t["prim_data"][1:]])
    t["prim_tag"] = np.full(n + 2, _ffi.PRIM_BOX, np.int32)
    t["prim_material"] = np.concatenate([np.arange(n, dtype=np.int32) % 8, t["prim_material"][1:]])
    t["lights"][1].prim = n + 1
    return scene_from_tables(camera_for((0, 0, -140), (0, 0, 0), 64, 64), width=64, height=64, n_strata=2, **t)


@pytest.mark.parametrize("n_side,dup", [(2, 1), (4, 1), (12, 1), (3, 40)])
def test_equal_centres_are_decided_like_the_host_sort(gpu, n_side, dup):
    sc = _lattice(n_side, dup)
    hs, ds = _assert_same_tables(gpu, sc, f"lattice {n_side}^3 x{dup}")
    if n_side ** 3 * dup > 16:
        assert ds.host_tie_segments > 0


def test_device_leaf_order_equals_the_reference_live_tree_on_ties(gpu):
    """lattice.pluscene carries the ranks of the reference's own bvh_tree (equal centres everywhere): the device
    build, given no ranks, must arrive at the same ones."""
    sc = golden_scene("lattice")
    ref_rank = np.ctypeslib.as_array(sc.desc.prim_tie_rank, shape=(sc.n_prims,)).copy()
    bs = gpu.upload(_NoRank(sc), build="device")
    meta = gpu.download_table(_ffi.TABLE_PRIM_META).view(np.int32).reshape(-1, 4)
    assert bs.built_on_device == 1 and bs.host_tie_segments > 0
    assert np.array_equal(meta[:, 3], ref_rank)
    _assert_same_tables(gpu, sc, "lattice (ranks given)")


def test_renders_through_both_builds_are_identical(gpu):
    from tools.gen_synth import make_scene
    sc = make_scene(5000, 7, "mixed", 96, 64, 2)
    imgs = []
    for build in ("host", "device"):
        gpu.upload(sc, build=build)
        for flags in (_ffi.RENDER_BVH_FAST, _ffi.RENDER_BVH_REFERENCE):
            img = np.zeros((64, 96, 3), np.float32)
            gpu.render(GpuContext.params(96, 64, 2, seed=11, flags=flags), img)
            imgs.append(img)
    for im in imgs[1:]:
        assert np.array_equal(imgs[0].view(np.uint32), im.view(np.uint32))


def test_build_argument_is_checked(gpu):
    sc = golden_scene("cornell-box")
    bs = _ffi.BuildStats()
    rc = gpu._lib.plu_gpu_upload_scene_ex(gpu._h, sc.desc_ptr, 7, C.byref(bs))
    assert rc == -1   # PLU_E_ARG
    gpu.upload(sc)
