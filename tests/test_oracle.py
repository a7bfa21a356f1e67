"""CPU tests: the oracle (oracle/oracle_pt.cpp) against the golden vectors produced by the
unmodified reference (tests/golden/make_golden.py).  Bit-exact: these pin the oracle."""
import os

import numpy as np
import pytest

from tests.conftest import SCENES
from tests.orc import GOLDEN, golden_scene


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def same_bits(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    return (bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))


def test_philox_known_answers(oracle):
    # Random123 kat_vectors, philox4x32 10 rounds
    assert oracle.philox([0, 0, 0, 0], [0, 0]) == [0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8]
    assert oracle.philox([0xffffffff] * 4, [0xffffffff] * 2) == [0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd]
    assert oracle.philox([0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344], [0xa4093822, 0x299f31d0]) == [0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1]


def test_det_trig_close_to_libm(oracle):
    x = np.linspace(0, 2 * np.pi * 1.01, 200001, dtype=np.float32)
    for which, f in ((0, np.sin), (1, np.cos)):
        err = np.abs(oracle.det_trig(which, x).astype(np.float64) - f(x.astype(np.float64)))
        assert err.max() < 2.5e-7
    a = np.linspace(-1, 1, 200001, dtype=np.float32)
    err = np.abs(oracle.det_trig(2, a).astype(np.float64) - np.arccos(a.astype(np.float64)))
    assert err.max() < 6e-7
    assert np.isnan(oracle.det_trig(2, np.array([1.0000001, -1.5, np.nan], np.float32))).all()


@pytest.mark.parametrize("name", SCENES)
def test_primary_hits_bit_exact(oracle, name):
    g = np.load(os.path.join(GOLDEN, name + ".golden.npz"))
    w, h = (int(v) for v in g["primary_wh"])
    sc = golden_scene(name, w, h)
    for trig in (0, 1):
        ids, t, n = oracle.trace_primary(sc, w, h, trig)
        assert np.array_equal(ids, g["primary_id"])
        assert np.array_equal(bits(t), bits(g["primary_t"]))
        assert np.array_equal(bits(n), bits(g["primary_n"]))


@pytest.mark.parametrize("name", SCENES)
def test_arbitrary_rays_bit_exact(oracle, name):
    g = np.load(os.path.join(GOLDEN, name + ".golden.npz"))
    sc = golden_scene(name)
    ids, t, n = oracle.trace_rays(sc, g["rays_o"], g["rays_d"], 1)
    assert np.array_equal(ids, g["rays_id"])
    assert same_bits(t, g["rays_t"]).all() and same_bits(n, g["rays_n"]).all()


@pytest.mark.parametrize("name", SCENES)
def test_per_sample_radiance_bit_exact_under_reference_rng(oracle, name):
    """ray_color for every sample of a 32x24 /smp 2 image, driven by the reference's mt19937 stream."""
    g = np.load(os.path.join(GOLDEN, name + ".golden.npz"))
    w, h, n, seed = (int(v) for v in g["st_cfg"])
    sc = golden_scene(name, w, h, n)
    img, ps = oracle.render_mt(sc, w, h, n, seed, 0, 0, w, h, trig=1, x_first=int(g["x_first"]))
    assert same_bits(ps, g["st_samples"]).all()
    assert same_bits(img, g["st_img"]).all()
    assert np.abs(g["st_samples"]).sum() > 0


@pytest.mark.parametrize("name", SCENES)
def test_leaf_order_matches_reference_tree(oracle, name):
    sc = golden_scene(name)
    ref_rank = np.ctypeslib.as_array(sc.desc.prim_tie_rank, shape=(sc.n_prims,))   # dumped from the live bvh_tree
    assert np.array_equal(oracle.leaf_order(sc), ref_rank)
    assert np.array_equal(sc.reference_leaf_order(), ref_rank)   # the product's own restatement of the build


def test_leaf_order_with_equal_centres_matches_reference_tree(oracle):
    """tests/golden/lattice.pluscene (make_lattice_golden.py): 1789 surfaces whose centres tie along every axis;
    its ranks were dumped from the reference's live bvh_tree, i.e. they carry std::sort's order of equal elements."""
    sc = golden_scene("lattice")
    ref_rank = np.ctypeslib.as_array(sc.desc.prim_tie_rank, shape=(sc.n_prims,))
    assert sorted(ref_rank.tolist()) == list(range(sc.n_prims))
    assert np.array_equal(oracle.leaf_order(sc), ref_rank)
    assert np.array_equal(sc.reference_leaf_order(), ref_rank)


@pytest.mark.parametrize("n_side,dup", [(4, 1), (12, 1), (3, 40)])
def test_leaf_order_with_equal_centres_oracle_vs_product(oracle, n_side, dup):
    """Equal box centres in segments of more than 16 surfaces: the order is whatever std::sort's introsort leaves
    (bvh_tree.cpp:23 is not a stable sort).  The oracle sorts surface ids through the comparator, the product
    sorts (centre, id) pairs: same algorithm, same comparator outcomes, so the same permutation."""
    from tests.test_device_build import _lattice
    sc = _lattice(n_side, dup)
    assert np.array_equal(oracle.leaf_order(sc), sc.reference_leaf_order())


@pytest.mark.parametrize("name", ["refrac0", "room"])
def test_philox_stream_converges_to_reference_renders(oracle, name):
    """Same path code, the Philox stream instead of mt19937: 1024 spp at 128x96 against the two
    independent-seed renders of the unmodified reference (the test the GPU path has to pass too)."""
    from plutracer_b200 import GpuContext
    from tests.metrics import check_converged, converged_metrics
    g = np.load(os.path.join(GOLDEN, name + ".converged.npz"))
    w, h, n = (int(v) for v in g["cfg"])
    sc = golden_scene(name, w, h, n)
    img, _ = oracle.render(sc, GpuContext.params(w, h, n, seed=424242, flags=1), trig=0)
    m = converged_metrics(img, g["a"], g["b"])
    from tests.conftest import record_parity
    record_parity(f"oracle:{name}:{w}x{h}:/smp{n}", m)
    check_converged(m, w * h, name)
