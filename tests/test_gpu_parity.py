"""GPU parity tests: the CUDA path through the C ABI against the golden vectors of the reference and
against the CPU oracle on the same Philox stream.  Bit-exact unless stated."""
import os

import numpy as np
import pytest

from tests.conftest import REF_SCENES, SCENES
from tests.orc import GOLDEN, golden_scene
from tests.conftest import record_parity
from tests.metrics import check_converged as _check_converged, converged_metrics as _converged_metrics
from tests.test_oracle import bits, same_bits

pytestmark = pytest.mark.gpu

REF, FAST = 2, 8   # PLU_RENDER_BVH_REFERENCE, PLU_RENDER_BVH_FAST


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("mode", [REF, FAST])
def test_primary_hits_match_reference(gpu, name, mode):
    """north_star check 1: hit surface ids of deterministic pinhole rays, bit-exact (t and normal too)."""
    g = np.load(os.path.join(GOLDEN, name + ".golden.npz"))
    w, h = (int(v) for v in g["primary_wh"])
    gpu.upload(golden_scene(name, w, h))
    ids, t, n = gpu.trace_primary(w, h, mode)
    assert np.array_equal(ids, g["primary_id"])
    assert np.array_equal(bits(t), bits(g["primary_t"]))
    assert np.array_equal(bits(n), bits(g["primary_n"].reshape(-1, 3)))


@pytest.mark.parametrize("name", SCENES)
@pytest.mark.parametrize("mode", [REF, FAST])
def test_arbitrary_rays_match_reference(gpu, oracle, name, mode):
    """axis-aligned, zero, non-unit directions and origins on surfaces (no-epsilon self hits)."""
    g = np.load(os.path.join(GOLDEN, name + ".golden.npz"))
    sc = golden_scene(name)
    gpu.upload(sc)
    ids, t, n = gpu.trace_rays(g["rays_o"], g["rays_d"], mode)
    keep = np.ones(len(ids), bool)
    if mode == FAST:
        # the one documented deviation of the fast BVH: rays on which the reference's own
        # origin-inflated node boxes compute 0 * inf (DESIGN.md "fast BVH")
        keep = ~oracle.nan_slab_rays(sc, g["rays_o"], g["rays_d"])
        assert keep.mean() > 0.9
    assert np.array_equal(ids[keep], g["rays_id"][keep]), np.nonzero(ids[keep] != g["rays_id"][keep])
    assert same_bits(t[keep], g["rays_t"][keep]).all() and same_bits(n[keep], g["rays_n"][keep]).all()


@pytest.mark.parametrize("name", SCENES)
def test_per_sample_radiance_matches_oracle(gpu, oracle, name):
    """ray_color of every sample of a 48x36 /smp 3 image on the shared Philox stream: bit-exact."""
    from plutracer_b200 import GpuContext
    w, h, n = 48, 36, 3
    sc = golden_scene(name, w, h, n)
    gpu.upload(sc)
    p = GpuContext.params(w, h, n, seed=1234567)
    got = gpu.sample_radiance(p, 0, w * h)
    want = oracle.sample_radiance(sc, p, 0, w * h, trig=0)
    eq = same_bits(got, want)
    assert eq.all(), f"{(~eq).sum()} of {eq.size} components differ; max abs diff {np.nanmax(np.abs(got - want))}"
    assert np.nansum(np.abs(want)) > 0


@pytest.mark.parametrize("name", ["cornell-box", "room"])
def test_image_matches_oracle_bit_exact(gpu, oracle, name):
    """renderer::render: accumulation order, += into existing pixels, the overflow-column quirk, several
    waves (strata_per_wave=1) and a sharded sample range."""
    from plutracer_b200 import GpuContext
    w, h, n = 40, 30, 4
    sc = golden_scene(name, w, h, n)
    gpu.upload(sc)
    for flags, kw in ((0, {}), (1, {}), (0, dict(strata_per_wave=1)), (0, dict(sample_begin=1, sample_stride=4, sample_count=4))):
        p = GpuContext.params(w, h, n, seed=77, flags=flags, **kw)
        base = np.full((h, w, 3), 0.25, np.float32)
        got = base.copy()
        st = gpu.render(p, got)
        want, ost = oracle.render(sc, p, out=base.copy(), trig=0, n_threads=4)
        assert same_bits(got, want).all()
        assert st.samples == ost.samples and st.rays_extend == ost.rays_extend
        assert 0 < st.rays_shadow <= ost.rays_shadow   # zero-radiance MIS rays are not cast on the GPU


def test_fast_bvh_equals_reference_walk_in_full_paths(gpu):
    from plutracer_b200 import GpuContext
    w, h, n = 64, 48, 2
    gpu.upload(golden_scene("cornell-box", w, h, n))
    a = gpu.sample_radiance(GpuContext.params(w, h, n, seed=5, flags=REF), 0, w * h)
    b = gpu.sample_radiance(GpuContext.params(w, h, n, seed=5, flags=FAST), 0, w * h)
    assert same_bits(a, b).all()


def test_sample_sharding_is_linear(gpu):
    """multi-GPU decomposition: the image is the sum of its sample shards (float summation order aside)."""
    from plutracer_b200 import GpuContext
    w, h, n = 64, 64, 4
    gpu.upload(golden_scene("cornell-box", w, h, n))
    full = np.zeros((h, w, 3), np.float32)
    gpu.render(GpuContext.params(w, h, n, seed=9), full)
    parts = np.zeros((h, w, 3), np.float32)
    for g in range(4):
        gpu.render(GpuContext.params(w, h, n, seed=9, sample_begin=g, sample_stride=4, sample_count=4), parts)
    ok = np.isfinite(full)
    assert np.allclose(full[ok], parts[ok], rtol=2e-5, atol=1e-6)


def test_errors(gpu):
    from plutracer_b200 import GpuContext, _ffi
    gpu.upload(golden_scene("minimal0"))
    with pytest.raises(_ffi.PluError):
        gpu.render(GpuContext.params(0, 10, 2), np.zeros((10, 1, 3), np.float32)[:, :0])
    with pytest.raises(_ffi.PluError):
        gpu.sample_radiance(GpuContext.params(8, 8, 2, sample_begin=0, sample_stride=0, sample_count=3), 0, 4)


@pytest.mark.parametrize("name", REF_SCENES)
def test_converged_image_matches_reference_renders(gpu, name):
    """1024 spp at 128x96 against two independent-seed runs of the reference's multithreaded renderer::render
    (overflow-column quirk reproduced with PLU_RENDER_COMPAT_OVERFLOW_COLUMN so the full frame compares)."""
    from plutracer_b200 import GpuContext
    g = np.load(os.path.join(GOLDEN, name + ".converged.npz"))
    w, h, n = (int(v) for v in g["cfg"])
    gpu.upload(golden_scene(name, w, h, n))
    img = np.zeros((h, w, 3), np.float32)
    gpu.render(GpuContext.params(w, h, n, seed=424242, flags=1), img)
    m = _converged_metrics(img, g["a"], g["b"])
    record_parity(f"gpu:{name}:{w}x{h}:/smp{n}", m)
    _check_converged(m, w * h, name)


def test_cornell_box_4096spp_matches_reference(gpu):
    """BASELINE configs[1] at its sample count: cornell-box, 4096 spp (/smp 64), converged at 256x256 because the
    reference needs 13 min per render at that size on 8 cores (2.7 h at 1024x1024)."""
    from plutracer_b200 import GpuContext
    g = np.load(os.path.join(GOLDEN, "cornell-box.converged256.npz"))
    w, h, n = (int(v) for v in g["cfg"])
    gpu.upload(golden_scene("cornell-box", w, h, n))
    img = np.zeros((h, w, 3), np.float32)
    st = gpu.render(GpuContext.params(w, h, n, seed=20261104, flags=1), img)
    m = _converged_metrics(img, g["a"], g["b"])
    record_parity(f"gpu:cornell-box:{w}x{h}:/smp{n}", m)
    assert st.samples == w * h * n * n + (h - 1) * n * n
    _check_converged(m, w * h, "cornell-box")


def test_cornell_box_1024x1024_4096spp_matches_reference(gpu):
    """BASELINE configs[1] AT ITS SIZE (north_star's target): cornell-box with the 1024x1024 camera, 4096 spp, against
    the reference's own renders of eight 32-row bands (a quarter of the image; two independent seeds, plu_ref
    render-st, tests/golden/make_golden.py --bands1024).  The whole 2^32-path image is rendered through the C ABI;
    the statistics are taken on the bands."""
    from plutracer_b200 import GpuContext
    g = np.load(os.path.join(GOLDEN, "cornell-box.converged1024.npz"))
    w, h, n = (int(v) for v in g["cfg"])
    rows = int(g["rows"])
    gpu.upload(golden_scene("cornell-box", w, h, n))
    img = np.zeros((h, w, 3), np.float32)
    st = gpu.render(GpuContext.params(w, h, n, seed=20261104), img)
    assert st.samples == w * h * n * n
    mine = np.concatenate([img[y0:y0 + rows] for y0 in g["y0"]])
    a, b = g["a"].reshape(-1, w, 3), g["b"].reshape(-1, w, 3)
    m = _converged_metrics(mine, a, b)
    m["per_band_ratio_clipped"] = []
    for k, y0 in enumerate(g["y0"]):
        mk = _converged_metrics(img[y0:y0 + rows], g["a"][k], g["b"][k])
        m["per_band_ratio_clipped"].append(mk["rmse_gpu_clipped"] / mk["rmse_floor_clipped"])
    record_parity(f"gpu:cornell-box:{w}x{h}:/smp{n}:bands", m)
    _check_converged(m, mine.shape[0] * w, "cornell-box")


def test_cornell_box_eight_gpu_seeds_against_eight_reference_seeds(gpu):
    """The bound is not a constant but the reference's own spread (tests/golden/make_noise_floor.py).  Eight GPU seeds
    against eight independent renders of the unmodified reference (128x96, 1024 spp): every (GPU image; reference pair)
    statistic lies inside the range the reference renders produce against each other, and the ensemble means agree
    per channel within north_star's 1 % (one render alone scatters by ~0.7 %, the reference against itself up to 1.3 %)."""
    import itertools
    import json
    from plutracer_b200 import GpuContext
    from tests.metrics import noise_floor
    g = np.load(os.path.join(GOLDEN, "cornell-box.seeds.npz"))
    w, h, n = (int(v) for v in g["cfg"])
    nf = noise_floor("cornell-box")
    refs = g["imgs"]
    gpu.upload(golden_scene("cornell-box", w, h, n))
    mine = []
    for seed in range(8):
        img = np.zeros((h, w, 3), np.float32)
        gpu.render(GpuContext.params(w, h, n, seed=1000 + seed), img)
        mine.append(img)
    rc, rp = [], []
    for img in mine:
        for i, j in itertools.combinations(range(len(refs)), 2):
            m = _converged_metrics(img, refs[i], refs[j])
            rc.append(m["rmse_gpu_clipped"] / m["rmse_floor_clipped"]); rp.append(m["rmse_gpu"] / m["rmse_floor"])
    ref_mean = np.nanmean(refs, axis=0)
    cap = 4.0 * float(np.median(ref_mean[ref_mean > 0]))   # the clip level of tests/metrics.py
    ok = np.isfinite(refs).all(axis=(0, 3)) & np.isfinite(np.array(mine)).all(axis=(0, 3))
    mean_ref = np.array([np.minimum(r, cap)[ok].mean(axis=0) for r in refs])      # [8, 3] clipped image means
    mean_gpu = np.array([np.minimum(r, cap)[ok].mean(axis=0) for r in mine])
    ens_err = np.abs(mean_gpu.mean(axis=0) - mean_ref.mean(axis=0)) / mean_ref.mean(axis=0)
    plain_err = np.abs(np.array([r[ok].mean(axis=0) for r in mine]).mean(axis=0) - np.array([r[ok].mean(axis=0) for r in refs]).mean(axis=0)) / np.array([r[ok].mean(axis=0) for r in refs]).mean(axis=0)
    record_parity(f"gpu:cornell-box:{w}x{h}:/smp{n}:8-gpu-seeds-vs-8-reference-seeds",
                  {"ratio_clipped": {"min": min(rc), "median": float(np.median(rc)), "max": max(rc), "n": len(rc)},
                   "ratio_plain": {"min": min(rp), "median": float(np.median(rp)), "max": max(rp)},
                   "ensemble_mean_err_clipped": ens_err.tolist(), "ensemble_mean_err_plain": plain_err.tolist(),
                   "single_render_scatter_clipped_ref": (mean_ref.std(axis=0) / mean_ref.mean(axis=0)).tolist(),
                   "single_render_scatter_clipped_gpu": (mean_gpu.std(axis=0) / mean_gpu.mean(axis=0)).tolist(),
                   "reference_triples": nf})
    assert max(rc) <= nf["ratio_clipped"]["max"] * 1.03 and min(rc) >= nf["ratio_clipped"]["min"] * 0.97
    assert abs(float(np.median(rc)) - nf["ratio_clipped"]["median"]) < 0.01
    assert ens_err.max() < 0.01


# ---- synthetic large scene (BASELINE configs[4]): the fast BVH against the reference's own tree ---------------
@pytest.fixture(scope="module")
def synth():
    from tools.gen_synth import make_scene
    return make_scene(20000, seed=3, kind="mixed", width=192, height=192, n_strata=2)


def test_synth_fast_bvh_primary_equals_reference_tree(gpu, oracle, synth):
    """20 002 boxes + triangles: tight SAH tree with culling == exhaustive walk of the reference's origin-inflated
    median-split tree (bit-exact id, t, normal), and == the CPU oracle on a subset of the rays."""
    gpu.upload(synth)
    w, h = 192, 192
    a = gpu.trace_primary(w, h, REF)
    b = gpu.trace_primary(w, h, FAST)
    assert (a[0] >= 0).mean() > 0.3
    assert np.array_equal(a[0], b[0]) and np.array_equal(bits(a[1]), bits(b[1])) and np.array_equal(bits(a[2]), bits(b[2]))
    ids, t, n, o, d = oracle.primary_with_rays(synth, w, h)   # oracle = reference walk on the CPU
    assert np.array_equal(ids, b[0]) and np.array_equal(bits(t), bits(b[1]))
    # secondary-like rays: random origins inside the lattice, random + axis-aligned + non-unit directions
    rng = np.random.default_rng(5)
    ro = rng.uniform(-55, 55, (20000, 3)).astype(np.float32)
    rd = rng.normal(size=(20000, 3)).astype(np.float32)
    rd /= np.linalg.norm(rd, axis=1, keepdims=True)
    rd[:2000] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 2000)] * rng.uniform(0.2, 2.0, (2000, 1)).astype(np.float32)
    rd[2000:4000] *= rng.uniform(0.2, 2.0, (2000, 1)).astype(np.float32)
    x = gpu.trace_rays(ro, rd, REF)
    y = gpu.trace_rays(ro, rd, FAST)
    keep = ~oracle.nan_slab_rays(synth, ro, rd)
    assert keep.mean() > 0.99
    assert np.array_equal(x[0][keep], y[0][keep]) and same_bits(x[1][keep], y[1][keep]).all()
    ids, t, n = oracle.trace_rays(synth, ro[:3000], rd[:3000])
    assert np.array_equal(ids, x[0][:3000]) and same_bits(t, x[1][:3000]).all()


def test_synth_paths_fast_equals_reference_tree_and_oracle(gpu, oracle, synth):
    from plutracer_b200 import GpuContext
    gpu.upload(synth)
    w, h, n = 192, 192, 2
    p_ref = GpuContext.params(w, h, n, seed=11, flags=REF)
    p_fast = GpuContext.params(w, h, n, seed=11, flags=FAST)
    first = 192 * 90
    a = gpu.sample_radiance(p_ref, first, 192 * 6)
    b = gpu.sample_radiance(p_fast, first, 192 * 6)
    assert same_bits(a, b).all()
    assert np.nansum(a) > 0
    want = oracle.sample_radiance(synth, p_fast, first, 96, trig=0)
    assert same_bits(b[:96], want).all()


def test_lane_refill_schedule_changes_nothing_but_the_order(gpu, synth):
    """k_extend_pt (persistent warps that refill idle lanes, the default on fast-BVH scenes) against k_extend (one
    ray per lane per iteration, PLU_RENDER_NO_LANE_REFILL): same image to the bit, same ray and vertex counts."""
    from plutracer_b200 import GpuContext, _ffi
    gpu.upload(synth)
    w, h, n = 160, 120, 3
    out = []
    for flags in (FAST, FAST | _ffi.RENDER_NO_LANE_REFILL, FAST | _ffi.RENDER_COUNT_BVH, FAST | _ffi.RENDER_COUNT_BVH | _ffi.RENDER_NO_LANE_REFILL):
        img = np.zeros((h, w, 3), np.float32)
        st = gpu.render(GpuContext.params(w, h, n, seed=5, flags=flags, strata_per_wave=2), img)
        out.append((img, st))
    for img, st in out[1:]:
        assert same_bits(out[0][0], img).all()
        assert (st.samples, st.rays_extend, st.rays_shadow, st.shade_vertices, st.shadow_entries) == \
               (out[0][1].samples, out[0][1].rays_extend, out[0][1].rays_shadow, out[0][1].shade_vertices, out[0][1].shadow_entries)
    assert out[2][1].bvh_nodes_visited == out[3][1].bvh_nodes_visited > 0 and out[2][1].prims_tested == out[3][1].prims_tested > 0
    assert np.nansum(out[0][0]) > 0


@pytest.mark.parametrize("name", ["cornell-box", "glass0"])
def test_waves_in_flight_do_not_change_the_image(lib, name, monkeypatch):
    """plu_gpu_render_device keeps up to four waves in flight on four streams (and the specular shade kernel on a
    side stream); the framebuffer adds are chained in wave order, so the image must equal the serial schedule
    (PLU_PIPES=1, PLU_NO_FORK) bit for bit — including the accumulation into a non-zero target."""
    from plutracer_b200 import GpuContext
    sc = golden_scene(name, 64, 48, 4)
    imgs = []
    for pipes, no_fork in (("4", False), ("1", True), ("3", False)):
        monkeypatch.setenv("PLU_PIPES", pipes)
        if no_fork:
            monkeypatch.setenv("PLU_NO_FORK", "1")
        else:
            monkeypatch.delenv("PLU_NO_FORK", raising=False)
        ctx = GpuContext(0)
        try:
            ctx.upload(sc)
            img = np.full((48, 64, 3), 0.25, np.float32)
            st = ctx.render(GpuContext.params(64, 48, 4, seed=3, strata_per_wave=1), img)   # 16 waves
            assert st.waves == 16
            imgs.append(img)
        finally:
            ctx.close()
    assert same_bits(imgs[0], imgs[1]).all() and same_bits(imgs[0], imgs[2]).all()
    assert np.nansum(imgs[0]) > 0.25 * imgs[0].size


# ---- scenes with spheres on the fast BVH: sphere::hit is fed non-unit directions (SURVEY.md 8d config 5 as worded) ----
@pytest.fixture(scope="module")
def synth_spheres():
    from tools.gen_synth import make_scene
    return make_scene(20000, seed=4, kind="spheres", width=160, height=160, n_strata=2)


def _sphere_rays(rng, n, extent):
    """origins in and around the lattice; unit, shortened (|d| 0.2..1) and stretched (|d| 1..2.5) directions"""
    ro = rng.uniform(-extent, extent, (n, 3)).astype(np.float32)
    rd = rng.normal(size=(n, 3)).astype(np.float32)
    rd /= np.linalg.norm(rd, axis=1, keepdims=True)
    k = n // 5
    rd[k:2 * k] *= rng.uniform(0.2, 1.0, (k, 1)).astype(np.float32)
    rd[2 * k:4 * k] *= rng.uniform(1.0, 2.5, (2 * k, 1)).astype(np.float32)
    rd[4 * k:4 * k + k // 2] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, k // 2)] * rng.uniform(0.3, 2.0, (k // 2, 1)).astype(np.float32)
    return ro, rd


def test_sphere_scene_general_walk_equals_reference_tree_on_rays(gpu, oracle, synth_spheres):
    """20 002 surfaces, 20 000 of them spheres: the pruned walk (cone criterion + reference ancestor check) against the
    exhaustive walk of the reference's own tree, bit-exact (id, t, normal), on 100 000 rays of which 80 % are non-unit
    (for |d| > 1 sphere::hit accepts spheres far off the ray; for |d| < 1 it misses spheres the ray goes through)."""
    gpu.upload(synth_spheres)
    rng = np.random.default_rng(21)
    ro, rd = _sphere_rays(rng, 100000, 60.0)
    x = gpu.trace_rays(ro, rd, REF)
    y = gpu.trace_rays(ro, rd, 0)
    keep = ~oracle.nan_slab_rays(synth_spheres, ro, rd)
    assert keep.mean() > 0.99
    hit = x[0] >= 0
    assert 0.2 < hit.mean() < 0.999
    # stretched directions really do "hit" spheres the ray misses geometrically (what the pruned walk has to follow)
    assert np.array_equal(x[0][keep], y[0][keep]) and same_bits(x[1][keep], y[1][keep]).all() and same_bits(x[2][keep], y[2][keep]).all()
    ids, t, n = oracle.trace_rays(synth_spheres, ro[:2000], rd[:2000])
    assert np.array_equal(ids, y[0][:2000]) and same_bits(t, y[1][:2000]).all()
    # primary rays
    a, b = gpu.trace_primary(160, 160, REF), gpu.trace_primary(160, 160, 0)
    assert np.array_equal(a[0], b[0]) and same_bits(a[1], b[1]).all() and (a[0] >= 0).mean() > 0.3


def test_sphere_scene_full_paths_equal_reference_tree(gpu, oracle, synth_spheres):
    from plutracer_b200 import GpuContext
    gpu.upload(synth_spheres)
    w, h, n = 160, 160, 2
    first = 160 * 70
    a = gpu.sample_radiance(GpuContext.params(w, h, n, seed=13, flags=REF), first, 160 * 8)
    b = gpu.sample_radiance(GpuContext.params(w, h, n, seed=13), first, 160 * 8)
    assert same_bits(a, b).all() and np.nansum(a) > 0
    want = oracle.sample_radiance(synth_spheres, GpuContext.params(w, h, n, seed=13), first, 64, trig=0)
    assert same_bits(b[:64], want).all()
    img_ref, img = np.zeros((h, w, 3), np.float32), np.zeros((h, w, 3), np.float32)
    s1 = gpu.render(GpuContext.params(w, h, n, seed=13, flags=REF), img_ref)
    s2 = gpu.render(GpuContext.params(w, h, n, seed=13), img)
    assert same_bits(img_ref, img).all()
    assert (s1.rays_extend, s1.rays_shadow, s1.shade_vertices) == (s2.rays_extend, s2.rays_shadow, s2.shade_vertices)


def test_sphere_scene_200k_rays(gpu, oracle):
    """200 002 surfaces (the size at which the exhaustive walk visited 5453 nodes per ray): 20 000 rays, both walks."""
    from tools.gen_synth import make_scene
    sc = make_scene(200000, seed=5, kind="spheres", width=64, height=64, n_strata=1)
    gpu.upload(sc)
    rng = np.random.default_rng(22)
    ro, rd = _sphere_rays(rng, 20000, 60.0)
    x = gpu.trace_rays(ro, rd, REF)
    y = gpu.trace_rays(ro, rd, 0)
    keep = ~oracle.nan_slab_rays(sc, ro, rd)
    assert np.array_equal(x[0][keep], y[0][keep]) and same_bits(x[1][keep], y[1][keep]).all()
    assert (x[0] >= 0).mean() > 0.2


@pytest.mark.parametrize("name", ["glass0", "room"])
def test_graph_replay_and_wave_size_do_not_change_the_image(lib, name, monkeypatch):
    """One CUDA graph launch per wave (default), the while-loop graph (PLU_GRAPH_LOOP), direct launches (PLU_NO_GRAPH),
    fixed and in-call adapted wave sizes: bit-identical images and identical ray counts."""
    from plutracer_b200 import GpuContext
    w, h, n = 640, 480, 16   # 256 strata x 307 200 pixels = 78.6 M paths: 20 waves of 2^22, fewer once the probe enlarges them
    sc = golden_scene(name, w, h, n)
    out = []
    for env in ({}, {"PLU_GRAPH_LOOP": "1"}, {"PLU_NO_GRAPH": "1"}, {"PLU_NO_ADAPTIVE_WAVES": "1"}):
        for k in ("PLU_GRAPH_LOOP", "PLU_NO_GRAPH", "PLU_NO_ADAPTIVE_WAVES"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        ctx = GpuContext(0)
        try:
            ctx.upload(sc)
            img = np.full((h, w, 3), 0.5, np.float32)
            st = ctx.render(GpuContext.params(w, h, n, seed=17), img)
            out.append((img, st))
        finally:
            ctx.close()
    for img, st in out[1:]:
        assert same_bits(out[0][0], img).all()
        assert (st.samples, st.rays_extend, st.rays_shadow, st.shade_vertices) == (out[0][1].samples, out[0][1].rays_extend, out[0][1].rays_shadow, out[0][1].shade_vertices)
    assert out[0][1].waves < out[3][1].waves   # the adapted render used larger waves


def test_fast_walks_equal_the_reference_tree_on_the_million_surface_scenes(gpu, oracle):
    """BASELINE configs[4] at ITS size (10^6 surfaces): the exhaustive walk of the reference's own tree
    (PLU_RENDER_BVH_REFERENCE: ~2 x 10^6 nodes per ray) against the shipped walks — the SAH tree for boxes + triangles, the
    pruned reference tree for spheres — on camera rays and on rays from inside the lattice incl. non-unit directions.
    The count of unmasked rays that differ must be 0 (masked: rays with d_k == 0 whose origin lies on a plane of one of the
    reference's node boxes, 0 * inf — counted)."""
    from tools.gen_synth import make_scene
    rng = np.random.default_rng(31)
    n_rays = 3000
    ro = rng.uniform(-55, 55, (n_rays, 3)).astype(np.float32)
    rd = rng.normal(size=(n_rays, 3)).astype(np.float32)
    rd /= np.linalg.norm(rd, axis=1, keepdims=True)
    rd[:600] *= rng.uniform(0.3, 2.0, (600, 1)).astype(np.float32)                                   # non-unit
    rd[600:900] = np.eye(3, dtype=np.float32)[rng.integers(0, 3, 300)] * rng.choice([-1.0, 1.0], (300, 1)).astype(np.float32)   # axis-aligned
    ro[900:1200] = np.array([0, 0, -140], np.float32)                                               # from the camera
    rd[900:1200] = rd[900:1200] * 0.2 + np.array([0, 0, 1], np.float32)
    rd[900:1200] /= np.linalg.norm(rd[900:1200], axis=1, keepdims=True)
    out = {}
    for kind in ("mixed", "spheres"):
        sc = make_scene(1000000, seed=1, kind=kind, width=64, height=64, n_strata=1)
        gpu.upload(sc)
        x = gpu.trace_rays(ro, rd, REF)
        y = gpu.trace_rays(ro, rd, 0)
        keep = ~oracle.nan_slab_rays(sc, ro, rd)
        differ = int(((x[0] != y[0]) | ~same_bits(x[1], y[1]))[keep].sum())
        out[kind] = {"rays": n_rays, "masked": int((~keep).sum()), "hits": int((x[0] >= 0).sum()), "unmasked_rays_that_differ": differ}
        assert keep.mean() > 0.98 and (x[0] >= 0).mean() > 0.3
        assert differ == 0, out
    record_parity("gpu:synthetic-1M:REF-vs-shipped-walk", out)


@pytest.mark.parametrize("name", ["cornell-box", "room", "refrac0"])
def test_rays_on_interior_node_planes_take_the_exact_walk(gpu, oracle, name):
    """0 * inf at an INTERIOR node of the reference tree (origin on a plane of the node's box, direction component 0 on that
    axis): the staged uniform walk only notes the NaN and repeats the walk in glm's select order — same (surface, t) as the
    CPU oracle, bit for bit, as for every other ray of the batch."""
    sc = golden_scene(name)
    gpu.upload(sc)
    ro, rd = _interior_plane_rays(oracle, sc)
    assert oracle.nan_slab_rays(sc, ro, rd).mean() > 0.9     # these rays do compute 0 * inf in the reference's boxes
    ids, t, n = gpu.trace_rays(ro, rd, REF)
    wi, wt, wn = oracle.trace_rays(sc, ro, rd)
    assert np.array_equal(ids, wi) and same_bits(t, wt).all() and same_bits(n, wn).all()
    assert (ids >= 0).any() and (ids < 0).any()


def _interior_plane_rays(oracle, sc):
    lro, b6 = oracle.bvh_nodes(sc)
    interior = b6[lro[:, 2] < 0]
    assert len(interior) >= 1
    rng = np.random.default_rng(17)
    ro, rd = [], []
    for nd in interior:
        lo, hi = nd[0:3], nd[3:6]
        for axis in range(3):
            for plane in (lo[axis], hi[axis]):
                for _ in range(6):
                    o = (lo + (hi - lo) * rng.uniform(0.05, 0.95, 3)).astype(np.float32)
                    o[axis] = plane
                    d = rng.normal(size=3).astype(np.float32)
                    d[axis] = 0.0
                    if rng.random() < 0.3:
                        d[(axis + 1) % 3] = 0.0   # two zero components
                    ro.append(o); rd.append(d / max(np.linalg.norm(d), 1e-6))
    return np.array(ro, np.float32), np.array(rd, np.float32)
