#!/usr/bin/env python
"""bench.py — Msamples/s of the render path on N B200s, with the reference's own CPU renderer timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--config NAME|all]

Default workload (the one BASELINE.json's metric is quoted on): cornell-box.urn at 1024x1024, 4096 spp
(configs[1]).  --config selects another BASELINE.json configuration (same JSON contract, one line each):

    minimal0      scenes/minimal0.urn     512x512    /smp 8   (configs[0], the reference's own CPU-runnable case)
    cornell-box   scenes/cornell-box.urn  1024x1024  /smp 64  (configs[1], default)
    glass0        scenes/glass0.urn       1280x960   /smp 8   (configs[2])
    refrac0       scenes/refrac0.urn      1280x960   /smp 8   (configs[2])
    room          scenes/room.urn         1920x1080  /smp 32  (configs[3])
    synth-boxes   10^6 boxes              1024x1024  /smp 8   (configs[4], tools/gen_synth.py)
    synth-mixed   10^6 boxes + triangles  1024x1024  /smp 8   (configs[4])
    synth-spheres 10^6 spheres            1024x1024  /smp 8   (configs[4] as SURVEY.md 8d words it)
    all           every line above, one process group, one JSON line per configuration

One step = one full render of the configuration (--spp-per-step renders a slice of the strata instead).  At N > 1
(torchrun) rank r renders the strata s = r (mod N) of every pixel into its own device framebuffer and the
framebuffers are combined by ONE NCCL reduce inside the timed region (strong scaling: the job is fixed).

value  = samples/s with everything resident in HBM (plu_gpu_render_device + reduce, CUDA events on the
         launching stream, max over ranks).
e2e    = the same through the reference-facing call with HOST buffers: scene upload, framebuffer H2D (`+=`
         semantics), render, framebuffer D2H inside the timed region.  N = 1: renderer(...).render(target);
         N > 1: ONE process driving all N GPUs through the C ABI (plu_gpu_create_multi: the path the C++ drop-in
         integration/plu_gpu_renderer.h uses, NCCL reduce owned by the context), run on rank 0 while the other ranks
         wait; the one-process-per-GPU host (sharded_renderer over torch.distributed) is reported as e2e_torchrun.
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

SEED = 20261104
L2_BYTES = 126 * 2 ** 20
DEFAULT_CONFIG = "cornell-box"

# name -> scene source, image size, strata; cpu = (W, H, /smp) of the bounded CPU sample (None: the full configuration)
CONFIGS = {
    "minimal0":      dict(scene="minimal0", w=512, h=512, n=8, cpu=None),
    "cornell-box":   dict(scene="cornell-box", w=1024, h=1024, n=64, cpu=(1024, 1024, 4)),
    "glass0":        dict(scene="glass0", w=1280, h=960, n=8, cpu=None),
    "refrac0":       dict(scene="refrac0", w=1280, h=960, n=8, cpu=None),
    "room":          dict(scene="room", w=1920, h=1080, n=32, cpu=(1920, 1080, 4)),
    "synth-boxes":   dict(synth="boxes", prims=1000000, w=1024, h=1024, n=8, cpu=(96, 96, 1)),
    "synth-mixed":   dict(synth="mixed", prims=1000000, w=1024, h=1024, n=8, cpu=(96, 96, 1)),
    "synth-spheres": dict(synth="spheres", prims=1000000, w=1024, h=1024, n=8, cpu=(96, 96, 1)),
}


_JSON_FD = None


def emit(line):
    """One JSON line on the process's ORIGINAL stdout.  Everything else that writes to fd 1 (NCCL's version banner, a
    library's printf) is sent to stderr for the whole run (main() re-points fd 1), so stdout carries the JSON lines only."""
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)


def config_names(arg):
    names = list(CONFIGS) if arg == "all" else [x for x in arg.split(",") if x]
    for x in names:
        if x not in CONFIGS:
            raise SystemExit(f"unknown --config {x!r}: one of {', '.join(CONFIGS)}, all")
    return names


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


def describe(name):
    c = CONFIGS[name]
    if "scene" in c:
        return f"{c['scene']}.urn at {c['w']}x{c['h']}, {c['n'] ** 2} spp"
    return f"synthetic {c['prims']}-surface {c['synth']} scene at {c['w']}x{c['h']}, {c['n'] ** 2} spp"


def data_label(name):
    c = CONFIGS[name]
    if "scene" in c:
        return (f"named scene: tests/golden/{c['scene']}.pluscene = the reference's scenes/{c['scene']}.urn flattened by its own loader "
                "(no dataset; Philox sample stream)")
    return f"synthetic: tools/gen_synth.py make_scene({c['prims']}, seed 1, '{c['synth']}')"


def load_scene(name):
    from plutracer_b200 import Scene
    c = CONFIGS[name]
    if "scene" in c:
        return Scene.load_flat(os.path.join(ROOT, "tests", "golden", c["scene"] + ".pluscene")).override(c["w"], c["h"], c["n"])
    from tools.gen_synth import make_scene
    return make_scene(c["prims"], 1, c["synth"], c["w"], c["h"], c["n"])


class ClockSampler:
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.rows, self.proc, self.index = [], None, index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._read, daemon=True).start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                self.proc.kill()
        sm, mx, reasons = [], [], set()
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx.append(float(r[1]))
            except (ValueError, IndexError):
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), r[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs, burst copy)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


def measured_traffic(config, kernel):
    """dram__bytes_read + dram__bytes_write per launch of `kernel` from the committed `ncu --set full` capture of this
    configuration (profiles/traffic.json for the default workload, profiles/traffic_<config>.json otherwise; written by
    tools/ncu_traffic.py from the .ncu-rep of this same command)."""
    try:
        fn = "traffic.json" if config == DEFAULT_CONFIG else f"traffic_{config}.json"
        with open(os.path.join(ROOT, "profiles", fn)) as f:
            t = json.load(f)
        # large scenes extend / connect with the persistent-warp kernels: the capture lists them under their own names
        e = t[kernel] if kernel in t else t[kernel + "_pt"]
        keys = ("issue_active_pct", "threads_per_instruction", "warps_active_pct", "dram_throughput_pct", "l2_read_gbs", "dram_read_gbs", "l1_hit_pct", "l2_hit_pct")
        return float(e["dram_bytes_per_launch"]), e.get("source", "profiles/" + fn), {k: e[k] for k in keys if k in e}
    except Exception:
        return None, None, {}


# ---- the reference's own CPU renderer (oracle/_ref/plu_ref = unmodified reference sources) -----------------
_flat_cache = {}


def flat_fixture(name):
    """Path of a .pluscene of the configuration (named scenes: the committed fixture; synthetic: written to /tmp)."""
    c = CONFIGS[name]
    if "scene" in c:
        return os.path.join(ROOT, "tests", "golden", c["scene"] + ".pluscene")
    if name not in _flat_cache:
        import tempfile
        path = os.path.join(tempfile.gettempdir(), f"plu_bench_{name}.pluscene")
        load_scene(name).save_flat(path)
        _flat_cache[name] = path
    return _flat_cache[name]


def run_reference_cpu(name, w, h, smp, timeout=3000):
    """renderer::render of the compiled reference on the host cores; returns (samples/s, info)."""
    from tests.orc import GOLDEN  # noqa: F401  (test infrastructure import is allowed in this leg only)
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "plu_ref")
    if not os.path.exists(ref_bin):
        return None
    scene = flat_fixture(name)
    # render-flat: the reference's own classes rebuilt from the flat tables, its own bvh_tree, then its unmodified renderer::render
    t0 = time.perf_counter()
    out = subprocess.run([ref_bin, "render-flat", scene, str(w), str(h), str(smp), "-"], capture_output=True, text=True, timeout=timeout)
    if out.returncode != 0:
        return None
    info = json.loads(out.stdout.strip().splitlines()[-1])
    info["process_s"] = time.perf_counter() - t0   # incl. the reference's own scene + bvh_tree construction
    return info["samples"] / (info["ms"] * 1e-3), info


def run_oracle_cpu(name, w, h, smp):
    """fallback CPU arm: the oracle port (oracle/oracle_pt.cpp), all host threads."""
    import numpy as np
    from plutracer_b200 import GpuContext
    from tests.orc import Oracle
    sc = load_scene(name).override(w, h, smp)
    p = GpuContext.params(w, h, smp, seed=SEED)
    t0 = time.perf_counter()
    _, st = Oracle().render(sc, p, out=np.zeros((h, w, 3), np.float32), trig=0, n_threads=0)
    dt = time.perf_counter() - t0
    return st.samples / dt, {"samples": int(st.samples), "ms": dt * 1e3, "threads": os.cpu_count()}


def cpu_arm(name, sample=None):
    c = CONFIGS[name]
    w, h, smp = sample if sample else (c["cpu"] if c["cpu"] else (c["w"], c["h"], c["n"]))
    r = run_reference_cpu(name, w, h, smp)
    kind = "reference"
    if r is None:
        r = run_oracle_cpu(name, w, h, smp); kind = "port"
    sps, info = r
    full = (w, h, smp) == (c["w"], c["h"], c["n"])
    what = "the full configuration" if full else "a bounded sample of the same workload (a rate: same scene, camera and path loop)"
    label = f"{c['scene']}.urn" if "scene" in c else f"synthetic {c['synth']} n={c['prims']}"
    return {"value": sps / 1e6, "unit": "Msamples/s", "cores": int(info.get("threads", os.cpu_count() or 1)), "kind": kind,
            "sample": f"{label} {w}x{h} /smp {smp} = {info['samples']} samples in {info['ms'] / 1e3:.1f} s of renderer::render: {what}"}


def main_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    for name in config_names(args.config):
        c = CONFIGS[name]
        # a bounded sample per step: ~4 M paths of the same workload (synthetic: 16 K primary samples, the reference's
        # tree makes every ray visit thousands of nodes)
        sample = (c["w"], c["h"], 2) if "scene" in c else c["cpu"]
        for _ in range(args.warmup):
            cpu_arm(name, (c["w"], c["h"], 1) if "scene" in c else (32, 32, 1))
        vals = []
        t0 = time.perf_counter()
        for _ in range(args.steps):
            vals.append(cpu_arm(name, sample))
        dt = time.perf_counter() - t0
        v = statistics.mean(x["value"] for x in vals)
        base = dict(vals[-1]); base["value"] = v
        line = {"impl": "reference", "metric": "Msamples/s on " + describe(name), "value": v, "unit": "Msamples/s",
                "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / max(args.steps, 1) * 1e3,
                "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": data_label(name),
                "config": {"workload": f"{describe(name)}; each step a bounded sample: {sample[0]}x{sample[1]} /smp {sample[2]}", "name": name},
                "cpu_baseline": base, "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
        print(json.dumps(line), flush=True)
    return 0


def prim_mix_bytes(scene):
    """mean bytes of primitive data per tested surface (sphere 16, box 32, triangle 48), by scene composition"""
    import numpy as np
    n = scene.n_prims
    if n <= 0:
        return 32.0
    tags = np.ctypeslib.as_array(scene.desc.prim_tag, shape=(n,))
    cnt = np.bincount(tags, minlength=3).astype(np.float64)
    return float((cnt[0] * 16.0 + cnt[1] * 32.0 + cnt[2] * 48.0) / n)


def has_specular(scene):
    """any perfect-reflection / perfect-refraction / glass material: k_shade then runs once per material class and bounce"""
    d = scene.desc
    return any(1 <= d.materials[i].tag <= 3 for i in range(d.n_materials))


def bench_config(name, args, rank, world, local, host_group):
    import numpy as np  # noqa: F401
    import torch
    import torch.distributed as dist
    from plutracer_b200 import GpuContext, _ffi, renderer, shard_plan, sharded_renderer, stratified_sampler, texture2d

    c = CONFIGS[name]
    W, H, NSTRATA = c["w"], c["h"], c["n"]
    warm = max(args.warmup, 3)
    scene = load_scene(name)
    ctx = GpuContext(local)
    t_up = time.perf_counter()
    bs = ctx.upload(scene)
    t_up = time.perf_counter() - t_up
    spp = NSTRATA * NSTRATA if args.spp_per_step <= 0 else min(args.spp_per_step, NSTRATA * NSTRATA)
    assert spp >= world, "fewer samples per pixel than ranks"
    # rank r takes strata r, r + world, ... of the first `spp` strata (plutracer_b200/shard.py)
    s_begin, s_stride, s_count = shard_plan(NSTRATA, world, rank, spp)
    params = GpuContext.params(W, H, NSTRATA, seed=SEED, sample_begin=s_begin, sample_stride=s_stride, sample_count=s_count, strata_per_wave=args.strata_per_wave)
    fb = torch.zeros(H, W, 3, dtype=torch.float32, device="cuda")
    flush = torch.empty(2 * L2_BYTES, dtype=torch.uint8, device="cuda")
    stream = torch.cuda.current_stream()

    def step():
        fb.zero_()
        st = ctx.render_device(params, fb.data_ptr(), stream.cuda_stream)
        if world > 1:
            dist.reduce(fb, dst=0, op=dist.ReduceOp.SUM)
        return st

    for _ in range(warm):
        st = step()
    sampler = ClockSampler(local)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    total_ms, launches = 0.0, 0
    for _ in range(args.steps):
        flush.fill_(1)   # L2 flush between timed iterations (path state >> L2 anyway)
        e0.record(stream)
        st = step()
        e1.record(stream)
        torch.cuda.synchronize()
        total_ms += e0.elapsed_time(e1)
        launches += int(st.kernel_launches)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    t = torch.tensor([total_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms = float(t.item())
    ms_per_step = total_ms / args.steps
    samples_per_step = W * H * spp
    value = samples_per_step / (ms_per_step * 1e-3) / 1e6
    rays_per_sample = (st.rays_extend + st.rays_shadow) / max(st.samples, 1)

    # ---- e2e: the reference-facing call with HOST buffers ----------------------------------------------------------
    fb_bytes = W * H * 3 * 4
    n_e2e = max(1, min(args.steps, 2))
    # one zero-filled target per call, as main.cpp:175 makes a fresh texture2d per image (render accumulates, `+=`); the pixels are
    # page-locked host memory (plu_host_alloc), which is what the e2e contract asks the host side of the copies to be
    targets = [texture2d((W, H), pinned=True) for _ in range(n_e2e + 1)]
    next_target = iter(targets)
    full = spp == NSTRATA * NSTRATA
    e2e_torchrun = None

    first_call = {}

    def timed(fn, collective):
        t_first = time.perf_counter()
        fn()   # first call: allocations, scene tables copied to the device(s) and traversal structures built
        torch.cuda.synchronize()
        first_call["ms"] = (time.perf_counter() - t_first) * 1e3
        if collective:
            dist.barrier()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for _ in range(n_e2e):
            fn()
        torch.cuda.synchronize()
        if collective:
            dist.barrier()
        dt = (time.perf_counter() - t0) / n_e2e
        if collective:
            te = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(te, op=dist.ReduceOp.MAX)
            dt = float(te.item())
        return dt

    if world == 1:
        r = renderer(scene, None, (32, 32), stratified_sampler((W, H), (NSTRATA, NSTRATA)), None, device=local, seed=SEED)

        def e2e_step():   # scene upload + framebuffer H2D (`+=`) + render + framebuffer D2H, all inside plu_gpu_render
            tx = next(next_target)
            if full:
                return r.render(tx)
            p = GpuContext.params(W, H, NSTRATA, seed=SEED, sample_begin=0, sample_stride=1, sample_count=spp)
            r.ctx.upload(scene)
            return r.ctx.render(p, tx.pixels)
        dt = timed(e2e_step, False)
        r.ctx.close()
        e2e = {"value": samples_per_step / dt / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": fb_bytes, "d2h_bytes_per_step": fb_bytes,
               "ms_per_step": dt * 1e3, "first_call_ms": first_call["ms"],
               "timing": "host wall clock around renderer(...).render(target): plu_gpu_upload_scene (the scene is unchanged between steps: its tables are "
                         "compared with those on the device and kept; first_call_ms is the call that copied and built them) + framebuffer H2D + render + D2H; "
                         "the target is a fresh zero texture2d per call in page-locked host memory (plu_host_alloc)",
               "path": "C ABI, one device"}
    else:
        r = sharded_renderer(scene, None, (32, 32), stratified_sampler((W, H), (NSTRATA, NSTRATA)), None, device=local, rank=rank, world=world, seed=SEED)

        def e2e_step():   # rank 0: pinned host pixels -> device; all: render own strata; NCCL reduce; rank 0: device -> host
            return r.render(next(next_target), spp=spp)
        dt = timed(e2e_step, True)
        r.ctx.close()
        e2e_torchrun = {"value": samples_per_step / dt / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": fb_bytes, "d2h_bytes_per_step": fb_bytes,
                        "ms_per_step": dt * 1e3, "timing": "host wall clock around sharded_renderer.render, max over ranks",
                        "path": "one process per GPU, torch.distributed NCCL reduce"}
        # the drop-in path: ONE process (rank 0) drives all N GPUs through plu_gpu_create_multi; the other ranks release
        # their wave buffers and wait on a host-side (gloo) barrier so that no NCCL kernel of theirs spins on the GPUs
        ctx.close(); ctx = None
        torch.cuda.synchronize()
        dist.barrier(group=host_group)
        e2e = None
        if rank == 0:
            m = GpuContext(list(range(world)))
            p = GpuContext.params(W, H, NSTRATA, seed=SEED, sample_begin=0, sample_stride=1, sample_count=spp)
            ms_red = []

            next_target = iter(targets)
            for t in targets:
                t.pixels.fill(0)

            def multi_step():
                tx = next(next_target)
                m.upload(scene)
                s = m.render(p, tx.pixels)
                ms_red.append(s.ms_reduce)
                return s
            dt = timed(multi_step, False)
            m.close()
            e2e = {"value": samples_per_step / dt / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": fb_bytes, "d2h_bytes_per_step": fb_bytes,
                   "ms_per_step": dt * 1e3, "first_call_ms": first_call["ms"], "ms_reduce": ms_red[-1],
                   "timing": "host wall clock around plu_gpu_upload_scene (unchanged scene: table compare per device; first_call_ms copied and built) + plu_gpu_render on a multi-device context",
                   "path": f"C ABI, ONE process, plu_gpu_create_multi over {world} devices, context-owned ncclReduce"}
        dist.barrier(group=host_group)
        if rank == 0:
            ctx = GpuContext(local); ctx.upload(scene)   # for the profiled passes below

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel, measured live (profiled pass, CUDA events between launches) ----
        peak, peak_src = measured_peak()
        prof_spp = min(spp, 256 if "scene" in c else 16)
        pp = GpuContext.params(W, H, NSTRATA, seed=SEED, flags=_ffi.RENDER_PROFILE, sample_begin=0, sample_stride=1, sample_count=prof_spp)
        fb.zero_()
        ps = ctx.render_device(pp, fb.data_ptr(), stream.cuda_stream)
        pc = GpuContext.params(W, H, NSTRATA, seed=SEED, flags=_ffi.RENDER_COUNT_BVH, sample_begin=0, sample_stride=1, sample_count=min(spp, 16))
        cs = ctx.render_device(pc, fb.data_ptr(), stream.cuda_stream)
        rays = cs.rays_extend + cs.rays_shadow
        on_chip = scene.n_prims <= 64   # the scene blob is staged in shared memory (plugpu.cu: blob_in_smem / fast_threshold)
        # B_bvh per ray: bytes of interior node per visit + bytes of surface data per test.  Reference-tree
        # scenes: 32-B nodes; fast-BVH scenes: the 64-B two-child node actually fetched per visit
        node_bytes = 32.0 if on_chip else 64.0
        b_bvh = (node_bytes * cs.bvh_nodes_visited + prim_mix_bytes(scene) * cs.prims_tested) / max(rays, 1)
        alg = {   # algorithmic bytes per kernel class (DESIGN.md "algorithmic bytes")
            "k_extend": ps.rays_extend * (48.0 + b_bvh),
            "k_shade": ps.shade_vertices * 144.0 + ps.rays_shadow * 48.0,
            "k_connect": ps.rays_shadow * (48.0 + b_bvh),
        }
        ms = {"k_extend": ps.ms_extend, "k_shade": ps.ms_shade, "k_connect": ps.ms_connect}
        dom = max(ms, key=ms.get)
        n_dom = int(ps.waves) * 8 * (2 if (dom == "k_shade" and has_specular(scene)) else 1)   # launches of that class in the profiled pass
        achieved = alg[dom] / (ms[dom] * 1e-3) / 1e9 if ms[dom] > 0 else 0.0
        bytes_per_sample = (alg["k_extend"] + alg["k_shade"] + alg["k_connect"]) / max(ps.samples, 1) + 12.0
        # a scene staged in shared memory never fetches its BVH from HBM, so the whole-path HBM figure leaves those bytes
        # out (the SURVEY.md 8d formula, which counts them, is reported beside it); large scenes fetch them through L2
        if on_chip:
            hbm_per_sample = (ps.rays_extend * 48.0 + alg["k_shade"] + ps.rays_shadow * 48.0) / max(ps.samples, 1) + 12.0
        else:
            hbm_per_sample = bytes_per_sample
        traffic, traffic_src, ncu_extra = measured_traffic(name, dom)
        trav_ms = ps.ms_extend + ps.ms_connect
        roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                    "traffic_source": traffic_src, "ncu": ncu_extra,   # issue-slot / lane / occupancy / L2 figures of the same capture
                    "algorithmic_bytes_per_launch": alg[dom] / max(n_dom, 1),
                    "peak_source": peak_src, "launches_profiled": int(n_dom), "avg_launch_ms": ms[dom] / max(n_dom, 1),
                    "ms_by_kernel": {"k_generate": ps.ms_generate, **ms, "k_accumulate": ps.ms_accumulate},
                    "algorithmic_bytes_per_sample": hbm_per_sample,
                    ("b_bvh_per_ray_on_chip" if on_chip else "b_bvh_per_ray"): b_bvh,
                    "nodes_per_ray": cs.bvh_nodes_visited / max(rays, 1), "surfaces_tested_per_ray": cs.prims_tested / max(rays, 1),
                    "bvh_fetch_achieved_gbs": (b_bvh * (ps.rays_extend + ps.rays_shadow)) / (trav_ms * 1e-3) / 1e9 if trav_ms > 0 else None,
                    "bvh_fetch_frac_of_hbm_peak": (b_bvh * (ps.rays_extend + ps.rays_shadow)) / (trav_ms * 1e-3) / 1e9 / peak if trav_ms > 0 else None,
                    "survey_formula_bytes_per_sample_incl_on_chip_bvh": bytes_per_sample,
                    "whole_path_achieved_gbs": hbm_per_sample * value / world * 1e6 / 1e9,   # per GPU
                    "whole_path_frac": hbm_per_sample * value / world * 1e6 / 1e9 / peak,
                    "note": "kernel durations come from a serial profiled pass (one event per launch); the timed steps keep several "
                            "waves in flight on separate streams, so `value` is higher than the sum of these durations implies"}
        cpu = None if (args.no_cpu_baseline or world > 1) else cpu_arm(name)
        par = f"spp-sharded x{world}, one NCCL reduce of the float framebuffer per step" if world > 1 else "single GPU"
        line = {"metric": "Msamples/s on " + describe(name), "value": value, "unit": "Msamples/s", "n_gpus": world,
                "steps": args.steps, "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": data_label(name),
                "config": {"workload": f"{describe(name).split(' at ')[0]} {W}x{H} /smp {NSTRATA}: {spp} spp per step ({samples_per_step} paths), <=8 vertices, NEE+MIS",
                           "name": name, "parallelism": par, "surfaces": int(scene.n_prims),
                           "l2": "flushed between timed steps (2x126 MiB write); path state per wave (~0.75 GB) exceeds L2",
                           "seed": SEED},
                "mrays_per_s": value * rays_per_sample, "rays_per_sample": rays_per_sample,
                "gpu_launches": launches, "clocks": clocks, "roofline": roofline,
                "scene_upload_ms": t_up * 1e3, "bvh_build_ms": bs.ms_total}
        line["e2e"] = e2e
        if e2e_torchrun is not None:
            line["e2e_torchrun"] = e2e_torchrun
        if cpu is not None:
            line["cpu_baseline"] = cpu
        emit(line)
    if ctx is not None:
        ctx.close()
    del fb, flush
    torch.cuda.empty_cache()
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--config", default=DEFAULT_CONFIG, help="a configuration name, a comma-separated list, or 'all' (see the module docstring)")
    ap.add_argument("--spp-per-step", type=int, default=0, help="render only the first k strata per step (0: the configuration's n^2)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strata-per-wave", type=int, default=0, help="paths in flight = W*H*this (0: library default)")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)

    global _JSON_FD
    sys.stdout.flush()
    _JSON_FD = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: plutracer_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    host_group = None
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the JSON lines only
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        host_group = dist.new_group(backend="gloo")
    for name in config_names(args.config):
        bench_config(name, args, rank, world, local, host_group)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
