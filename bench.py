#!/usr/bin/env python
"""bench.py — Msamples/s of the render path on cornell-box.urn at 1024x1024, 4096 spp (BASELINE.json
configs[1]) on N B200s, with the reference's own CPU renderer timed beside it.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

One step = one full render of the named configuration (1024x1024 pixels x 4096 samples per pixel =
2^32 paths; --spp-per-step renders a slice of the strata instead).  At N > 1 (torchrun) rank r renders
the strata s = r (mod N) of every pixel into its own device framebuffer and the framebuffers are
combined by ONE NCCL reduce inside the timed region (strong scaling: the job is fixed).

value  = samples/s with everything resident in HBM (plu_gpu_render_device + reduce, CUDA events on the
         launching stream, max over ranks).
e2e    = the same through the reference-facing call renderer(...).render(target) with HOST buffers:
         scene upload, framebuffer H2D (`+=` semantics), render, framebuffer D2H inside the timed region.
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

SCENE = "cornell-box"
W, H, NSTRATA = 1024, 1024, 64
SEED = 20261104
L2_BYTES = 126 * 2 ** 20


def env_int(name, default):
    try:
        return int(os.environ.get(name, default))
    except ValueError:
        return default


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


def measured_traffic(kernel):
    """dram__bytes_read + dram__bytes_write per launch of `kernel` from the committed `ncu --set full` capture
    (profiles/traffic.json, written by tools/ncu_traffic.py from the .ncu-rep of this same command)."""
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            e = json.load(f)[kernel]
        extra = {k: e[k] for k in ("issue_active_pct", "threads_per_instruction", "warps_active_pct", "dram_throughput_pct") if k in e}
        return float(e["dram_bytes_per_launch"]), e.get("source", "profiles/traffic.json"), extra
    except Exception:
        return None, None, {}


# ---- the reference's own CPU renderer (oracle/_ref/plu_ref = unmodified reference sources) -----------------
def run_reference_cpu(smp, timeout=900):
    """renderer::render of the compiled reference on the host cores; returns (samples/s, info)."""
    from tests.orc import GOLDEN  # noqa: F401  (test infrastructure import is allowed in this leg only)
    ref_bin = os.path.join(ROOT, "oracle", "_ref", "plu_ref")
    scene = os.path.join(ROOT, "tests", "golden", SCENE + ".pluscene")
    if not (os.path.exists(ref_bin) and os.path.exists(scene)):
        return None
    # render-flat: the reference's own classes rebuilt from the flat fixture, then its unmodified renderer::render
    out = subprocess.run([ref_bin, "render-flat", scene, str(W), str(H), str(smp), "-"], capture_output=True, text=True, timeout=timeout)
    if out.returncode != 0:
        return None
    info = json.loads(out.stdout.strip().splitlines()[-1])
    return info["samples"] / (info["ms"] * 1e-3), info


def run_oracle_cpu(smp):
    """fallback CPU arm: the oracle port (oracle/oracle_pt.cpp), all host threads."""
    import numpy as np
    from plutracer_b200 import GpuContext
    from tests.orc import Oracle, golden_scene
    sc = golden_scene(SCENE, W, H, smp)
    p = GpuContext.params(W, H, smp, seed=SEED)
    t0 = time.perf_counter()
    _, st = Oracle().render(sc, p, out=np.zeros((H, W, 3), np.float32), trig=0, n_threads=0)
    dt = time.perf_counter() - t0
    return st.samples / dt, {"samples": int(st.samples), "ms": dt * 1e3, "threads": os.cpu_count()}


def cpu_arm(smp):
    r = run_reference_cpu(smp)
    kind = "reference"
    if r is None:
        r = run_oracle_cpu(smp); kind = "port"
    sps, info = r
    return {"value": sps / 1e6, "unit": "Msamples/s", "cores": int(info.get("threads", os.cpu_count() or 1)), "kind": kind,
            "sample": f"{SCENE}.urn {W}x{H} /smp {smp} = {info['samples']} samples of the same workload in {info['ms'] / 1e3:.1f} s"}


def main_reference(args):
    rank = env_int("RANK", 0)
    if rank != 0:
        return 0
    smp = 2   # 4.19 M samples per step: a bounded sample of the workload
    for _ in range(args.warmup):
        cpu_arm(1)
    vals = []
    t0 = time.perf_counter()
    for _ in range(args.steps):
        vals.append(cpu_arm(smp))
    dt = time.perf_counter() - t0
    v = statistics.mean(x["value"] for x in vals)
    base = dict(vals[-1]); base["value"] = v
    line = {"impl": "reference", "metric": "Msamples/s on cornell-box.urn at 1024x1024, 4096 spp", "value": v, "unit": "Msamples/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": dt / max(args.steps, 1) * 1e3,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"{SCENE}.urn {W}x{H} /smp {NSTRATA} (4096 spp); each step a bounded sample: /smp {smp}"},
            "cpu_baseline": base, "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line))
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--spp-per-step", type=int, default=NSTRATA * NSTRATA)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--strata-per-wave", type=int, default=0, help="paths in flight = W*H*this (0: library default)")
    args = ap.parse_args()
    if args.impl == "reference":
        return main_reference(args)

    import numpy as np
    import torch
    import torch.distributed as dist
    from plutracer_b200 import GpuContext, Scene, _ffi, renderer, shard_plan, sharded_renderer, stratified_sampler, texture2d

    rank, world, local = env_int("RANK", 0), env_int("WORLD_SIZE", 1), env_int("LOCAL_RANK", 0)
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: plutracer_b200 has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # stdout carries the JSON line only
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    warm = max(args.warmup, 3)

    scene = Scene.load_flat(os.path.join(ROOT, "tests", "golden", SCENE + ".pluscene")).override(W, H, NSTRATA)
    ctx = GpuContext(local)
    ctx.upload(scene)
    spp = args.spp_per_step
    assert 0 < spp <= NSTRATA * NSTRATA and spp >= world
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

    # ---- e2e: the reference-facing call with HOST buffers (every rank takes part when N > 1) ------------------
    tx = texture2d((W, H))
    fb_bytes = W * H * 3 * 4
    n_e2e = max(1, min(args.steps, 2))
    if world == 1:
        r = renderer(scene, None, (32, 32), stratified_sampler((W, H), (NSTRATA, NSTRATA)), None, device=local, seed=SEED)

        def e2e_step():   # scene upload + framebuffer H2D (`+=`) + render + framebuffer D2H, all inside plu_gpu_render
            tx.pixels.fill(0)
            if spp == NSTRATA * NSTRATA:
                return r.render(tx)
            p = GpuContext.params(W, H, NSTRATA, seed=SEED, sample_begin=0, sample_stride=1, sample_count=spp)
            r.ctx.upload(scene)
            return r.ctx.render(p, tx.pixels)
        h2d, d2h = fb_bytes + 1024, fb_bytes
    else:
        r = sharded_renderer(scene, None, (32, 32), stratified_sampler((W, H), (NSTRATA, NSTRATA)), None, device=local, rank=rank, world=world, seed=SEED)

        def e2e_step():   # rank 0: pinned host pixels -> device; all: render own strata; NCCL reduce; rank 0: device -> host
            tx.pixels.fill(0)
            return r.render(tx, spp=spp)
        h2d, d2h = fb_bytes, fb_bytes
    e2e_step()
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(n_e2e):
        e2e_step()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = (time.perf_counter() - t0) / n_e2e
    te = torch.tensor([dt], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    dt = float(te.item())
    e2e = {"value": samples_per_step / dt / 1e6, "unit": "Msamples/s", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "ms_per_step": dt * 1e3, "timing": "host wall clock around the call, max over ranks"}
    r.ctx.close()

    line = None
    if rank == 0:
        # ---- roofline of the dominant kernel, measured live (profiled pass, CUDA events between launches) ----
        peak, peak_src = measured_peak()
        pp = GpuContext.params(W, H, NSTRATA, seed=SEED, flags=_ffi.RENDER_PROFILE, sample_begin=0, sample_stride=1, sample_count=min(spp, 256))
        fb.zero_()
        ps = ctx.render_device(pp, fb.data_ptr(), stream.cuda_stream)
        pc = GpuContext.params(W, H, NSTRATA, seed=SEED, flags=_ffi.RENDER_COUNT_BVH, sample_begin=0, sample_stride=1, sample_count=16)
        cs = ctx.render_device(pc, fb.data_ptr(), stream.cuda_stream)
        rays = cs.rays_extend + cs.rays_shadow
        # B_bvh per ray = 32 B per interior node visited + tested primitives (sphere 16, box 32, triangle 48);
        # cornell-box: 6 boxes + 2 spheres -> mean 28 B per primitive test
        b_bvh = (32.0 * cs.bvh_nodes_visited + 28.0 * cs.prims_tested) / max(rays, 1)
        alg = {   # algorithmic bytes per kernel class (DESIGN.md "algorithmic bytes")
            "k_extend": ps.rays_extend * (48.0 + b_bvh),
            "k_shade": ps.shade_vertices * 144.0 + ps.rays_shadow * 48.0,
            "k_connect": ps.rays_shadow * (48.0 + b_bvh),
        }
        ms = {"k_extend": ps.ms_extend, "k_shade": ps.ms_shade, "k_connect": ps.ms_connect}
        dom = max(ms, key=ms.get)
        n_dom = int(ps.waves) * 8 * (2 if dom == "k_shade" else 1)   # launches of that kernel class in the profiled pass
        achieved = alg[dom] / (ms[dom] * 1e-3) / 1e9 if ms[dom] > 0 else 0.0
        bytes_per_sample = (alg["k_extend"] + alg["k_shade"] + alg["k_connect"]) / max(ps.samples, 1) + 12.0
        # the scene of this workload is staged in shared memory: its BVH bytes never touch HBM, so the whole-path HBM
        # figure leaves them out (the SURVEY.md 8d formula, which counts them, is reported beside it)
        hbm_per_sample = (ps.rays_extend * 48.0 + alg["k_shade"] + ps.rays_shadow * 48.0) / max(ps.samples, 1) + 12.0
        traffic, traffic_src, ncu_extra = measured_traffic(dom)
        roofline = {"bound": "hbm", "kernel": dom, "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                    "traffic_source": traffic_src, "ncu": ncu_extra,   # issue-slot / lane / occupancy figures of the same capture
                    "algorithmic_bytes_per_launch": alg[dom] / max(n_dom, 1),
                    "peak_source": peak_src, "launches_profiled": int(n_dom), "avg_launch_ms": ms[dom] / max(n_dom, 1),
                    "ms_by_kernel": {"k_generate": ps.ms_generate, **ms, "k_accumulate": ps.ms_accumulate},
                    "algorithmic_bytes_per_sample": hbm_per_sample, "b_bvh_per_ray_on_chip": b_bvh,
                    "survey_formula_bytes_per_sample_incl_on_chip_bvh": bytes_per_sample,
                    "whole_path_achieved_gbs": hbm_per_sample * value / world * 1e6 / 1e9,   # per GPU, HBM bytes only
                    "whole_path_frac": hbm_per_sample * value / world * 1e6 / 1e9 / peak,
                    "note": "kernel durations come from a serial profiled pass (one event per launch); the timed steps keep several "
                            "waves in flight on separate streams, so `value` is higher than the sum of these durations implies"}
        cpu = None if (args.no_cpu_baseline or world > 1) else cpu_arm(4)
        line = {"metric": "Msamples/s on cornell-box.urn at 1024x1024, 4096 spp", "value": value, "unit": "Msamples/s", "n_gpus": world,
                "steps": args.steps, "warmup": warm, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {"workload": f"{SCENE}.urn {W}x{H} /smp {NSTRATA}: {spp} spp per step ({samples_per_step} paths), <=8 vertices, NEE+MIS",
                           "parallelism": f"spp-sharded x{world}, one NCCL reduce of the float framebuffer per step" if world > 1 else "single GPU",
                           "l2": "flushed between timed steps (2x126 MiB write); path state per wave (~0.75 GB) exceeds L2",
                           "seed": SEED},
                "mrays_per_s": value * rays_per_sample, "rays_per_sample": rays_per_sample,
                "gpu_launches": launches, "clocks": clocks, "roofline": roofline}
        line["e2e"] = e2e
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
