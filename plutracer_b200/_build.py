"""Build libplugpu.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

The flags that matter for parity with the reference's x86-64 build (no FMA, IEEE div/sqrt, no
flush-to-zero) are part of the recipe, not an option: see csrc/device/pmath.cuh.

One measured option changes them for ONE translation unit: PLU_SHADE_FAST=1 builds csrc/shade_tu.cu
(k_shade: the BSDF / light / texture / MIS arithmetic between two scene->hit calls) with -fmad=true
-prec-div=false -prec-sqrt=false into a second library, libplugpu_shadefast.so (selected at run time with
PLU_LIB=shadefast).  Intersection, ray set-up, camera and accumulation keep the exact flags in both.
The default library is the exact one; every bit-exact parity test runs on it.
"""
import os
import shutil
import subprocess
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "_lib")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")


def lib_path(flavour=None):
    flavour = os.environ.get("PLU_LIB", "") if flavour is None else flavour
    return os.path.join(LIB_DIR, f"libplugpu_{flavour}.so" if flavour else "libplugpu.so")   # PLU_LIB=shadefast, or an A/B build


LIB_PATH = lib_path()

SOURCES = ["plugpu.cu", "shade_tu.cu", "bvh_gpu.cu", "host/scene_store.cpp", "host/bvh_build.cpp", "host/urn_loader.cpp"]
EXACT_FP = ["-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false"]
FAST_FP = ["-fmad=true", "-prec-div=false", "-prec-sqrt=false", "-ftz=false", "-DPLU_SHADE_FAST=1"]
COMMON = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    *(["-DPLU_LEAF_SMALL=" + os.environ["PLU_LEAF_SMALL"]] if os.environ.get("PLU_LEAF_SMALL") else []),
    *(["-DPLU_SSTK=" + os.environ["PLU_SSTK"]] if os.environ.get("PLU_SSTK") else []),
    *(["-DPLU_ANYHIT=" + os.environ["PLU_ANYHIT"]] if os.environ.get("PLU_ANYHIT") else []),
    *(["-DPLU_PREFETCH_FAR=" + os.environ["PLU_PREFETCH_FAR"]] if os.environ.get("PLU_PREFETCH_FAR") else []),
    *(["-DPLU_PT_BLOCKS=" + os.environ["PLU_PT_BLOCKS"]] if os.environ.get("PLU_PT_BLOCKS") else []),
    *(["-DPLU_SEED_AREA=" + os.environ["PLU_SEED_AREA"]] if os.environ.get("PLU_SEED_AREA") else []),
    *(["-DPLU_SHADE_BLOCKS=" + os.environ["PLU_SHADE_BLOCKS"]] if os.environ.get("PLU_SHADE_BLOCKS") else []),
    *(["-DPLU_UNIFORM_UNROLL=" + os.environ["PLU_UNIFORM_UNROLL"]] if os.environ.get("PLU_UNIFORM_UNROLL") else []),
    *(["-DPLU_STAGED_WARP_SCOPE=" + os.environ["PLU_STAGED_WARP_SCOPE"]] if os.environ.get("PLU_STAGED_WARP_SCOPE") else []),
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-Xcompiler", "-pthread",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libplugpu.so cannot be built (there is no CPU fallback)")


def _stale(path):
    if not os.path.exists(path):
        return True
    t = os.path.getmtime(path)
    deps = [INCLUDE + "/plugpu.h", os.path.abspath(__file__)]
    for root, _, files in os.walk(CSRC):
        deps += [os.path.join(root, f) for f in files]
    return any(os.path.getmtime(d) > t for d in deps)


def build_lib(force=False, verbose=False, shade_fast=None):
    """Compile csrc/ into plutracer_b200/_lib/libplugpu.so (shade_fast: libplugpu_shadefast.so); returns its path."""
    if shade_fast is None:
        shade_fast = os.environ.get("PLU_SHADE_FAST", "") not in ("", "0")
    out = lib_path("shadefast" if shade_fast else "")
    if not force and not _stale(out):
        return out
    os.makedirs(LIB_DIR, exist_ok=True)
    obj_dir = os.path.join(LIB_DIR, "obj_shadefast" if shade_fast else "obj")
    os.makedirs(obj_dir, exist_ok=True)
    nvcc = _nvcc()
    logs = []

    def compile_one(src):
        fp = FAST_FP if (shade_fast and src == "shade_tu.cu") else EXACT_FP
        obj = os.path.join(obj_dir, src.replace("/", "_") + ".o")
        cmd = [nvcc] + COMMON + fp + ["-I" + INCLUDE, "-I" + CSRC, "-c", os.path.join(CSRC, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
        logs.append(res.stderr)
        return obj

    with ThreadPoolExecutor(len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    res = subprocess.run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-pthread", "-o", out + ".tmp"] + objs,
                         capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("link failed:\n" + res.stdout + res.stderr)
    os.replace(out + ".tmp", out)
    if verbose:
        print("".join(logs))
    return out


if __name__ == "__main__":
    print(build_lib(force=True, verbose=True))
