"""Build libplugpu.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

The flags that matter for parity with the reference's x86-64 build (no FMA, IEEE div/sqrt, no
flush-to-zero) are part of the recipe, not an option: see csrc/device/pmath.cuh.
"""
import os
import shutil
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB_DIR = os.path.join(HERE, "_lib")
LIB_PATH = os.path.join(LIB_DIR, "libplugpu.so")
INCLUDE = os.path.join(os.path.dirname(HERE), "include")

SOURCES = ["plugpu.cu", "bvh_gpu.cu", "host/scene_store.cpp", "host/bvh_build.cpp", "host/urn_loader.cpp"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "--threads", "4",
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    *(["-DPLU_LEAF_SMALL=" + os.environ["PLU_LEAF_SMALL"]] if os.environ.get("PLU_LEAF_SMALL") else []),
    *(["-DPLU_SSTK=" + os.environ["PLU_SSTK"]] if os.environ.get("PLU_SSTK") else []),
    *(["-DPLU_ANYHIT=" + os.environ["PLU_ANYHIT"]] if os.environ.get("PLU_ANYHIT") else []),
    "-Xcompiler", "-fPIC", "-Xcompiler", "-ffp-contract=off", "-Xcompiler", "-pthread", "-shared",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libplugpu.so cannot be built (there is no CPU fallback)")


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [INCLUDE + "/plugpu.h"]
    for root, _, files in os.walk(CSRC):
        deps += [os.path.join(root, f) for f in files]
    return any(os.path.getmtime(d) > t for d in deps)


def build_lib(force=False, verbose=False):
    """Compile csrc/ into plutracer_b200/_lib/libplugpu.so; returns its path."""
    if not force and not _stale():
        return LIB_PATH
    os.makedirs(LIB_DIR, exist_ok=True)
    srcs = [os.path.join(CSRC, s) for s in SOURCES if os.path.exists(os.path.join(CSRC, s))]
    cmd = [_nvcc()] + NVCC_FLAGS + ["-I" + INCLUDE, "-I" + CSRC, "-o", LIB_PATH + ".tmp"] + srcs
    if verbose:
        cmd.insert(1, "-Xptxas=-v")
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    os.replace(LIB_PATH + ".tmp", LIB_PATH)
    if verbose:
        print(res.stderr)
    return LIB_PATH


if __name__ == "__main__":
    print(build_lib(force=True, verbose=True))
