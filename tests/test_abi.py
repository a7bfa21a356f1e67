"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol that
include/plugpu.h declares, and the host-side scene plumbing works.  No compute calls."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from tests.orc import GOLDEN, ROOT, golden_scene


def test_header_symbols_all_exported(lib):
    from plutracer_b200 import _ffi
    hdr = open(os.path.join(ROOT, "include", "plugpu.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = set(re.findall(r"\b(plu_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 18
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in plugpu.h but not exported by libplugpu.so"
    assert declared == set(_ffi.SYMBOLS), declared ^ set(_ffi.SYMBOLS)
    assert lib.plu_abi_version() == _ffi.ABI_VERSION == 2


def test_struct_sizes_match_header(lib):
    from plutracer_b200 import _ffi
    assert C.sizeof(_ffi.CameraDesc) == 17 * 4
    assert C.sizeof(_ffi.MaterialDesc) == 13 * 4
    assert C.sizeof(_ffi.TextureDesc) == 12 * 4
    assert C.sizeof(_ffi.LightDesc) == 10 * 4
    assert C.sizeof(_ffi.RenderParams) == 48
    assert C.sizeof(_ffi.RenderStats) == 136


def test_no_cpu_fallback(lib):
    """Without a device plu_gpu_create must fail with PLU_E_CUDA (this test only runs its assertion on CPU boxes)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from plutracer_b200 import GpuContext, _ffi
    with pytest.raises(_ffi.PluError) as e:
        GpuContext(0)
    assert e.value.code == -2


def test_flat_scene_round_trip(lib, tmp_path):
    sc = golden_scene("cornell-box")
    assert sc.n_prims == 8 and sc.n_lights == 1 and sc.defaults == (512, 512, 8)
    p = tmp_path / "x.pluscene"
    sc.save_flat(p)
    assert open(p, "rb").read() == open(os.path.join(GOLDEN, "cornell-box.pluscene"), "rb").read()
    sc.override(1024, 256, 64)
    assert sc.defaults == (1024, 256, 64)
    assert sc.desc.cam.inv_image_size[0] == np.float32(1) / np.float32(1024)
    assert sc.desc.cam.inv_image_size[1] == np.float32(1) / np.float32(256)


def test_bad_inputs_rejected(lib, tmp_path):
    from plutracer_b200 import Scene, _ffi
    with pytest.raises(_ffi.PluError):
        Scene.load_flat(tmp_path / "missing.pluscene")
    bad = tmp_path / "bad.pluscene"
    bad.write_bytes(b"not a scene")
    with pytest.raises(_ffi.PluError):
        Scene.load_flat(bad)
    d = _ffi.SceneDesc()
    d.abi_version = 99
    assert lib.plu_scene_save_flat(C.byref(d), 1, 1, 1, b"/tmp/x") == -1


def test_host_memory_entry_points_reject_bad_arguments(lib):
    """plu_host_*: argument errors need no device; without one the allocation itself is refused (no pageable stand-in)."""
    import torch
    from plutracer_b200 import _ffi
    p = C.c_void_p()
    assert lib.plu_host_alloc(0, C.byref(p)) == _ffi.PLU_E_ARG and not p.value
    assert lib.plu_host_alloc(64, None) == _ffi.PLU_E_ARG
    assert lib.plu_host_register(None, 64) == _ffi.PLU_E_ARG
    assert lib.plu_host_unregister(None) == _ffi.PLU_E_ARG
    assert lib.plu_host_free(None) == _ffi.PLU_OK
    if not torch.cuda.is_available():
        assert lib.plu_host_alloc(64, C.byref(p)) == _ffi.PLU_E_CUDA and not p.value
        from plutracer_b200 import texture2d
        with pytest.raises(_ffi.PluError):
            texture2d((4, 4), pinned=True)
        assert not texture2d((4, 4), pinned=None).pinned


def test_header_is_plain_c_and_links_against_the_library(lib, tmp_path):
    """include/plugpu.h is the drop-in boundary: it must compile as C99 (no C++ or torch types in the signatures) and a C
    program must link against libplugpu.so and call a host-side entry point."""
    import shutil
    import subprocess
    from plutracer_b200 import _ffi
    gcc = shutil.which("gcc")
    if not gcc:
        pytest.skip("no gcc")
    src = tmp_path / "c_abi.c"
    src.write_text('#include <stdio.h>\n#include "plugpu.h"\n'
                   'int main(void) { plu_scene* s = 0; char err[64]; int rc = plu_scene_load_flat("/nonexistent.pluscene", &s, err, 64);\n'
                   '  printf("%u %d %lld\\n", plu_abi_version(), rc, (long long)plu_bmp_size(4, 4)); return 0; }\n')
    exe = tmp_path / "c_abi"
    libdir = os.path.dirname(_ffi.LIB_PATH)
    subprocess.run([gcc, "-std=c99", "-Wall", "-Wextra", "-pedantic", "-Werror", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe),
                    "-L", libdir, "-l:" + os.path.basename(_ffi.LIB_PATH), "-Wl,-rpath," + libdir], check=True, capture_output=True)
    out = subprocess.run([str(exe)], check=True, capture_output=True, text=True).stdout.split()
    assert int(out[0]) == _ffi.ABI_VERSION and int(out[1]) != 0 and int(out[2]) == lib.plu_bmp_size(4, 4)
