import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def lib():
    """libplugpu.so, built in-tree (nvcc cross-compiles without a GPU)."""
    from plutracer_b200 import _ffi, build_lib
    build_lib()
    return _ffi.lib()


@pytest.fixture(scope="session")
def oracle(lib):
    from tests.orc import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def gpu(lib):
    from plutracer_b200 import GpuContext
    ctx = GpuContext(0)   # raises PLU_E_CUDA without a device: no fallback
    yield ctx
    ctx.close()


_PARITY = {}


def record_parity(key, metrics):
    """Keep the converged-image statistics of a test run: written at session end to $PLU_PARITY_OUT (default
    gpurun_out/parity_metrics.json, the directory gpurun brings back; copied to profiles/r02_parity.json)."""
    _PARITY[key] = metrics


def pytest_sessionfinish(session, exitstatus):
    if not _PARITY:
        return
    import json
    path = os.environ.get("PLU_PARITY_OUT", os.path.join(ROOT, "gpurun_out", "parity_metrics.json"))
    try:
        os.makedirs(os.path.dirname(path), exist_ok=True)
        old = {}
        if os.path.exists(path):
            with open(path) as f:
                old = json.load(f)
        old.update(_PARITY)
        with open(path, "w") as f:
            json.dump(old, f, indent=1, sort_keys=True)
    except OSError:
        pass


REF_SCENES = ["minimal0", "cornell-box", "glass0", "refrac0", "room"]   # the reference's scenes (BASELINE.json configs)
OWN_SCENES = ["atrium", "colonnade", "shard", "gallery", "trilight"]                         # scenes/ of this repo
SCENES = REF_SCENES + OWN_SCENES
