import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as entry  # noqa: E402


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def pkg():
    return entry.load_package()


@pytest.fixture(scope="session")
def ref():
    """oracle/_ref: the reference's own sources compiled for the host (built by __graft_entry__.build())."""
    r = entry.load_oracle()
    if not r.available():
        pytest.skip("oracle/_ref/libradiance_ref.so not built")
    return r


@pytest.fixture(scope="session")
def plt(pkg):
    return pkg.rd.Platform.GetPlatform(0)


class _ParityLog:
    """Achieved parity numbers (mismatch fractions, max relative t error, RMSE) per configuration; written to
    gpurun_out/parity_r2.json at the end of the session and copied under profiles/ by hand."""

    def __init__(self):
        self.rows = {}

    def record(self, config, **numbers):
        row = self.rows.setdefault(config, {})
        for k, v in numbers.items():
            row[k] = float(v) if hasattr(v, "__float__") and not isinstance(v, (bool, int, str)) else v

    def dump(self):
        import json
        if not self.rows:
            return
        out = os.path.join(ROOT, "gpurun_out")
        try:
            os.makedirs(out, exist_ok=True)
            path = os.path.join(out, "parity_r2.json")
            old = {}
            if os.path.exists(path):
                try:
                    old = json.load(open(path))
                except Exception:
                    old = {}
            old.update(self.rows)
            json.dump(old, open(path, "w"), indent=1, sort_keys=True)
        except OSError:
            pass


@pytest.fixture(scope="session")
def parity_log():
    log = _ParityLog()
    yield log
    log.dump()


# ---- reference structures of the 10 M-triangle configurations (tests/test_gpu_fullsize.py) -------------------------------
# The reference's host builder is single-threaded and takes minutes at 10 M triangles (radiance/src/bvh.cpp:46-341). When the
# full-size parity tests are part of the session the builds start in background threads as soon as collection is done, so
# they overlap with the rest of the GPU suite (ctypes releases the GIL); the tests join them.
_FULLSIZE = {}


def fullsize_builds():
    """{"height": future -> RefBlas of heightfield(2237), "patches": future -> [RefBlas] of the 4096 C4 patch meshes}."""
    if _FULLSIZE:
        return _FULLSIZE
    from concurrent.futures import ThreadPoolExecutor
    ref = entry.load_oracle()
    scenes = entry.load_package().scenes
    if not ref.available():
        return _FULLSIZE
    pool = ThreadPoolExecutor(max_workers=max(2, (os.cpu_count() or 2)))

    def height():
        v, t = scenes.heightfield(2237, 1)
        return ref.RefBlas(v, t)

    def patches():
        b = scenes.groups_scene(2240, 64, 64, 1920, 1080)
        futs = [pool.submit(ref.RefBlas, v, t) for (v, t) in b.meshes]
        return b, futs
    _FULLSIZE["height"] = pool.submit(height)
    _FULLSIZE["patches"] = pool.submit(patches)
    _FULLSIZE["pool"] = pool
    return _FULLSIZE


def pytest_collection_finish(session):
    try:
        import torch
        gpu = torch.cuda.is_available()
    except Exception:
        gpu = False
    if gpu and any("test_gpu_fullsize" in item.nodeid for item in session.items):
        fullsize_builds()
