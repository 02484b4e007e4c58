import os
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def _cuda_device_count():
    try:
        import ctypes
        lib = ctypes.CDLL("libcuda.so.1")
        if lib.cuInit(0) != 0:
            return 0
        n = ctypes.c_int(0)
        if lib.cuDeviceGetCount(ctypes.byref(n)) != 0:
            return 0
        return n.value
    except OSError:
        return 0


def pytest_collection_modifyitems(config, items):
    # `-m gpu` on a box without a GPU must fail loudly, not skip: the marker selects, nothing else.
    pass


@pytest.fixture(scope="session")
def golden_meshes():
    d = np.load(os.path.join(ROOT, "tests", "golden", "meshes.npz"))
    return {k: d[k] for k in d.files}


@pytest.fixture(scope="session")
def default_scene():
    from webgpu_path_tracer_b200 import Scene, Camera
    s = Scene()
    s.init_mesh_data()
    s.create_meshes()
    s.create_bvh()
    cam = Camera()
    cam.set_camera([0.5, 0, 2.5], [0.5, 0, 0], [0, 1, 0])      # index.js:39
    return s, cam
