"""B200-native path-tracing core behind the reference's buffer/dispatch boundary.

`api.PathTracer` is the ctypes face of libwgpt.so (include/wgpt.h); `renderer.Renderer`,
`scene.Scene`, `scene.Camera` mirror the reference's host classes.  The CUDA library is the only
compute path: importing this package never touches oracle/.
"""
from .api import PathTracer, Group, WgptError, MODE_EXACT, MODE_FAST, MODE_FAST_TOL, load_library, nccl_unique_id  # noqa: F401
from .scene import Scene, Camera, Transform, Sphere, Quad, Mesh, ObjReader  # noqa: F401
from .renderer import Renderer  # noqa: F401
