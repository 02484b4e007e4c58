"""Scene builders of the tests: the shared ones live in the package (webgpu_path_tracer_b200/standins.py)."""
from webgpu_path_tracer_b200.standins import *  # noqa: F401,F403
