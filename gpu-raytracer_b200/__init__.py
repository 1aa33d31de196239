"""gpu-raytracer_b200 — B200-native path-tracing core behind the scene contract of Olowerr/GPU-Raytracer.

The directory name carries a hyphen (it is the project's name); import it with
    okrt = importlib.import_module("gpu-raytracer_b200")
Everything computational lives in libokrt.so (csrc/, built by build.py) behind include/okrt.h.
"""
from . import _abi as abi                      # noqa: F401
from .raytracer import Mesh, OctTree, Geometry, RayTracer, ResourceManager, Scene  # noqa: F401
from .raytracer import environmentCrossToFaces, scaleTexturesToArray            # noqa: F401
from .raytracer import loadObjects, ObjectDescription                           # noqa: F401
from . import scenes                           # noqa: F401
from . import sharding                         # noqa: F401
from . import harness                          # noqa: F401
