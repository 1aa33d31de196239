import importlib
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def okrt():
    """The product package (ctypes over libokrt.so).  Builds the library if it is missing."""
    lib = os.path.join(ROOT, "gpu-raytracer_b200", "libokrt.so")
    if not os.path.exists(lib):
        import __graft_entry__
        __graft_entry__.build()
    return importlib.import_module("gpu-raytracer_b200")


@pytest.fixture(scope="session")
def oracle():
    from tests import oracle_binding
    return oracle_binding.load()


@pytest.fixture(scope="session")
def ref():
    """The reference's own BvhBuilder.cpp + glm behind oracle/_ref (prebuilt; rebuilt where /root/reference exists)."""
    from tests import oracle_binding
    lib = oracle_binding.load_ref()
    if lib is None:
        pytest.skip("oracle/_ref/libokrt_ref.so not available")
    return lib


@pytest.fixture(scope="session")
def ref_octree():
    """The reference's own oct-tree builder (RayTracer.cpp:503-788) behind oracle/_ref/libokrt_ref_octree.so."""
    from oracle import binding
    lib = binding.load_ref_octree()
    if lib is None:
        pytest.skip("oracle/_ref/libokrt_ref_octree.so not available")
    return lib
