"""Shared scene plumbing for the tests: the SAME arrays go to the product (C ABI, gpu-raytracer_b200/harness.py)
and to the oracle (oracle/binding.py)."""
import importlib

from oracle.binding import oracle_scene  # noqa: F401

_h = importlib.import_module("gpu-raytracer_b200.harness")
build_host_side, scene_digest, Renderer = _h.build_host_side, _h.scene_digest, _h.Renderer
harness = _h                            # tests that switch pipelines set harness.DEFAULT_PIPELINE
