"""path_tracer_b200 — B200-native (sm_100a) backend for the render hot path of CouncilmanJeremyJamm/path_tracer.

Only what the path needs lives here: `csrc/` (CUDA kernels, host-side Scene::new, the C ABI of include/ptb.h built into
libptb.so), `api.py` (ctypes binding + mirror of the reference's Scene/Camera/Model/material interface) and `scenes.py`
(procedural stand-ins for the reference's unshipped assets).
"""
from . import scenes  # noqa: F401
from .api import (Camera, Context, Dielectric, Emissive, GGX, Lambertian, Model, PtbError, Scene, Specular, Volume, lib, lib_path)  # noqa: F401
