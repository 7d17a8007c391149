"""bachelor-thesis-pathtracer_b200 — B200 (sm_100a) wavefront implementation of the hot path of
Marsimplodation/Bachelor-Thesis-Pathtracer behind a C ABI (include/ptb.h).

The directory name carries a hyphen (it mirrors the reference repository's name), so import it with
    importlib.import_module("bachelor-thesis-pathtracer_b200")
or through the repo-root shim `ptb_b200.py`.
"""
from .api import (  # noqa: F401
    LIB_PATH, REF_BVH_NODE, REF_GRID, MODE_ALL, MODE_BVH, MODE_NAMES, MODE_SUB_BEAMS, MODE_TWO_PLANE, SYMBOLS, Camera, Params, PathTracer, PathTracerGroup, PtbError,
    convert_scene_file, load_library,
)
from .build import build_native  # noqa: F401
from . import multigpu, scenes, testing_protocol  # noqa: F401,E402
