"""cuda_ray_tracer_b200 — B200-native LBVH build + ray traversal/shading core behind a C ABI.

Host-side mirror of the reference's interface (parseInput / build_lbvh_karas / render, see api.py);
all compute happens in libb200rt.so (hand-written sm_100a CUDA).  There is no CPU fallback: importing
works without a GPU, every compute call raises when the library or a B200 is missing."""
from .api import (RtError, Scene, load_library, parse_scene_file, triangle_setup, png_write, png_read,  # noqa: F401
                  BOUNDS_FROM_DESC, BOUNDS_COMPUTED, BUILD_TOPDOWN, SPP_REF, SPP_TRUE, TRAVERSE_NEAR_FIRST,
                  ROUGH_ZYX, COUNTERS, TIME_TRAVERSE, TRAVERSE_REF_ORDER, TileSet, Counters, SceneDesc)
