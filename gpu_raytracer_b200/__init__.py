"""gpu_raytracer_b200 — the reference's per-pixel ray tracing hot path as sm_100a
CUDA kernels behind the reference's own dispatch surface.
The compute lives in librtb200.so (csrc/); this package is the ctypes binding."""
from . import abi  # noqa: F401
from .native import LIB_PATH, load  # noqa: F401
from .host import (Scene, build_bvh, bvh_has_cycle, default_scene, terrain_scene, cornell_scene, field_scene, pack_scene_metadata,  # noqa: F401
                   camera_default, push_constants_new, load_gltf, GltfError, write_image, write_pfm)
from .renderer import Context, BufferManager, ComputeRenderer, RtError  # noqa: F401
