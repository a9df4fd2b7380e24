"""raytracing_cuda_b200 -- B200-native drop-in for the ray-tracing hot path of
Miki96/raytracing_cuda (the `launchKernel` boundary, reference kernel.cu:406-462).

The compute path is hand-written CUDA for sm_100a in csrc/ behind the C-ABI of
include/rtb_api.h; this package is the ctypes host mirror used by tests, bench and tools.
"""
from . import _lib, structs  # noqa: F401
from .renderer import PinnedFrames, Renderer, default_stride, make_shard, rays_per_frame  # noqa: F401
from .scene import (POSES, TIME_PRESETS, VIEW_ISLAND, VIEW_MOUNTAINS, Scene, heightmap_objects,  # noqa: F401
                    load_obj, load_sky_pngs, read_png, synthetic_sky, terrain_objects)
