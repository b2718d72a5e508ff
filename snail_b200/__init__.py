"""snail_b200 — B200-native (sm_100a CUDA) hot path of nadult/Snail behind the reference's Scene/Render interface.

The product is libsnail_b200.so (C ABI: include/snail_b200.h). This package is the ctypes host binding.
"""
from .api import (ABI_SYMBOLS, BUILD_DEVICE, BUILD_SAH_BINNED, BUILDER_LBVH, BUILDER_SAH, GPUNODE_DTYPE, host_build_binned_sah, FAST_BUILD, MAT_SIMPLE, MAT_TEX, MAT_TRANSPARENT, MAT_UBER, NO_SHADING_DATA, NODE_DTYPE,
                  SHTRI_DTYPE, TRI_DTYPE, USE_SAH, Camera, Comm, HostScene, Light, Scene, SnlError, Stats, bench_camera, comm_plan, host_dbvh_build, load_library, measure_peaks)

__all__ = ["ABI_SYMBOLS", "Camera", "Comm", "HostScene", "Light", "Scene", "SnlError", "Stats", "bench_camera", "load_library", "comm_plan", "host_dbvh_build", "measure_peaks", "NODE_DTYPE",
           "TRI_DTYPE", "SHTRI_DTYPE", "NO_SHADING_DATA", "FAST_BUILD", "USE_SAH", "BUILD_DEVICE", "BUILD_SAH_BINNED", "BUILDER_LBVH", "BUILDER_SAH", "GPUNODE_DTYPE", "host_build_binned_sah", "MAT_SIMPLE", "MAT_UBER", "MAT_TEX",
           "MAT_TRANSPARENT"]
