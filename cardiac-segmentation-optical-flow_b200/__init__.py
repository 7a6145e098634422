"""cardioflow_b200: the optical-flow hot path of nicolas1805961/Cardiac-Segmentation-Optical-flow on B200.

Host side = this thin Python mirror of the reference's module API; device side = hand-written sm_100a CUDA
behind the C ABI of include/cardioflow.h (libcardioflow.so, loaded with ctypes).  No Triton, no backend
dispatch, no CPU fallback.  Import name: ``cardioflow_b200`` (alias package at the repo root; the directory
name carries a hyphen and cannot be imported directly).
"""
from .corr import CorrBlock, coords_grid, lookup_taps
from .local_corr import CorrVolume
from .ncc import NCC
from .shard import Lanes, lane_bounds
from .upsample import upsample_flow, upsample_seg
from .warp import SpatialTransformer, SpatialTransformerContour, VecInt, utils_warp, warp, warp_onehot_argmax, warp_taps

__all__ = ["CorrBlock", "coords_grid", "lookup_taps", "SpatialTransformer", "SpatialTransformerContour", "VecInt",
           "warp", "warp_onehot_argmax", "warp_taps", "upsample_flow", "upsample_seg", "utils_warp", "Lanes", "lane_bounds", "NCC", "CorrVolume"]
