"""hectorgrapher_b200 — B200-native 3D scan-matching hot path of HectorGrapher.

The product is libhg_b200.so (hand-written sm_100a CUDA behind the C-ABI of
include/hg_b200.h) plus the C++ adapters in csrc/host/. This Python package is
only a ctypes view of that C-ABI for tests and bench.py; it contains no compute
and has no fallback: if the shared library is missing or no CUDA device is
present, calls fail loudly.
"""
from . import capi  # noqa: F401
from .capi import (  # noqa: F401
    Batch, ConstraintBatch, Cost, Fcsm, Grid, HgError, PbStream, Stack, build, grid_proto_decode, constraints_run_batch, get_cell_index,
    adaptive_voxel_filter, kernel_launch_count, lib, low_resolution_scores, scan_match, scan_match_evaluate, set_device, voxel_filter,
)
