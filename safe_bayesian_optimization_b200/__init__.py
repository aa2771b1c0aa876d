"""B200-native safe-BO step: GP posterior over every terrain cell, safe-set mask and
acquisition argmax, behind the C-ABI of include/sbo.h (libsbo.so, sm_100a only).

No CPU fallback: importing is cheap, but creating a context needs the built CUDA
library and an sm_100 GPU.
"""
from ._lib import LIB_PATH, SboError, build, load
from .node import OptimizerCore, find_safety_contour_indices
from .optimizer import SboContext, SboGroup, StepResult, merge_results, plan_row_bands, shard_rows

__all__ = ["LIB_PATH", "SboError", "build", "load", "OptimizerCore", "find_safety_contour_indices",
           "SboContext", "SboGroup", "StepResult", "merge_results", "plan_row_bands", "shard_rows"]
