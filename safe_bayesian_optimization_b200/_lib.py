"""ctypes binding of libsbo.so (C-ABI in include/sbo.h).

There is no CPU fallback: if the CUDA library is missing this module raises at
load time, and ``sbo_create`` fails on a machine without an sm_100 GPU.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB_PATH = os.path.join(CSRC, "libsbo.so")

SBO_FP32, SBO_FP64 = 0, 1
SBO_KERNEL_AUTO, SBO_KERNEL_SIMT, SBO_KERNEL_TCGEN05 = 0, 1, 2

# sbo_option
OPTIONS = {"tc_cta_group": 1, "tc_skip": 2, "tc_pair": 3, "tc_eager": 4, "tc_inc": 5, "thin": 6, "thin_2d": 7, "tc_graded": 8,
           "tc_strip": 9, "refresh_every": 10, "drift_check_every": 11, "drift_rows": 12,
           "dist_fit": 13, "tc_phases": 14, "mask_delta": 15}

SBO_OK = 0
SBO_ERR_INVALID_ARG, SBO_ERR_NO_DEVICE, SBO_ERR_CUDA, SBO_ERR_NOT_POSDEF = -1, -2, -3, -4
SBO_ERR_CAPACITY, SBO_ERR_STATE, SBO_ERR_COMM = -5, -6, -7


class SboConfig(C.Structure):
    _fields_ = [("sigma_f", C.c_double), ("length_scale", C.c_double), ("noise_var", C.c_double),
                ("prior_mean", C.c_double), ("include_noise_in_sigma", C.c_int32), ("precision", C.c_int32),
                ("kernel", C.c_int32), ("device", C.c_int32), ("max_obs", C.c_int32), ("reserved0", C.c_int32),
                ("max_cells", C.c_int64)]


class SboArgmax(C.Structure):
    _fields_ = [("idx", C.c_int64), ("value", C.c_double)]


class SboStepOut(C.Structure):
    _fields_ = [("n_cells", C.c_int64), ("safe_count", C.c_int64), ("argmax_ucb", SboArgmax),
                ("argmax_width", SboArgmax), ("clamped_sigma2", C.c_int64), ("n_obs", C.c_int32),
                ("kernel_used", C.c_int32), ("ms_posterior", C.c_float), ("ms_sets", C.c_float),
                ("ms_comm", C.c_float), ("ms_total", C.c_float),
                ("low_sigma2", C.c_int64), ("refreshed", C.c_int32), ("drift", C.c_float)]


class SboSetsRecord(C.Structure):
    _fields_ = [("ucb_value", C.c_double), ("ucb_idx", C.c_int64), ("width_value", C.c_double), ("width_idx", C.c_int64),
                ("safe_count", C.c_int64), ("clamped", C.c_int64), ("low_sigma2", C.c_int64)]


class SboMeasurement(C.Structure):
    _fields_ = [("x", C.c_double), ("y", C.c_double), ("z", C.c_double), ("value", C.c_double),
                ("uncertainty", C.c_double), ("stamp_sec", C.c_int32), ("stamp_nanosec", C.c_uint32)]


SBO_SUBGOAL_NONE, SBO_SUBGOAL_GOAL, SBO_SUBGOAL_PROJECTED = 0, 1, 2

# every symbol include/sbo.h declares: name -> (restype, argtypes)
_P = C.c_void_p
_D = C.POINTER(C.c_double)
_U8 = C.POINTER(C.c_uint8)
SYMBOLS = {
    "sbo_abi_version": (C.c_int32, []),
    "sbo_create": (C.c_int32, [C.POINTER(SboConfig), C.POINTER(_P)]),
    "sbo_destroy": (C.c_int32, [_P]),
    "sbo_last_error": (C.c_char_p, [_P]),
    "sbo_status_string": (C.c_char_p, [C.c_int32]),
    "sbo_set_stream": (C.c_int32, [_P, _P]),
    "sbo_synchronize": (C.c_int32, [_P]),
    "sbo_set_grid": (C.c_int32, [_P, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, C.c_int32,
                                 C.c_int32, C.c_int32]),
    "sbo_set_observations": (C.c_int32, [_P, _D, _D, C.c_int32]),
    "sbo_append_observations": (C.c_int32, [_P, _D, _D, C.c_int32]),
    "sbo_num_observations": (C.c_int32, [_P]),
    "sbo_fit_info": (C.c_int32, [_P, C.POINTER(C.c_float), C.POINTER(C.c_int32)]),
    "sbo_mask_copy_info": (C.c_int32, [_P, C.POINTER(C.c_int64), C.POINTER(C.c_int32)]),
    "sbo_copy_order": (C.c_int32, [_P, C.POINTER(C.c_int32)]),
    "sbo_copy_linv": (C.c_int32, [_P, _D]),
    "sbo_copy_alpha": (C.c_int32, [_P, _D]),
    "sbo_step": (C.c_int32, [_P, C.c_double, C.c_double, _U8, C.POINTER(SboStepOut)]),
    "sbo_step_incremental": (C.c_int32, [_P, C.c_double, C.c_double, _U8, C.POINTER(SboStepOut)]),
    "sbo_compute_sets": (C.c_int32, [_P, _D, _D, C.c_int64, C.c_double, C.c_double, _U8, _D, _D,
                                     C.POINTER(SboStepOut)]),
    "sbo_select_subgoal": (C.c_int32, [_P, C.POINTER(C.c_int32), C.c_int64, C.c_double, C.c_double,
                                       C.POINTER(C.c_int32)]),
    "sbo_frontier_pixel_maps": (C.c_int32, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "sbo_frontier_image": (C.c_int32, [_P, _U8, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "sbo_compact_cells": (C.c_int32, [_P, C.c_int32, C.POINTER(C.c_int32), C.c_int64, C.POINTER(C.c_int64)]),
    "sbo_point_within_polygon": (C.c_int32, [_D, C.POINTER(C.c_int32), C.c_int32, C.c_double, C.c_double,
                                             C.POINTER(C.c_int32)]),
    "sbo_polydist": (C.c_int32, [_D, C.c_int32, C.c_double, C.c_double, _D, _D]),
    "sbo_subgoal_point": (C.c_int32, [_P, C.POINTER(C.c_int32), C.c_int64, C.c_double, C.c_double, _D,
                                      C.POINTER(C.c_int32), C.c_int32, _D, C.POINTER(C.c_int32),
                                      C.POINTER(C.c_int32), _D]),
    "sbo_ingest_measurements": (C.c_int32, [_P, C.POINTER(SboMeasurement), C.c_int32, C.POINTER(C.c_int32),
                                            C.POINTER(C.c_int32)]),
    "sbo_tc_plan": (C.c_int32, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_double)]),
    "sbo_tc_plan_detail": (C.c_int32, [_P, _D]),
    "sbo_tc_executed_flops": (C.c_int32, [_P, _D]),
    "sbo_set_option": (C.c_int32, [_P, C.c_int32, C.c_int32]),
    "sbo_get_option": (C.c_int32, [_P, C.c_int32, C.POINTER(C.c_int32)]),
    "sbo_plan_row_bands": (C.c_int32, [C.POINTER(SboConfig), _D, C.c_int32, C.c_double, C.c_double, C.c_double, C.c_double,
                                       C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32), _D]),
    "sbo_copy_mu": (C.c_int32, [_P, _P]),
    "sbo_copy_sigma2": (C.c_int32, [_P, _P]),
    "sbo_copy_mu_f64": (C.c_int32, [_P, _D]),
    "sbo_copy_sigma2_f64": (C.c_int32, [_P, _D]),
    "sbo_copy_mask": (C.c_int32, [_P, _U8]),
    "sbo_step_device": (C.c_int32, [_P, C.c_double, C.c_double, C.POINTER(SboStepOut)]),
    "sbo_compute_sets_device": (C.c_int32, [_P, _P, _P, C.c_int64, C.c_double, C.c_double, _P,
                                            C.POINTER(SboStepOut)]),
    "sbo_compute_sets_noalloc": (C.c_int32, [_P, _P, C.c_int64, C.c_double, C.c_double, _P, _P, _P]),
    "sbo_comm_unique_id": (C.c_int32, [_U8]),
    "sbo_comm_init": (C.c_int32, [_P, _U8, C.c_int32, C.c_int32]),
    "sbo_comm_destroy": (C.c_int32, [_P]),
    "sbo_merge_step_out": (C.c_int32, [C.POINTER(SboStepOut), C.c_int32, C.POINTER(SboStepOut)]),
    "sbo_group_create": (C.c_int32, [C.POINTER(SboConfig), C.POINTER(C.c_int32), C.c_int32, C.POINTER(_P)]),
    "sbo_group_destroy": (C.c_int32, [_P]),
    "sbo_group_last_error": (C.c_char_p, [_P]),
    "sbo_group_size": (C.c_int32, [_P]),
    "sbo_group_member": (_P, [_P, C.c_int32]),
    "sbo_group_bands": (C.c_int32, [_P, C.POINTER(C.c_int32)]),
    "sbo_group_set_option": (C.c_int32, [_P, C.c_int32, C.c_int32]),
    "sbo_group_set_grid": (C.c_int32, [_P, C.c_double, C.c_double, C.c_double, C.c_double, C.c_int32, C.c_int32]),
    "sbo_group_set_observations": (C.c_int32, [_P, _D, _D, C.c_int32]),
    "sbo_group_append_observations": (C.c_int32, [_P, _D, _D, C.c_int32]),
    "sbo_group_num_observations": (C.c_int32, [_P]),
    "sbo_group_step": (C.c_int32, [_P, C.c_double, C.c_double, _U8, C.POINTER(SboStepOut)]),
    "sbo_group_step_incremental": (C.c_int32, [_P, C.c_double, C.c_double, _U8, C.POINTER(SboStepOut)]),
    "sbo_group_select_subgoal": (C.c_int32, [_P, C.POINTER(C.c_int32), C.c_int64, C.c_double, C.c_double,
                                             C.POINTER(C.c_int32)]),
    "sbo_group_copy_mu_f64": (C.c_int32, [_P, _D]),
    "sbo_group_copy_sigma2_f64": (C.c_int32, [_P, _D]),
    "sbo_launch_count": (C.c_int64, [_P]),
    "sbo_device_info": (C.c_int32, [_P, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
}


def build(force: bool = False, extra: str = "") -> str:
    """Compile libsbo.so in-tree with nvcc for sm_100a (csrc/Makefile)."""
    if force:
        subprocess.run(["make", "-C", CSRC, "-s", "clean"], check=True)
    cmd = ["make", "-C", CSRC, "-s", "-j8"]
    if extra:
        cmd.append(f"EXTRA={extra}")
    subprocess.run(cmd, check=True)
    return LIB_PATH


_lib = None


def load() -> C.CDLL:
    """Load the CUDA library; raise loudly if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a).  There is no CPU fallback.")
    # make torch's bundled NCCL findable by the library's dlopen when torch is in the process
    if "SBO_NCCL_LIB" not in os.environ:
        try:
            import importlib.util
            spec = importlib.util.find_spec("nvidia.nccl")
            if spec and spec.submodule_search_locations:
                cand = os.path.join(list(spec.submodule_search_locations)[0], "lib", "libnccl.so.2")
                if os.path.exists(cand):
                    os.environ["SBO_NCCL_LIB"] = cand
        except Exception:
            pass
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the header and the library disagree
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


class SboError(RuntimeError):
    def __init__(self, status: int, message: str):
        super().__init__(f"sbo status {status}: {message}")
        self.status = status
