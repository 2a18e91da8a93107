"""ctypes binding of libgs2d.so (include/gs2d.h).  No fallback of any kind: if the library
is missing or the device is not sm_100, every entry point raises."""
from __future__ import annotations

import ctypes
import os
from ctypes import POINTER, Structure, c_char_p, c_float, c_int, c_size_t, c_uint32, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
# GS2D_LIB: kernel-variant A/B runs only (profiles/build_variant.py); the product loads the in-tree build
LIB_PATH = os.environ.get("GS2D_LIB") or os.path.join(_HERE, "lib", "libgs2d.so")

GS2D_OK = 0


class StateLayout(Structure):
    """mirror of gs2d_state_layout (include/gs2d.h)"""

    _fields_ = [
        ("total_bytes", c_size_t),
        ("off_status", c_size_t),
        ("off_tile_bins", c_size_t),
        ("off_ids_sorted", c_size_t),
        ("off_keys_sorted", c_size_t),
        ("off_packed", c_size_t),
        ("off_scratch", c_size_t),
        ("zero_begin", c_size_t),
        ("zero_bytes", c_size_t),
        ("key_bits", c_int),
        ("sort_passes", c_int),
    ]


class FitLayout(Structure):
    """mirror of gs2d_fit_layout (include/gs2d.h)"""

    _fields_ = [
        ("total_bytes", c_size_t),
        ("off_status", c_size_t),
        ("off_loss", c_size_t),
        ("off_tile_count", c_size_t),
        ("off_tile_off", c_size_t),
        ("off_grads", c_size_t),
        ("off_packed", c_size_t),
        ("zero_bytes", c_size_t),
        ("tiles", c_int),
    ]


class GradExchange(Structure):
    """mirror of gs2d_grad_exchange (include/gs2d.h)"""

    _fields_ = [("accumulate", c_void_p), ("peers", c_void_p), ("num_peers", c_int), ("zero", c_void_p),
                ("multicast", c_void_p), ("shard_begin", c_int), ("shard_end", c_int), ("xyz_peers", c_void_p),
                ("chol_peers", c_void_p), ("feat_peers", c_void_p), ("status_peers", c_void_p),
                ("work_peers", c_void_p), ("band_edges", c_void_p),
                ("barrier_pads", c_void_p), ("rank", c_int), ("barrier_epoch", c_uint32), ("push_per", c_int)]


class FitOptions(Structure):
    """mirror of gs2d_fit_options (include/gs2d.h)"""

    _fields_ = [("rgb_w", c_void_p), ("adan_state_w", c_void_p), ("rgb_w_grad_out", c_void_p), ("adan_step", c_int),
                ("adan_fresh", c_int), ("skip_update", c_int), ("slots", c_int), ("all_per_tile", c_int),
                ("analytic_grad", c_int), ("params8", c_void_p)]


class QuantizedFrame(Structure):
    """mirror of gs2d_quantized_frame (include/gs2d.h)"""

    _fields_ = [("xyz_half", c_void_p), ("chol_code", c_void_p), ("color_index", c_void_p), ("codebooks", c_void_p),
                ("chol_scale", c_float * 3), ("chol_beta", c_float * 3)]


FIT_LOSS_SLOTS = 64
TARGET_F32_CHW, TARGET_U8_HWC, TARGET_I420, TARGET_TILED = 0, 1, 2, 3
FIT_PHASE_GRADS, FIT_PHASE_UPDATE = 1, 2

_P = c_void_p
_SIGNATURES = {
    "gs2d_version": (c_int, []),
    "gs2d_last_error": (c_char_p, []),
    "gs2d_device_arch": (c_int, [c_int]),
    "gs2d_launch_count": (ctypes.c_longlong, []),
    "gs2d_prof_enable": (c_int, [c_int]),
    "gs2d_prof_read": (c_int, [POINTER(ctypes.c_double), POINTER(c_int), c_int]),
    "gs2d_project_chol_fwd": (c_int, [c_int, _P, _P, c_int, c_int, c_int, c_int, c_float, _P, _P, _P, _P, _P, _P]),
    "gs2d_project_scalerot_fwd": (c_int, [c_int, _P, _P, _P, c_int, c_int, c_int, c_int, c_float, _P, _P, _P, _P, _P, _P]),
    "gs2d_project_chol_bwd": (c_int, [c_int, _P, c_int, c_int, _P, _P, _P, c_int, _P, c_int, _P, _P, _P, _P]),
    "gs2d_project_scalerot_bwd": (c_int, [c_int, _P, _P, c_int, c_int, _P, _P, _P, c_int, _P, c_int, _P, _P, _P, _P, _P]),
    "gs2d_project_chol_bwd_analytic": (c_int, [c_int, _P, c_int, c_int, _P, _P, _P, c_int, _P, c_int, _P, _P, _P, _P]),
    "gs2d_project_scalerot_bwd_analytic": (c_int, [c_int, _P, _P, c_int, c_int, _P, _P, _P, c_int, _P, c_int, _P, _P, _P, _P, _P]),
    "gs2d_cov2d_bounds": (c_int, [c_int, _P, _P, _P, _P]),
    "gs2d_scan_scratch_bytes": (c_size_t, [c_int]),
    "gs2d_cumsum_tiles_hit": (c_int, [c_int, _P, _P, _P, _P, _P, c_size_t, _P]),
    "gs2d_cumsum_tiles_hit_static": (c_int, [c_int, _P, _P, _P, _P, _P]),
    "gs2d_map_intersects": (c_int, [c_int, c_int, _P, _P, _P, _P, c_int, c_int, _P, _P, _P]),
    "gs2d_sort_scratch_bytes": (c_size_t, [c_int]),
    "gs2d_sort_pairs": (c_int, [c_int, _P, _P, _P, _P, c_int, c_int, _P, c_size_t, _P]),
    "gs2d_tile_bin_edges": (c_int, [c_int, _P, _P, c_int, _P]),
    "gs2d_state_layout_for": (c_int, [c_int, c_int, c_int, c_int, POINTER(StateLayout)]),
    "gs2d_rasterize_sum_fwd": (c_int, [c_int, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P, c_size_t, _P, _P, _P, _P, _P]),
    "gs2d_raster_binned_scratch_bytes": (c_size_t, [c_int]),
    "gs2d_rasterize_sum_fwd_binned": (c_int, [c_int, c_int, _P, _P, _P, _P, _P, _P, c_int, c_int, _P, _P, _P, _P, c_size_t, _P]),
    "gs2d_rasterize_sum_bwd": (c_int, [c_int, c_int, c_int, c_int, _P, c_size_t, _P, _P, _P, _P]),
    "gs2d_rasterize_sum_bwd_binned": (c_int, [c_int, c_int, _P, _P, _P, _P, _P, _P, c_int, c_int, _P, _P, _P, _P, c_size_t, _P]),
    "gs2d_opstate_bytes": (c_size_t, [c_int, c_int, c_int]),
    "gs2d_opstate_ranges": (c_int, [c_int, c_int, c_int, POINTER(c_size_t), POINTER(c_size_t), POINTER(c_size_t)]),
    "gs2d_rasterize_sum_fwd_ranked": (c_int, [c_int, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, _P, c_size_t, _P, _P,
                                              _P, c_int, _P]),
    "gs2d_rasterize_sum_bwd_ranked": (c_int, [c_int, c_int, c_int, c_int, _P, c_size_t, _P, _P, _P, _P]),
    "gs2d_rasterize_sum_nd_fwd": (c_int, [c_int, c_int, _P, _P, _P, _P, _P, _P, _P, c_int, c_int, c_int, c_int, _P, c_size_t, _P, _P, _P, _P, _P]),
    "gs2d_rasterize_sum_nd_bwd": (c_int, [c_int, c_int, c_int, c_int, c_int, c_int, _P, c_size_t, _P, _P, _P, _P, _P, _P]),
    "gs2d_fit_layout_for": (c_int, [c_int, c_int, c_int, c_int, POINTER(FitLayout)]),
    "gs2d_fit_target_bytes": (c_size_t, [c_int, c_int]),
    "gs2d_fit_target_prepare": (c_int, [c_int, c_int, _P, c_int, _P, _P]),
    "gs2d_fit_init": (c_int, [c_int, c_int, c_int, c_int, _P, c_size_t, _P, _P]),
    "gs2d_fit_step": (c_int, [c_int, c_int, c_int, c_int, _P, _P, _P, _P, _P, c_int, _P, c_size_t, c_int,
                              c_float, c_float, c_float, c_float, c_float, c_float, c_int, _P, _P, _P, _P, _P]),
    "gs2d_fit_step_band": (c_int, [c_int, c_int, c_int, c_int, _P, _P, _P, _P, _P, c_int, _P, c_size_t, c_int,
                                   c_float, c_float, c_float, c_float, c_float, c_float, c_int, c_int, c_int, c_int,
                                   _P, _P, _P, _P, _P, _P, _P]),
    "gs2d_i420_to_rgb": (c_int, [c_int, c_int, _P, _P, _P]),
    "gs2d_render_quantized": (c_int, [c_int, c_int, c_int, c_int, _P, _P, c_size_t, _P, c_int, c_int, _P, _P, _P]),
    "gs2d_dequantize": (c_int, [c_int, _P, _P, _P, _P, _P]),
    "gs2d_frame_pipe_create": (c_int, [_P, c_size_t, c_int, POINTER(c_void_p)]),
    "gs2d_frame_pipe_destroy": (c_int, [_P]),
    "gs2d_frame_pipe_set_prepare": (c_int, [_P, c_int, c_int, c_int, _P]),
    "gs2d_frame_pipe_push": (c_int, [_P, _P, POINTER(c_int)]),
    "gs2d_frame_pipe_acquire": (c_int, [_P, c_int, _P, POINTER(c_void_p)]),
    "gs2d_frame_pipe_release": (c_int, [_P, c_int, _P]),
    "gs2d_adan_step": (c_int, [c_int, _P, _P, _P, _P, _P, _P, _P, c_float, c_float, c_float, c_float, c_float, c_float,
                               c_int, _P, c_int, _P]),
    "gs2d_prune_scratch_bytes": (c_size_t, [c_int]),
    "gs2d_prune_lowest": (c_int, [c_int, c_int, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, c_size_t, _P]),
    "gs2d_render_chol": (c_int, [c_int, c_int, c_int, c_int, _P, _P, _P, _P, c_int, _P, c_size_t, _P, c_int, c_int,
                                 _P, _P, _P]),
}

_lib = None
_arch_ok: dict = {}


def exported_symbols():
    """every symbol include/gs2d.h declares (used by the CPU load test)"""
    return sorted(_SIGNATURES)


def load() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"gsvc_b200: {LIB_PATH} is missing -- build it with `python -m gsvc_b200.build` "
                "(there is no CPU or PyTorch fallback)"
            )
        lib = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)   # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(rc: int, what: str = "") -> None:
    if rc != GS2D_OK:
        msg = load().gs2d_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"gs2d {what} failed (code {rc}): {msg}")


def require_device(index: int) -> None:
    """The kernels are sm_100a only; refuse anything else loudly (no dispatch, no fallback)."""
    ok = _arch_ok.get(index)
    if ok is None:
        arch = load().gs2d_device_arch(int(index))
        ok = arch == 100
        _arch_ok[index] = ok
        if not ok:
            raise RuntimeError(f"gsvc_b200 needs an sm_100 (B200) device, cuda:{index} reports sm_{arch}")
    elif not ok:
        raise RuntimeError(f"gsvc_b200 needs an sm_100 (B200) device, cuda:{index} is not")


STAGES = ("project_fwd", "memset", "scan_map", "sort", "edges_pack", "raster_fwd", "raster_bwd", "project_bwd",
          "fit_tiles", "adan", "other")   # GS2D_STAGE_* of include/gs2d.h


def prof_enable(on: bool) -> None:
    check(load().gs2d_prof_enable(1 if on else 0), "prof_enable")


def prof_read() -> dict:
    """{stage: (total_ms, spans)} of everything recorded since the last read (synchronises)"""
    n = len(STAGES)
    ms = (ctypes.c_double * n)()
    cnt = (c_int * n)()
    check(load().gs2d_prof_read(ms, cnt, n), "prof_read")
    return {STAGES[i]: (float(ms[i]), int(cnt[i])) for i in range(n)}


def launch_count() -> int:
    return int(load().gs2d_launch_count())
