"""ctypes binding of libhcg.so (the C-ABI declared in include/hcg.h).

There is no CPU implementation behind this module: if the library cannot be loaded (or built
with nvcc), importing the product's compute path raises.
"""
from __future__ import annotations

import ctypes as C
import os
import threading

from . import build as _build

HCG_MAX_CONV = 8
HCG_MAX_READOUT = 8
HCG_COLLATE_TRANSPOSE = 1
HCG_COLLATE_GROUPED = 2
HCG_ERR_NOT_GROUPED = 7
HCG_GRAPH_CROSS_EDGES = 1

ERR_NAMES = {1: "INVALID_ARG", 2: "UNSUPPORTED", 3: "WORKSPACE", 4: "CUDA", 5: "BAD_INDEX", 6: "BAD_BATCH", 7: "NOT_GROUPED"}

_fp = C.c_void_p  # device pointers travel as integers


class HcgGraph(C.Structure):
    _fields_ = [
        ("N", C.c_int64), ("E", C.c_int64), ("B", C.c_int64),
        ("rowptr", _fp), ("col", _fp), ("eid", _fp), ("dis", _fp), ("graph_ptr", _fp),
        ("t_rowptr", _fp), ("t_col", _fp), ("t_eid", _fp),
        ("fill", C.c_float),
        ("ew0", _fp), ("t_ew0", _fp), ("selfw0", _fp), ("dis0", _fp),
        ("max_graph_nodes", C.c_int32), ("flags", C.c_int32),
        ("edge_mask", _fp),
        ("max_graph_edges", C.c_int32), ("reserved0", C.c_int32),
    ]


class HcgParams(C.Structure):
    _fields_ = [
        ("F", C.c_int32), ("H", C.c_int32), ("L", C.c_int32), ("R", C.c_int32), ("C", C.c_int32),
        ("slope", C.c_float),
        ("conv_w", _fp * HCG_MAX_CONV), ("conv_b", _fp * HCG_MAX_CONV),
        ("ro_w", _fp * HCG_MAX_READOUT), ("ro_b", _fp * HCG_MAX_READOUT),
    ]


class HcgGrads(C.Structure):
    _fields_ = [
        ("conv_w", _fp * HCG_MAX_CONV), ("conv_b", _fp * HCG_MAX_CONV),
        ("ro_w", _fp * HCG_MAX_READOUT), ("ro_b", _fp * HCG_MAX_READOUT),
        ("edge_mask", _fp),
    ]


class HcgSaved(C.Structure):
    _fields_ = [("acts", _fp * HCG_MAX_CONV), ("hid", _fp * HCG_MAX_READOUT)]


class HcgEnsemble(C.Structure):
    _fields_ = [("M", C.c_int32), ("F", C.c_int32), ("H", C.c_int32), ("C", C.c_int32), ("slope", C.c_float),
                ("reserved0", C.c_int32), ("conv_w0", _fp), ("conv_b0", _fp), ("conv_w1", _fp), ("conv_b1", _fp),
                ("ro_w0", _fp), ("ro_b0", _fp), ("ro_w1", _fp), ("ro_b1", _fp)]


class HcgError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"libhcg error {code} ({ERR_NAMES.get(code, '?')}): {msg}")
        self.code = code


_lock = threading.Lock()
class HcgPackedStore(C.Structure):
    _fields_ = [("G", C.c_int64), ("N", C.c_int64), ("E", C.c_int64),
                ("F", C.c_int32), ("Fu", C.c_int32), ("Ff", C.c_int32), ("edge_kind", C.c_int32),
                ("col_slot", _fp), ("node_ptr", _fp), ("edge_ptr", _fp), ("y", _fp), ("xu8", _fp), ("xf32", _fp),
                ("edges", _fp), ("Fb", C.c_int32), ("bit_stride", C.c_int32), ("xbits", _fp)]


_lib = None

# name -> (restype, argtypes); the test-suite checks this table against include/hcg.h
SIGNATURES = {
    "hcg_last_error": (C.c_char_p, []),
    "hcg_version": (C.c_char_p, []),
    "hcg_launch_count": (C.c_uint64, []),
    "hcg_profile_enable": (C.c_int, [C.c_int]),
    "hcg_profile_read": (C.c_int, [C.c_char_p, C.POINTER(C.c_float), C.POINTER(C.c_int64), C.c_int]),
    "hcg_config_set": (C.c_int, [C.c_int, C.c_int]),
    "hcg_config_get": (C.c_int, [C.c_int]),
    "hcg_selftest_tc_gemm": (C.c_int, [_fp, _fp, C.c_int, _fp, _fp]),
    "hcg_selftest_tc_xty": (C.c_int, [_fp, _fp, C.c_int, _fp, _fp]),
    "hcg_sizeof_graph": (C.c_size_t, []),
    "hcg_sizeof_params": (C.c_size_t, []),
    "hcg_sizeof_grads": (C.c_size_t, []),
    "hcg_sizeof_saved": (C.c_size_t, []),
    "hcg_sizeof_ensemble": (C.c_size_t, []),
    "hcg_ensemble_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int32]),
    "hcg_gcn_forward_coo_ensemble": (C.c_int, [_fp, _fp, _fp, _fp, C.c_int64, C.c_int64, C.c_int64, C.c_float,
                                               C.POINTER(HcgEnsemble), _fp, _fp, _fp, C.c_size_t, _fp]),
    "hcg_collate_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int64, C.c_int]),
    "hcg_collate": (C.c_int, [_fp, C.c_int64, _fp, C.c_int64, C.c_int64, C.c_float, _fp, C.c_float, C.c_int,
                              _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp,
                              _fp, C.c_size_t, _fp]),
    "hcg_collate_offsets": (C.c_int, [_fp, C.c_int64, _fp, _fp, C.c_int64, C.c_int64, C.c_float, C.c_int, C.c_int32, C.c_int32,
                                      _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, _fp, C.c_size_t, _fp]),
    "hcg_collate_status": (C.c_int, [_fp, _fp, C.POINTER(C.c_int32)]),
    "hcg_forward_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int64, C.POINTER(HcgParams)]),
    "hcg_gcn_forward": (C.c_int, [C.POINTER(HcgGraph), _fp, C.POINTER(HcgParams), _fp, _fp, C.POINTER(HcgSaved),
                                  _fp, C.c_size_t, _fp]),
    "hcg_forward_coo_supported": (C.c_int, [C.POINTER(HcgParams)]),
    "hcg_forward_coo_workspace_bytes": (C.c_size_t, [C.c_int64, C.POINTER(HcgParams)]),
    "hcg_gcn_forward_coo": (C.c_int, [_fp, _fp, _fp, _fp, C.c_int64, C.c_int64, C.c_int64, C.c_float, C.POINTER(HcgParams),
                                      _fp, _fp, _fp, C.c_size_t, _fp]),
    "hcg_gcn_forward_coo_static": (C.c_int, [_fp, _fp, _fp, _fp, C.c_int64, C.c_int64, C.c_int64, C.c_float, C.POINTER(HcgParams),
                                             _fp, _fp, C.POINTER(C.c_int32), _fp]),
    "hcg_forward_coo_status": (C.c_int, [_fp, C.c_int64, _fp, C.POINTER(C.c_int32)]),
    "hcg_backward_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int64, C.POINTER(HcgParams)]),
    "hcg_gcn_backward": (C.c_int, [C.POINTER(HcgGraph), _fp, C.POINTER(HcgParams), C.POINTER(HcgSaved), _fp, _fp, _fp,
                                   C.POINTER(HcgGrads), _fp, _fp, C.c_size_t, _fp]),
    "hcg_batch_from_ptr": (C.c_int, [_fp, C.c_int64, C.c_int64, _fp, _fp]),
    "hcg_unpack_batch": (C.c_int, [C.POINTER(HcgPackedStore), _fp, C.c_int64, C.c_int64, _fp, _fp, _fp, _fp, C.c_int64,
                                   _fp, _fp, _fp]),
    "hcg_forward_packed_supported": (C.c_int, [C.POINTER(HcgPackedStore), C.POINTER(HcgParams)]),
    "hcg_gcn_forward_packed": (C.c_int, [C.POINTER(HcgPackedStore), C.c_int64, C.c_int64, C.c_float, C.POINTER(HcgParams),
                                         _fp, _fp, _fp, C.c_size_t, _fp]),
    "hcg_rmse_seed": (C.c_int, [_fp, _fp, C.c_int64, C.c_int, _fp, _fp, _fp]),
    "hcg_train_step_workspace_bytes": (C.c_size_t, [C.c_int64, C.c_int64, C.c_int64, C.POINTER(HcgParams)]),
    "hcg_train_step": (C.c_int, [_fp, _fp, C.c_int64, _fp, _fp, _fp, C.c_int64, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_float,
                                 C.POINTER(HcgParams), C.POINTER(HcgGrads), _fp, _fp, _fp, _fp, C.c_int64,
                                 C.c_double, C.c_double, C.c_double, C.c_double, C.c_int64, _fp, _fp, C.c_size_t, _fp]),
    "hcg_adam_step": (C.c_int, [_fp, _fp, _fp, _fp, C.c_int64, _fp, C.c_double, C.c_double, C.c_double, C.c_double,
                                C.c_int64, _fp, _fp]),
}


def lib_path() -> str:
    return _build.LIB_PATH


def load():
    """Load (building first if the sources are newer and nvcc exists) and return the CDLL."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        path = os.environ.get("HCG_LIB_PATH")  # an explicitly chosen build (e.g. an instrumented one)
        if path:
            if not os.path.exists(path):
                raise RuntimeError(f"HCG_LIB_PATH={path} does not exist")
        elif _build.needs_build():
            path = _build.LIB_PATH
            if _build.find_nvcc() is not None:
                _build.build()
            elif not os.path.exists(path):
                raise RuntimeError(
                    "hcatgnet_b200: libhcg.so is missing and nvcc is not available; the CUDA extension is "
                    "required (there is no CPU fallback). Run `python -m hcatgnet_b200.build` on a box with nvcc.")
        else:
            path = _build.LIB_PATH
        lib = C.CDLL(path)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError here = library/header mismatch: fail loudly
            fn.restype = res
            fn.argtypes = args
        _lib = lib
    return _lib


def check(code: int) -> None:
    if code != 0:
        raise HcgError(code, load().hcg_last_error().decode("utf-8", "replace"))


HCG_CFG_FORWARD_PATH = 0
HCG_CFG_BACKWARD_PATH = 1
HCG_CFG_GEMM_PATH = 2
HCG_CFG_AGG_PATH = 3
PATH_AUTO, PATH_LAYERWISE, PATH_GRAPH_RESIDENT, PATH_TENSOR_CORE = 0, 1, 2, 3


def config_set(key: int, value: int) -> int:
    """Library-wide switch (tests / benchmarks); returns the previous value."""
    prev = load().hcg_config_set(key, value)
    if prev < 0:
        raise ValueError(f"unknown libhcg config key {key}")
    return prev


def launch_count() -> int:
    return int(load().hcg_launch_count())


def profile_enable(on: bool) -> None:
    check(load().hcg_profile_enable(1 if on else 0))


def profile_read(cap: int = 64) -> dict:
    """{kernel name: (total device ms, launches)} since the last profile_enable(True)."""
    names = C.create_string_buffer(cap * 64)
    ms = (C.c_float * cap)()
    cnt = (C.c_int64 * cap)()
    rc = load().hcg_profile_read(names, ms, cnt, cap)
    if rc > 0:
        check(rc)
    out = {}
    for i in range(-rc):
        out[names.raw[i * 64:(i + 1) * 64].split(b"\0", 1)[0].decode()] = (float(ms[i]), int(cnt[i]))
    return out
