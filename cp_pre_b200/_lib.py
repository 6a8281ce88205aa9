"""ctypes binding of libcppre.so (the C ABI declared in include/cppre.h).

This is the *only* place the Python host code touches native code.  There is no fallback: if
the shared library is missing, ``load()`` raises, and every operator built on it raises with it.
"""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libcppre.so")

MAX_FIELDS, MAX_SLOTS, MAX_SCALARS, MAX_TAPS = 6, 6, 12, 512
NOALLOC_MAX_RANKS = 16      # CPPRE_NOALLOC_MAX_RANKS

OK, ERR_BAD_ARG, ERR_UNSUPPORTED, ERR_CUDA = 0, 1, 2, 3

(PROG_LINEAR, PROG_BURGERS, PROG_NS_CONTINUITY, PROG_NS_MOMENTUM, PROG_MHD_CONTINUITY,
 PROG_MHD_MOMENTUM, PROG_MHD_ENERGY, PROG_MHD_INDUCTION, PROG_MHD_GAUSS, PROG_JOREK_CONTINUITY,
 PROG_JOREK_TEMPERATURE) = range(11)
PROGRAM_FIELDS = {PROG_LINEAR: 1, PROG_BURGERS: 1, PROG_NS_CONTINUITY: 2, PROG_NS_MOMENTUM: 3,
                  PROG_MHD_CONTINUITY: 3, PROG_MHD_MOMENTUM: 6, PROG_MHD_ENERGY: 6,
                  PROG_MHD_INDUCTION: 4, PROG_MHD_GAUSS: 2, PROG_JOREK_CONTINUITY: 4, PROG_JOREK_TEMPERATURE: 4}
PROGRAM_SLOTS = {PROG_LINEAR: 1, PROG_BURGERS: 3, PROG_NS_CONTINUITY: 2, PROG_NS_MOMENTUM: 4,
                 PROG_MHD_CONTINUITY: 3, PROG_MHD_MOMENTUM: 3, PROG_MHD_ENERGY: 3,
                 PROG_MHD_INDUCTION: 3, PROG_MHD_GAUSS: 3, PROG_JOREK_CONTINUITY: 5, PROG_JOREK_TEMPERATURE: 5}

EPI_STORE, EPI_SCORE, EPI_COVER = 0, 1, 2
REDUCE_MAX, REDUCE_SUM = 0, 1
FLAG_ABS, FLAG_INTERIOR_0, FLAG_INTERIOR_1, FLAG_INTERIOR_2, FLAG_IN_BF16, FLAG_DECODE = 1, 2, 4, 8, 16, 32
PATH_AUTO, PATH_GENERIC, PATH_TMA, PATH_TIN, PATH_FLAT = 0, 1, 2, 3, 4

WALL_TOP, WALL_BOTTOM, WALL_LEFT, WALL_RIGHT = range(4)
BC_DIRICHLET, BC_NEUMANN, BC_PERIODIC, BC_SYMMETRIC, BC_NONE = range(5)
SPEC_CONJ_FIRST, SPEC_INVERSE, SPEC_CONJ_LAST = 1, 2, 4


class Field(C.Structure):
    _fields_ = [("d_data", C.c_void_p), ("stride", C.c_int64 * 4), ("scale", C.c_float), ("offset", C.c_float)]


class Program(C.Structure):
    _fields_ = [("kind", C.c_int32), ("n_fields", C.c_int32), ("n_slots", C.c_int32),
                ("ksize", (C.c_int32 * 3) * MAX_SLOTS), ("pad", (C.c_int32 * 3) * MAX_SLOTS),
                ("h_taps", C.c_void_p * MAX_SLOTS), ("scalars", C.c_float * MAX_SCALARS)]


class Epilogue(C.Structure):
    _fields_ = [("mode", C.c_int32), ("flags", C.c_int32), ("reduce", C.c_int32),
                ("band_q", C.c_float), ("d_ref", C.c_void_p), ("ref_stride", C.c_int64 * 4),
                ("d_weight", C.c_void_p), ("d_out", C.c_void_p), ("out_stride", C.c_int64 * 4),
                ("d_scores", C.c_void_p), ("d_band_q", C.c_void_p), ("weight_stride", C.c_int64 * 3)]


class BC(C.Structure):
    _fields_ = [("type", C.c_int32 * 4), ("width", C.c_int32 * 4), ("value", C.c_float * 4)]


# name -> (restype, argtypes); every symbol include/cppre.h declares
_P = C.c_void_p
_I64 = C.c_int64
SIGNATURES = {
    "cppre_version": (C.c_int, []),
    "cppre_last_error": (C.c_char_p, []),
    "cppre_launch_count": (C.c_uint64, []),
    "cppre_last_path": (C.c_char_p, []),
    "cppre_residual": (C.c_int, [C.POINTER(Program), C.POINTER(_I64), C.POINTER(Field),
                                 C.POINTER(Epilogue), C.c_int, _P]),
    "cppre_residual_tin_eligible": (C.c_int, [C.POINTER(Program), C.POINTER(_I64), C.POINTER(Field),
                                              C.POINTER(Epilogue)]),
    "cppre_residual_tma_eligible": (C.c_int, [C.POINTER(Program), C.POINTER(_I64), C.POINTER(Field),
                                              C.POINTER(Epilogue)]),
    "cppre_cell_moments": (C.c_int, [_P, _I64, _P, _I64, _I64, _I64, _P, _P, _P]),
    "cppre_cell_std": (C.c_int, [_P, C.c_double, C.c_float, _I64, _P, _P, _P]),
    "cppre_cell_modulation": (C.c_int, [_P, _I64, _P, _I64, _I64, _I64, C.c_float, _P, _P, _P]),
    "cppre_joint_score": (C.c_int, [_P, _I64, _P, _I64, _P, _I64, _I64, _P, _P]),
    "cppre_select_workspace_bytes": (C.c_size_t, [_I64, _I64]),
    "cppre_select_ranks": (C.c_int, [_P, _I64, _I64, _I64, C.POINTER(C.c_int32), C.c_int32, _P, _P, _P]),
    "cppre_ordered_keys": (C.c_int, [_P, _I64, _P, _P]),
    "cppre_window_compact": (C.c_int, [_P, _I64, _I64, _I64, _P, _P, C.c_int32, _I64, _P, _P, _P, _P, C.c_int32,
                                       C.POINTER(C.c_void_p), C.c_int32, C.c_int32, _P]),
    "cppre_select_from_top": (C.c_int, [_P, _I64, _I64, _I64, _P, C.c_int32, C.c_int32, _P, _P, _P, C.c_int32, _I64, _P]),
    "cppre_peer_copy": (C.c_int, [_P, _I64, _I64, C.POINTER(C.c_void_p), C.c_int32, _P]),
    "cppre_peer_bcast_keys": (C.c_int, [_P, _I64, _I64, C.POINTER(C.c_void_p), C.c_int32, C.c_int32, _P]),
    "cppre_window_exchange": (C.c_int, [_P, _I64, _I64, _I64, _P, _I64, C.c_int32, C.c_int32, _I64, C.POINTER(C.c_void_p),
                                        C.c_int32, C.c_int32, _P, _P, C.c_int32, _P]),
    "cppre_owner_select": (C.c_int, [_P, C.c_int32, C.c_int32, _I64, _I64, C.c_int32, C.c_int32, _I64,
                                     C.POINTER(C.c_void_p), C.c_int32, _P, _P, C.c_int32, _P]),
    "cppre_radix_hist": (C.c_int, [_P, _I64, C.c_uint32, C.c_uint32, C.c_int32, _P, _P]),
    "cppre_radix_select": (C.c_int, [_P, _I64, _I64, _P, _P]),
    "cppre_radix_select_begin": (C.c_int, [_P, _P, _I64, _P]),
    "cppre_radix_select_hist": (C.c_int, [_P, _I64, _P, C.c_int32, _P, _P]),
    "cppre_radix_select_pick": (C.c_int, [_P, _P, C.c_int32, _P, _P]),
    "cppre_radix_select_ranks_noalloc": (C.c_int, [_P, _I64, C.POINTER(_I64), C.c_int32, _P, _P]),
    "cppre_inside_count": (C.c_int, [_P, _I64, _P, _I64, _P, _I64, _I64, _I64, _P, _P]),
    "cppre_inside_count_band": (C.c_int, [_P, _I64, _P, _I64, _P, C.c_float, _I64, _I64, _P, _P]),
    "cppre_transpose_inner": (C.c_int, [_P, _I64, _I64, _I64, _I64, _P, _P]),
    "cppre_bc_padded_shape": (C.c_int, [C.POINTER(BC), _I64, _I64, C.POINTER(_I64), C.POINTER(_I64)]),
    "cppre_bc_index_map": (C.c_int, [C.POINTER(BC), _I64, _I64, C.POINTER(C.c_int32), C.POINTER(C.c_float),
                                     C.POINTER(C.c_int32), C.POINTER(C.c_float)]),
    "cppre_bc_pad": (C.c_int, [_P, _I64, _I64, _I64, C.POINTER(BC), _P, _P]),
    "cppre_bc_conv2d": (C.c_int, [_P, _I64, _I64, _I64, C.POINTER(BC), C.POINTER(C.c_float), C.c_int32,
                                  C.c_int32, _P, _P]),
    "cppre_edge_difference": (C.c_int, [_P, _I64, _I64, _I64, _I64, _I64, _I64, C.c_int32, C.c_float, _P, _P]),
    "cppre_spectral_pointwise": (C.c_int, [_P, _P, _I64, _I64, C.c_int32, C.c_float, _P, _P]),
}

_lib = None


class CppreError(RuntimeError):
    """A libcppre call failed; ``code`` is the CPPRE_ERR_* value."""

    def __init__(self, code, message):
        super().__init__(message)
        self.code = code


def load():
    """Loads libcppre.so once.  Raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -m cp_pre_b200.build` (needs nvcc). "
            "cp_pre_b200 has no CPU or PyTorch fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)     # AttributeError if the library lacks a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    """Maps a return code to the exception type the reference raises in the same situation."""
    if rc == OK:
        return
    msg = load().cppre_last_error().decode("utf-8", "replace")
    if rc == ERR_BAD_ARG:
        raise ValueError(msg)
    if rc == ERR_UNSUPPORTED:
        raise NotImplementedError(msg)
    raise CppreError(rc, msg)


def launch_count() -> int:
    return int(load().cppre_launch_count())


def last_path() -> str:
    return load().cppre_last_path().decode()
