"""ctypes binding of include/pbhc_b200.h.  There is NO fallback: if the CUDA library is missing or a
call fails, an exception is raised (the product path never routes through oracle/ or torch math)."""
import ctypes as C
import os

import torch  # noqa: F401  (loads libcudart.so.12 so the library resolves against torch's runtime)

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpbhc_b200.so")

c_f32p = C.c_void_p
MAX_JOINTS = 40


class PbhcError(RuntimeError):
    pass


class TableDesc(C.Structure):
    _fields_ = [("records", C.c_void_p), ("clip_num_frames", C.c_void_p), ("clip_frame_start", C.c_void_p),
                ("clip_length", C.c_void_p), ("clip_dt", C.c_void_p), ("total_frames", C.c_int64),
                ("num_clips", C.c_int32), ("num_joints", C.c_int32), ("num_bodies", C.c_int32),
                ("num_dof", C.c_int32), ("contact_size", C.c_int32), ("rec_stride", C.c_int32)]


LOOKUP_FIELDS = ["root_pos", "root_rot", "dof_pos", "root_vel", "root_ang_vel", "dof_vel", "motion_aa",
                 "motion_bodies", "rg_pos", "rb_rot", "body_vel", "body_ang_vel", "rg_pos_t", "rg_rot_t",
                 "body_vel_t", "body_ang_vel_t", "contact_mask", "f0l", "f1l", "blend"]


class LookupOut(C.Structure):
    _fields_ = [(k, C.c_void_p) for k in LOOKUP_FIELDS]


class Tree(C.Structure):
    _fields_ = [("num_joints", C.c_int32), ("parents", C.POINTER(C.c_int32)), ("offsets", C.POINTER(C.c_float)),
                ("local_rot_mat", C.POINTER(C.c_float))]


_lib = None

EXPORTS = ["pbhc_abi_version", "pbhc_last_error", "pbhc_lookup", "pbhc_fk", "pbhc_gae", "pbhc_adv_normalize", "pbhc_clip_prepare"]


def lib():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PbhcError(f"{LIB_PATH} is missing: run `python -m pbhc_b200.build` (there is no CPU fallback)")
    L = C.CDLL(LIB_PATH)
    L.pbhc_abi_version.restype = C.c_int
    L.pbhc_last_error.restype = C.c_char_p
    L.pbhc_lookup.restype = C.c_int
    L.pbhc_lookup.argtypes = [C.POINTER(TableDesc), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64,
                              C.POINTER(LookupOut), C.c_void_p]
    L.pbhc_fk.restype = C.c_int
    L.pbhc_fk.argtypes = [C.POINTER(Tree), C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                          C.c_void_p]
    L.pbhc_gae.restype = C.c_int
    L.pbhc_gae.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_int32,
                           C.c_float, C.c_float, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]
    L.pbhc_adv_normalize.restype = C.c_int
    L.pbhc_adv_normalize.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    L.pbhc_clip_prepare.restype = C.c_int
    L.pbhc_clip_prepare.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_int32,
                                    C.c_int32, C.c_int32, C.c_float, C.POINTER(C.c_double), C.c_int32, C.c_void_p,
                                    C.c_void_p, C.c_int32, C.c_void_p]
    if L.pbhc_abi_version() != 1:
        raise PbhcError("ABI version mismatch")
    _lib = L
    return L


def check(rc):
    if rc != 0:
        raise PbhcError(f"pbhc error {rc}: {lib().pbhc_last_error().decode()}")


def ptr(t, dtype=None):
    """device pointer of a contiguous CUDA tensor (None -> NULL)."""
    if t is None:
        return None
    if not t.is_cuda:
        raise PbhcError("expected a CUDA tensor (no CPU fallback)")
    if not t.is_contiguous():
        raise PbhcError("expected a contiguous tensor")
    if dtype is not None and t.dtype != dtype:
        raise PbhcError(f"expected dtype {dtype}, got {t.dtype}")
    return t.data_ptr()


def stream_ptr(device=None):
    return torch.cuda.current_stream(device).cuda_stream
