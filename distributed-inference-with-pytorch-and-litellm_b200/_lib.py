"""ctypes binding of libddp_b200.so (C ABI: include/ddp_b200.h).

There is no CPU fallback: if the library is missing, or a kernel entry point is reached without
a CUDA device, the caller gets a RuntimeError that says so.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

_PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_PKG, "libddp_b200.so")
CSRC = os.path.join(_PKG, "csrc")
HEADER = os.path.join(os.path.dirname(_PKG), "include", "ddp_b200.h")

ABI_VERSION = 2
F32, BF16 = 0, 1
TABLE_HOST, TABLE_DEVICE = 0, 1
NORM_EXACT_F32, NORM_TORCH_BF16 = 0, 1
ADAMW_ROUND_CLIPPED_GRAD, ADAMW_STORE_GRAD, ADAMW_MAXIMIZE = 1, 2, 4
TILE_ELEMS = 2048
STATIC_SCRATCH_MAX_CHUNKS = 262144


class Seg(ctypes.Structure):
    """struct ddpb200_seg"""
    _fields_ = [("ptr", ctypes.c_void_p * 4), ("numel", ctypes.c_int64), ("chunk_begin", ctypes.c_int64)]


assert ctypes.sizeof(Seg) == 48
MAX_PEERS = 8


class PeerMap(ctypes.Structure):
    """struct ddpb200_peer_map"""
    _fields_ = [("arena", ctypes.c_void_p * MAX_PEERS), ("partials", ctypes.c_void_p * MAX_PEERS),
                ("arena_multicast", ctypes.c_void_p), ("rank", ctypes.c_int32), ("world", ctypes.c_int32)]


_vp, _i64, _i32, _u32, _f32, _f64 = (ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_uint32,
                                     ctypes.c_float, ctypes.c_double)

# name -> (restype, argtypes); the one place the Python side states the ABI
SIGNATURES = {
    "ddpb200_abi_version": (ctypes.c_int, []),
    "ddpb200_error_string": (ctypes.c_char_p, [ctypes.c_int]),
    "ddpb200_sm_count": (ctypes.c_int, []),
    "ddpb200_bucket_assign": (ctypes.c_int, [_vp, _vp, _vp, _vp, _vp, _i64, _vp, _i64, _vp, _vp, _vp, _vp]),
    "ddpb200_bucket_layout": (ctypes.c_int, [_vp, _vp, _vp, _i64, _vp, _vp]),
    "ddpb200_plan_chunks": (ctypes.c_int, [_vp, _i64, _i64, _vp]),
    "ddpb200_suggest_chunk_elems": (_i64, [_i64, _i32]),
    "ddpb200_pack_scale_cast": (ctypes.c_int, [_vp, _i32, _i64, _i64, _i64, _i32, _i32, _f32, _vp, _vp, _vp, _vp]),
    "ddpb200_l2norm_sumsq": (ctypes.c_int, [_vp, _i32, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp]),
    "ddpb200_l2norm_sumsq_static": (ctypes.c_int, [_vp, _i32, _i64, _i64, _i64, _i32, _vp, _vp]),
    "ddpb200_l2norm_finalize": (ctypes.c_int, [_vp, _i64, _i32, _f32, _vp, _vp]),
    "ddpb200_scale_by_device_scalar": (ctypes.c_int, [_vp, _i32, _i64, _i64, _i64, _i32, _vp, _vp]),
    "ddpb200_fused_clip_adamw": (ctypes.c_int, [_vp, _i32, _i64, _i64, _i64, _i32, _i32, _i32, _f64, _vp,
                                                _f64, _f64, _f64, _f64, _i64, _vp, _vp, _vp, _vp, _u32, _vp]),
    "ddpb200_adamw_bias_corrections": (ctypes.c_int, [_f64, _f64, _i64, _vp]),
    "ddpb200_allreduce_sumsq_chunks": (ctypes.c_int, [_vp, _i64, _i64, _i64, _i32, _vp, _i32, _f32, _vp]),
    "ddpb200_fold_chunk_partials": (ctypes.c_int, [_vp, _i64, _i64, _vp, _vp, _vp]),
    "ddpb200_reduce_scatter_sumsq_chunks": (ctypes.c_int, [_vp, _i64, _i64, _i64, _i32, _vp, _i32, _f32, _vp]),
    "ddpb200_sharded_adamw_allgather": (ctypes.c_int, [_vp, _i64, _i64, _i64, _i32, _vp, _vp, _vp, _vp, _f64, _vp,
                                                       _f64, _f64, _f64, _f64, _i64, _vp, _u32, _i32, _vp]),
    "ddpb200_selfcheck_sqrt": (ctypes.c_int, [_u32, ctypes.c_uint64, _f32, _f32, _vp, _vp]),
    "ddpb200_selfcheck_div": (ctypes.c_int, [ctypes.c_uint64, ctypes.c_uint64, _vp, _vp]),
    "ddpb200_selfcheck_div_by_const": (ctypes.c_int, [_vp, _i32, _vp, _vp]),
}


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a with nvcc (csrc/Makefile) into libddp_b200.so, in tree."""
    cmd = ["make", "-C", CSRC, "-j4"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if verbose or res.returncode != 0:
        print(res.stdout)
    if res.returncode != 0:
        raise RuntimeError("building libddp_b200.so failed (nvcc, sm_100a); see output above")
    return LIB_PATH


_lib = None


def lib() -> ctypes.CDLL:
    """The loaded library. Raises if it has not been built — never falls back to anything."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a). This package has no CPU or eager fallback.")
        L = ctypes.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)  # AttributeError here = header and library disagree
            fn.restype = res
            fn.argtypes = args
        if L.ddpb200_abi_version() != ABI_VERSION:
            raise RuntimeError("libddp_b200.so ABI version mismatch")
        _lib = L
    return _lib


def check(code: int, what: str) -> None:
    if code != 0:
        msg = lib().ddpb200_error_string(code).decode()
        raise RuntimeError(f"{what} failed: {msg} (code {code})")
