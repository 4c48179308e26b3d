"""CPU oracle for the DDP grad-sync + clip + AdamW hot path — TEST INFRASTRUCTURE ONLY.

Thin numpy/ctypes front for ``ddp_oracle.c`` (see that file's header for what each function
restates, with torch file:line). The product package never imports this module; only
``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s CPU-baseline legs do.

bf16 data is carried as ``numpy.uint16`` bit patterns, fp32 as ``numpy.float32``.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

F32, BF16 = 0, 1
_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libddp_oracle.so")


def build(force: bool = False) -> str:
    """Compile ddp_oracle.c with gcc (the recipe is oracle/Makefile)."""
    src = os.path.join(_HERE, "ddp_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libddp_oracle.so"], stdout=subprocess.DEVNULL)
    return _SO


_lib = None


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        L = ctypes.CDLL(build())
        c = ctypes
        vp, i64, i32, f32, f64, u32 = c.c_void_p, c.c_int64, c.c_int32, c.c_float, c.c_double, c.c_uint32
        L.oracle_round_bf16.restype = f32
        L.oracle_round_bf16.argtypes = [f32]
        L.oracle_bucket_assign.restype = i64
        L.oracle_bucket_assign.argtypes = [vp, vp, vp, vp, vp, i64, vp, i64, vp, vp]
        L.oracle_scale_cast.restype = None
        L.oracle_scale_cast.argtypes = [vp, i32, vp, i32, i64, f32]
        L.oracle_sumsq.restype = f64
        L.oracle_sumsq.argtypes = [vp, i32, i64]
        L.oracle_total_norm_coef.restype = None
        L.oracle_total_norm_coef.argtypes = [vp, i64, i32, f32, vp]
        L.oracle_scale_inplace.restype = None
        L.oracle_scale_inplace.argtypes = [vp, i32, i64, f32]
        L.oracle_bias_corrections.restype = None
        L.oracle_bias_corrections.argtypes = [f64, f64, i64, vp]
        L.oracle_fused_clip_adamw.restype = None
        L.oracle_fused_clip_adamw.argtypes = [vp, i32, vp, i32, vp, vp, i32, i64, f64, f64, f64, f64,
                                              f64, i64, i32, f32, i32, f32, u32]
        L.oracle_single_tensor_adamw.restype = None
        L.oracle_single_tensor_adamw.argtypes = [vp, i32, vp, i32, vp, vp, i32, i64, f64, f64, f64,
                                                 f64, f64, i64]
        L.oracle_div_by_const_mismatches.restype = i64
        L.oracle_div_by_const_mismatches.argtypes = [f32]
        _lib = L
    return _lib


def _dt(a: np.ndarray) -> int:
    if a.dtype == np.float32:
        return F32
    if a.dtype == np.uint16:
        return BF16
    raise TypeError(f"oracle arrays are float32 or uint16 (bf16 bits), got {a.dtype}")


def _p(a: np.ndarray) -> int:
    assert a.flags["C_CONTIGUOUS"]
    return a.ctypes.data


# ---- bf16 helpers -------------------------------------------------------------------------
def f32_to_bf16_bits(x: np.ndarray) -> np.ndarray:
    """Round-to-nearest-even fp32 -> bf16 bit patterns (uint16)."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    out = np.empty(x.shape, dtype=np.uint16)
    lib().oracle_scale_cast(_p(x), F32, _p(out), BF16, x.size, 1.0)
    return out


def bf16_bits_to_f32(b: np.ndarray) -> np.ndarray:
    b = np.ascontiguousarray(b, dtype=np.uint16)
    return (b.astype(np.uint32) << 16).view(np.float32)


# ---- a-1 ------------------------------------------------------------------------------------
def bucket_assign(numel, elem_size, limits, keys=None, expect_sparse=None, tensor_indices=None):
    """-> (list of buckets, each a list of tensor indices; list of per-bucket limits)."""
    n = len(numel)
    numel_a = np.asarray(numel, dtype=np.int64)
    es_a = np.asarray(elem_size, dtype=np.int32)
    key_a = None if keys is None else np.asarray(keys, dtype=np.int64)
    sp_a = None if expect_sparse is None else np.asarray(expect_sparse, dtype=np.uint8)
    ti_a = None if tensor_indices is None else np.asarray(tensor_indices, dtype=np.int64)
    lim_a = np.asarray([min(int(x), 2**64 - 1) for x in limits], dtype=np.uint64)
    bucket_of = np.empty(n, dtype=np.int64)
    blim = np.zeros(n, dtype=np.uint64)
    nb = lib().oracle_bucket_assign(_p(numel_a), _p(es_a), None if key_a is None else _p(key_a),
                                    None if sp_a is None else _p(sp_a),
                                    None if ti_a is None else _p(ti_a), n, _p(lim_a), len(lim_a),
                                    _p(bucket_of), _p(blim))
    if nb < 0:
        raise ValueError("oracle_bucket_assign: bad input")
    buckets = [[] for _ in range(nb)]
    for i in range(n):
        ti = int(ti_a[i]) if ti_a is not None else i
        buckets[int(bucket_of[i])].append(ti)
    return buckets, [int(x) for x in blim[:nb]]


# ---- a-2 / a-4 ---------------------------------------------------------------------------------
def scale_cast(src: np.ndarray, dst_dtype, scale: float) -> np.ndarray:
    src = np.ascontiguousarray(src)
    dst = np.empty(src.shape, dtype=dst_dtype)
    lib().oracle_scale_cast(_p(src), _dt(src), _p(dst), _dt(dst), src.size, np.float32(scale))
    return dst


# ---- a-5 ---------------------------------------------------------------------------------------
def sumsq(g: np.ndarray) -> float:
    g = np.ascontiguousarray(g)
    return float(lib().oracle_sumsq(_p(g), _dt(g), g.size))


def total_norm_coef(sumsqs, max_norm: float, torch_bf16: bool):
    a = np.asarray(sumsqs, dtype=np.float64)
    out = np.zeros(2, dtype=np.float32)
    lib().oracle_total_norm_coef(_p(a), a.size, int(torch_bf16), np.float32(max_norm), _p(out))
    return float(out[0]), float(out[1])


def scale_inplace(g: np.ndarray, coef: float) -> None:
    lib().oracle_scale_inplace(_p(g), _dt(g), g.size, np.float32(coef))


# ---- a-6 ----------------------------------------------------------------------------------------
def bias_corrections(beta1: float, beta2: float, step: int):
    out = np.zeros(2, dtype=np.float32)
    lib().oracle_bias_corrections(beta1, beta2, step, _p(out))
    return float(out[0]), float(out[1])


def fused_clip_adamw(p, g, m, v, *, lr, beta1, beta2, eps, weight_decay, step, clip_coef=None,
                     grad_scale=None, round_clipped=True, store_grad=False, maximize=False) -> None:
    """In place on numpy arrays p, m, v (and g if store_grad)."""
    assert _dt(m) == _dt(v)
    flags = (1 if round_clipped else 0) | (2 if store_grad else 0) | (4 if maximize else 0)
    lib().oracle_fused_clip_adamw(_p(p), _dt(p), _p(g), _dt(g), _p(m), _p(v), _dt(m), p.size,
                                  lr, beta1, beta2, eps, weight_decay, step,
                                  int(clip_coef is not None), np.float32(clip_coef or 0.0),
                                  int(grad_scale is not None), np.float32(grad_scale or 1.0), flags)


def single_tensor_adamw(p, g, m, v, *, lr, beta1, beta2, eps, weight_decay, step) -> None:
    lib().oracle_single_tensor_adamw(_p(p), _dt(p), _p(g), _dt(g), _p(m), _p(v), _dt(m), p.size,
                                     lr, beta1, beta2, eps, weight_decay, step)


def div_by_const_mismatches(c: float) -> int:
    """How many of the 2^23 significands x give fma-sequence(x, c) != x / c (see ddp_oracle.c)."""
    return int(lib().oracle_div_by_const_mismatches(np.float32(c)))
