"""TEST-ONLY kernel backend: the interface of ddp_b200.kernels.CudaKernels answered by the CPU
oracle (oracle/ddp_oracle.c) on CPU tensors. It exists so that the host logic of
DistributedDataParallel / AdamW / clip_grad_norm_ (bucket scheduling, rebuild, hooks, table
handling) can be driven under gloo with world_size 2 on a machine without a GPU. It is never
imported by the product package."""
from __future__ import annotations

import numpy as np
import torch

import oracle
from ddp_b200.kernels import SegTable, dtype_code


class OracleKernels:
    name = "oracle-cpu (tests only)"

    def __init__(self):
        self.calls = {"pack": 0, "sumsq": 0, "finalize": 0, "scale_by": 0, "adamw": 0}

    def table(self, cols, numels, chunk_elems=None, device=None):
        return SegTable(cols, numels, chunk_elems)

    @staticmethod
    def _rows(tab, first, count):
        count = tab.n - first if count is None else count
        return [(tab.host["ptr"][r], int(tab.host["numel"][r])) for r in range(first, first + count)]

    def pack(self, tab, src_dtype, dst_dtype, scale, first=0, count=None, sumsq_out=None, device=None,
             prefer_device=True):
        self.calls["pack"] += 1
        L = oracle.lib()
        for i, (ptr, n) in enumerate(self._rows(tab, first, count)):
            L.oracle_scale_cast(int(ptr[0]), dtype_code(src_dtype), int(ptr[1]), dtype_code(dst_dtype), n,
                                np.float32(scale))
            if sumsq_out is not None:
                sumsq_out[i] = L.oracle_sumsq(int(ptr[1]), dtype_code(dst_dtype), n)

    def sumsq(self, tab, dtype, out, first=0, count=None, prefer_device=True):
        self.calls["sumsq"] += 1
        L = oracle.lib()
        for i, (ptr, n) in enumerate(self._rows(tab, first, count)):
            out[i] = L.oracle_sumsq(int(ptr[1]), dtype_code(dtype), n)

    def norm_finalize(self, sumsq, max_norm, torch_bf16, out=None):
        self.calls["finalize"] += 1
        total, coef = oracle.total_norm_coef(sumsq.double().numpy(), max_norm, torch_bf16)
        if out is None:
            out = torch.empty(2, dtype=torch.float32)
        out[0], out[1] = total, coef
        return out

    def scale_by(self, tab, dtype, coef, first=0, count=None, prefer_device=True):
        self.calls["scale_by"] += 1
        L = oracle.lib()
        for ptr, n in self._rows(tab, first, count):
            L.oracle_scale_inplace(int(ptr[1]), dtype_code(dtype), n, np.float32(float(coef[0])))

    def adamw(self, tab, param_dtype, grad_dtype, state_dtype, *, lr, beta1, beta2, eps, weight_decay, step,
              clip_coef=None, grad_scale=None, found_inf=None, round_clipped=True, store_grad=False,
              maximize=False, prefer_device=True, step_tensor=None):
        self.calls["adamw"] += 1
        L = oracle.lib()
        flags = (1 if round_clipped else 0) | (2 if store_grad else 0) | (4 if maximize else 0)
        for ptr, n in self._rows(tab, 0, None):
            L.oracle_fused_clip_adamw(int(ptr[0]), dtype_code(param_dtype), int(ptr[1]), dtype_code(grad_dtype),
                                      int(ptr[2]), int(ptr[3]), dtype_code(state_dtype), n, float(lr), beta1,
                                      beta2, eps, weight_decay, int(step), int(clip_coef is not None),
                                      np.float32(float(clip_coef[0]) if clip_coef is not None else 0.0),
                                      int(grad_scale is not None),
                                      np.float32(float(grad_scale) if grad_scale is not None else 1.0), flags)
