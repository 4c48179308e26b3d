"""CPU checks of the C-ABI library: it loads, exports every symbol include/ddp_b200.h declares,
and its host-only entry points (bucket planner, chunk planning, bias corrections) agree with
torch, the golden fixtures and the oracle. No kernel is launched here."""
import ctypes
import json
import os
import re
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist

import ddp_b200
import oracle
from ddp_b200 import _lib


def test_library_exports_every_declared_symbol(native_lib):
    header = open(_lib.HEADER).read()
    declared = set(re.findall(r"\b(ddpb200_\w+)\s*\(", header))
    declared.discard("ddpb200_seg")
    assert len(declared) >= 13
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in sorted(declared):
        assert hasattr(raw, name), f"{name} declared in include/ddp_b200.h but not exported"
    assert declared == set(_lib.SIGNATURES), "python binding and header disagree"
    assert native_lib.ddpb200_abi_version() == _lib.ABI_VERSION == 2
    assert native_lib.ddpb200_error_string(0) == b"ok"
    assert b"invalid" in native_lib.ddpb200_error_string(-1)


def test_seg_struct_layout():
    assert ctypes.sizeof(_lib.Seg) == 48
    assert ddp_b200.kernels.SEG_DTYPE.itemsize == 48
    assert ddp_b200.kernels.SEG_DTYPE.fields["numel"][1] == 32
    assert ddp_b200.kernels.SEG_DTYPE.fields["chunk_begin"][1] == 40


def test_planner_matches_golden_and_oracle(native_lib, golden_dir):
    cases = json.load(open(os.path.join(golden_dir, "bucket_layouts.json")))
    dts = {4: torch.float32, 2: torch.bfloat16}
    for c in cases:
        tensors = [torch.empty(n, dtype=dts[e], device="meta") for n, e in zip(c["numel"], c["elem_size"])]
        order = c["order"]
        if order is not None:
            tensors = [tensors[i] for i in order]
        got, lim = ddp_b200.compute_bucket_assignment_by_size(tensors, c["limits"], c["sparse"], order)
        assert got == c["buckets"], c["name"]
        assert lim == c["bucket_limits"], c["name"]


def test_planner_matches_live_torch_randomised(native_lib):
    rng = np.random.default_rng(5)
    dts = [torch.float32, torch.bfloat16, torch.float64]
    for trial in range(60):
        n = int(rng.integers(1, 70))
        tensors = [torch.empty(int(rng.integers(0, 9000)), dtype=dts[int(rng.integers(0, 3 if trial % 3 == 0 else 1))],
                               device="meta") for _ in range(n)]
        limits = sorted(int(x) for x in rng.integers(1, 60000, size=int(rng.integers(1, 4))))
        sparse = [bool(x) for x in rng.random(n) < 0.1] if trial % 4 == 0 else [False] * n
        use_order = trial % 2 == 1
        order = [int(x) for x in rng.permutation(n)] if use_order else None
        if use_order:
            want, wl = dist._compute_bucket_assignment_by_size(tensors, limits, sparse, order)
            got, gl = ddp_b200.compute_bucket_assignment_by_size(tensors, limits, sparse, order)
        else:
            want, wl = dist._compute_bucket_assignment_by_size(tensors, limits, sparse)
            got, gl = ddp_b200.compute_bucket_assignment_by_size(tensors, limits, sparse)
        assert got == [list(b) for b in want], trial
        assert gl == list(wl), trial


def test_planner_init_single_bucket_and_maxsize(native_lib):
    tensors = [torch.empty(1000, device="meta") for _ in range(5)]
    got, lim = ddp_b200.compute_bucket_assignment_by_size(tensors, [sys.maxsize])
    assert got == [[0, 1, 2, 3, 4]] and lim == [sys.maxsize]
    with pytest.raises(ValueError):
        ddp_b200.compute_bucket_assignment_by_size([], [1024])


def test_layout_offsets_no_padding(native_lib):
    # SURVEY.md §8 a-1 probe: four 256-element fp32 tensors in one bucket sit at bytes 0,1024,2048,3072
    specs = ddp_b200.layout_buckets([256] * 4, [[0, 1, 2, 3]], [4096])
    assert specs[0].offsets == [0, 256, 512, 768] and specs[0].numel == 1024
    specs = ddp_b200.layout_buckets([7, 0, 13, 5], [[3, 1], [2, 0]])
    assert [s.offsets for s in specs] == [[0, 5], [0, 13]]
    assert [s.numel for s in specs] == [5, 20]


def test_plan_chunks_and_slicing(native_lib):
    t = ddp_b200.SegTable([None] * 4, [5000, 0, 2048, 1, 70000], chunk_elems=2048)
    assert list(t.host["chunk_begin"]) == [0, 3, 3, 4, 5]
    assert t.total_chunks == 5 + 35
    # the C helper computes the same prefix sums
    segs = (_lib.Seg * 5)()
    for i, n in enumerate([5000, 0, 2048, 1, 70000]):
        segs[i].numel = n
    total = ctypes.c_int64()
    assert native_lib.ddpb200_plan_chunks(ctypes.addressof(segs), 5, 2048, ctypes.addressof(total)) == 0
    assert [s.chunk_begin for s in segs] == [0, 3, 3, 4, 5] and total.value == 40
    assert native_lib.ddpb200_plan_chunks(ctypes.addressof(segs), 5, 1000, ctypes.addressof(total)) == -1
    _, _, n, chunks = t.args(2, 2)
    assert (n, chunks) == (2, 2)


def test_suggest_chunk_elems(native_lib):
    f = native_lib.ddpb200_suggest_chunk_elems
    assert f(8_000_000_000, 148) == 65536
    assert f(10_000_000, 148) == 2048
    for total in (1, 10 ** 6, 10 ** 8, 10 ** 9):
        c = f(total, 148)
        assert c % 2048 == 0 and 2048 <= c <= 65536


def test_bias_corrections_match_oracle(native_lib):
    out = np.zeros(2, dtype=np.float32)
    for step in (1, 2, 10, 1000, 100000):
        for b1, b2 in ((0.9, 0.999), (0.8, 0.95), (0.0, 0.5)):
            assert native_lib.ddpb200_adamw_bias_corrections(b1, b2, step, out.ctypes.data) == 0
            want = oracle.bias_corrections(b1, b2, step)
            assert (float(out[0]), float(out[1])) == want
    assert native_lib.ddpb200_adamw_bias_corrections(0.9, 0.999, 0, out.ctypes.data) == -1   # step >= 1


def test_invalid_arguments_are_rejected_without_a_gpu(native_lib):
    # argument validation happens before any CUDA call, so these are safe on a CPU-only box
    assert native_lib.ddpb200_fused_clip_adamw(None, 0, 3, 3, 2048, 0, 0, 0, 1e-3, None, 0.9, 0.999, 1e-8, 0.01, 1,
                                               None, None, None, None, 0, None) == -1
    assert native_lib.ddpb200_fused_clip_adamw(None, 0, 0, 0, 1000, 0, 0, 0, 1e-3, None, 0.9, 0.999, 1e-8, 0.01, 1,
                                               None, None, None, None, 0, None) == -1   # chunk not a multiple of 2048
    assert native_lib.ddpb200_l2norm_finalize(None, 5, 0, 1.0, None, None) == -1
    assert native_lib.ddpb200_pack_scale_cast(None, 0, 0, 0, 2048, 7, 0, 1.0, None, None, None, None) == -2  # dtype 7


def test_product_has_no_cpu_fallback(native_lib):
    k = ddp_b200.CudaKernels()
    g = torch.zeros(10)
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        k.sumsq(k.table([None, [g], None, None], [10]), torch.float32, torch.zeros(1))
    # and the product package never imports the oracle
    pkg_dir = os.path.dirname(ddp_b200.__file__)
    for root, _, files in os.walk(pkg_dir):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh")):
                src = open(os.path.join(root, f)).read()
                assert "import oracle" not in src and "from oracle" not in src and "libddp_oracle" not in src, f


def test_torch_library_ops_are_registered_with_schemas_and_fake_impls(native_lib):
    """The dispatcher-facing layer (ddp_b200/ops.py) without a GPU: every op exists with the schema the docs
    state, its fake implementation runs under FakeTensorMode (what torch.compile traces with), and a CPU
    tensor is refused by the dispatcher — the ops are registered for CUDA only, there is no CPU kernel."""
    from torch._subclasses.fake_tensor import FakeTensorMode
    ops = torch.ops.ddp_b200
    want = {
        "pack_scale_cast": "(Tensor[] src, Tensor(a1!)[] dst, float scale) -> ()",
        "pack_scale_cast_sumsq": "(Tensor[] src, Tensor(a1!)[] dst, float scale, Tensor(a3!) sumsq_out) -> ()",
        "l2norm_sumsq": "(Tensor[] grads, Tensor(a1!) sumsq_out) -> ()",
        "l2norm_finalize": "(Tensor sumsq, float max_norm, bool torch_bf16) -> Tensor",
        "scale_by_": "(Tensor(a0!)[] grads, Tensor coef) -> ()",
    }
    for name, sig in want.items():
        assert str(getattr(ops, name).default._schema) == f"ddp_b200::{name}{sig}"
    adam = str(ops.fused_clip_adamw_.default._schema)
    assert adam.startswith("ddp_b200::fused_clip_adamw_(Tensor(a0!)[] params, Tensor(a1!)[] grads, Tensor(a2!)[] exp_avgs, "
                           "Tensor(a3!)[] exp_avg_sqs, float lr, float beta1, float beta2, float weight_decay, float eps, SymInt step")
    assert "Tensor? clip_coef=None, Tensor? grad_scale=None, Tensor? found_inf=None" in adam
    assert "Tensor(a0!) arena, Tensor[] views, Tensor(a2!) sumsq_out" in str(ops.allreduce_sumsq_.default._schema)
    with FakeTensorMode():
        g = [torch.empty(10, device="cuda", dtype=torch.bfloat16), torch.empty(5, device="cuda", dtype=torch.bfloat16)]
        ss = torch.empty(2, device="cuda")
        ops.l2norm_sumsq(g, ss)
        out = ops.l2norm_finalize(ss, 1.0, True)
        assert out.shape == (2,) and out.dtype == torch.float32 and out.device.type == "cuda"
        m = [torch.zeros_like(x) for x in g]
        ops.fused_clip_adamw_([x.clone() for x in g], g, m, [x.clone() for x in m], 1e-3, 0.9, 0.999, 0.01, 1e-8, 1,
                              clip_coef=out[1:2])
        ops.pack_scale_cast(g, [torch.empty_like(x) for x in g], 0.5)
    with pytest.raises(NotImplementedError):
        ops.l2norm_sumsq([torch.zeros(4)], torch.zeros(1))
