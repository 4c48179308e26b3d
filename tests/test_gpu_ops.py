"""The torch.library layer (ddp_b200/ops.py) on the GPU: every op passes torch.library.opcheck (schema,
fake implementation, dispatcher registration), gives the same bits as the direct C-ABI call it wraps,
shows up under torch.profiler by name, and traces without a graph break."""
import numpy as np
import pytest
import torch

import ddp_b200
import oracle

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rand(n, dt, scale, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n, generator=g) * scale).to(dt).to(DEV)


SIZES = [5, 2048, 4097, 70001]


@pytest.mark.parametrize("dt", [torch.float32, torch.bfloat16])
def test_ops_pass_opcheck(dt):
    ops = torch.ops.ddp_b200
    grads = [_rand(n, dt, 0.01, i) for i, n in enumerate(SIZES)]
    ss = torch.zeros(len(SIZES), dtype=torch.float32, device=DEV)
    torch.library.opcheck(ops.l2norm_sumsq.default, (grads, ss))
    torch.library.opcheck(ops.l2norm_finalize.default, (ss, 1.0, dt == torch.bfloat16))
    dst = [torch.empty_like(g) for g in grads]
    torch.library.opcheck(ops.pack_scale_cast.default, (grads, dst, 0.125))
    torch.library.opcheck(ops.pack_scale_cast_sumsq.default, (grads, dst, 0.125, ss))
    coef = torch.tensor([0.5], dtype=torch.float32, device=DEV)
    torch.library.opcheck(ops.scale_by_.default, ([g.clone() for g in grads], coef))
    p = [_rand(n, dt, 0.02, 10 + i) for i, n in enumerate(SIZES)]
    m = [torch.zeros_like(x) for x in p]
    v = [torch.zeros_like(x) for x in p]
    torch.library.opcheck(ops.fused_clip_adamw_.default, (p, grads, m, v, 1e-3, 0.9, 0.999, 1e-2, 1e-8, 1),
                          {"clip_coef": coef})


def test_ops_equal_direct_calls_and_oracle():
    ops = torch.ops.ddp_b200
    K = ddp_b200.CudaKernels()
    dt = torch.bfloat16
    grads = [_rand(n, dt, 0.01, i) for i, n in enumerate(SIZES)]
    ss_op = torch.zeros(len(SIZES), dtype=torch.float32, device=DEV)
    ss_k = torch.zeros_like(ss_op)
    ops.l2norm_sumsq(grads, ss_op)
    K.sumsq(K.table([None, grads, None, None], SIZES), dt, ss_k, prefer_device=False)
    assert torch.equal(ss_op, ss_k)
    out = ops.l2norm_finalize(ss_op, 0.05, True)
    assert torch.equal(out, K.norm_finalize(ss_k, 0.05, True))
    # fused clip + AdamW through the op, three steps, against the oracle bit for bit
    p = [_rand(n, dt, 0.02, 10 + i) for i, n in enumerate(SIZES)]
    m = [torch.zeros_like(x) for x in p]
    v = [torch.zeros_like(x) for x in p]

    def bits(t):
        return t.detach().cpu().view(torch.int16).numpy().view(np.uint16).copy()
    pn, mn, vn = [bits(x) for x in p], [bits(x) for x in m], [bits(x) for x in v]
    hp = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2)
    coef = out[1:2]
    for step in (1, 2, 3):
        ops.fused_clip_adamw_(p, grads, m, v, hp["lr"], hp["beta1"], hp["beta2"], hp["weight_decay"], hp["eps"], step,
                              clip_coef=coef)
        for i in range(len(SIZES)):
            oracle.fused_clip_adamw(pn[i], bits(grads[i]), mn[i], vn[i], step=step, clip_coef=float(coef), **hp)
    torch.cuda.synchronize()
    for i in range(len(SIZES)):
        assert np.array_equal(bits(p[i]), pn[i]) and np.array_equal(bits(m[i]), mn[i]) and np.array_equal(bits(v[i]), vn[i])


def test_public_wrappers_go_through_the_ops_and_profile_by_name():
    ps = [torch.nn.Parameter(_rand(n, torch.bfloat16, 0.02, i)) for i, n in enumerate(SIZES)]
    for i, q in enumerate(ps):
        q.grad = _rand(q.numel(), torch.bfloat16, 0.01, 50 + i)
    opt = ddp_b200.AdamW(ps, lr=1e-3)
    with torch.profiler.profile(activities=[torch.profiler.ProfilerActivity.CPU, torch.profiler.ProfilerActivity.CUDA]) as prof:
        ddp_b200.clip_grad_norm_(ps, 1.0, fused_optimizer=opt)
        opt.step()
        torch.cuda.synchronize()
    names = {e.key for e in prof.key_averages()}
    for want in ("ddp_b200::l2norm_sumsq", "ddp_b200::l2norm_finalize", "ddp_b200::fused_clip_adamw_"):
        assert want in names, (want, sorted(n for n in names if "ddp" in n))
    # the kernels ran under those ops
    assert any("adamw" in n and "kernel" in n for n in names) and any("stream_kernel" in n for n in names)


def test_ops_trace_without_graph_break():
    ops = torch.ops.ddp_b200
    dt = torch.float32
    p = [_rand(n, dt, 0.02, 10 + i) for i, n in enumerate(SIZES)]
    g = [_rand(n, dt, 0.01, i) for i, n in enumerate(SIZES)]
    m = [torch.zeros_like(x) for x in p]
    v = [torch.zeros_like(x) for x in p]
    p2, m2, v2 = [x.clone() for x in p], [x.clone() for x in m], [x.clone() for x in v]

    def tail(p, g, m, v, ss):
        ops.l2norm_sumsq(g, ss)
        out = ops.l2norm_finalize(ss, 0.05, False)
        ops.fused_clip_adamw_(p, g, m, v, 1e-3, 0.9, 0.999, 1e-2, 1e-8, 1, clip_coef=out[1:2])
        return out[0]

    ss = torch.zeros(len(SIZES), device=DEV)
    want = tail(p, g, m, v, ss)
    compiled = torch.compile(tail, backend="aot_eager", fullgraph=True)
    got = compiled(p2, g, m2, v2, torch.zeros_like(ss))
    torch.cuda.synchronize()
    assert torch.equal(got, want)
    for a, b in zip(p, p2):
        assert torch.equal(a, b)
