"""GPU parity: the sm_100a kernels, called through the C ABI (libddp_b200.so via ctypes), against
the CPU oracle on the same seeded inputs (bit-exact where the arithmetic is integer/elementwise,
stated tolerance for the one reduction), against the committed golden vectors from torch, against
torch's own CUDA ops on the box, and through size-independent properties at full sizes."""
import json
import os

import numpy as np
import pytest
import torch

import ddp_b200
import oracle

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
TDT = {"f32": torch.float32, "bf16": torch.bfloat16}


def to_np(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    return t.numpy().copy()


def from_np(a: np.ndarray) -> torch.Tensor:
    if a.dtype == np.uint16:
        return torch.from_numpy(a.view(np.int16).copy()).view(torch.bfloat16).to(DEV)
    return torch.from_numpy(a.copy()).to(DEV)


def rand(n, dt, scale, seed):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(n, generator=g) * scale).to(dt).to(DEV)


def bits_equal(a: torch.Tensor, b_np: np.ndarray) -> bool:
    return np.array_equal(to_np(a).view(np.uint16 if a.dtype == torch.bfloat16 else np.uint32),
                          b_np.view(np.uint16 if b_np.dtype == np.uint16 else np.uint32))


@pytest.fixture(scope="module")
def K():
    assert torch.cuda.is_available()
    return ddp_b200.CudaKernels()


SIZES = [0, 1, 7, 2047, 2048, 2049, 4096 + 5, 70001, 300000]


# ---- kernel (a): pack / unpack ---------------------------------------------------------------
@pytest.mark.parametrize("sdt,ddt", [("f32", "f32"), ("bf16", "bf16"), ("f32", "bf16"), ("bf16", "f32")])
@pytest.mark.parametrize("ws", [1, 2, 3, 8])
@pytest.mark.parametrize("loc", ["host", "device"])
def test_pack_scale_cast_bit_exact(K, sdt, ddt, ws, loc):
    srcs = [rand(n, TDT[sdt], 0.01, 10 + i) for i, n in enumerate(SIZES)]
    total = sum(SIZES)
    bucket = torch.full((total + 3,), 7.0, dtype=TDT[ddt], device=DEV)     # +3: canary past the end
    offs = np.cumsum([0] + SIZES[:-1])
    views = [bucket[o:o + n] for o, n in zip(offs, SIZES)]                 # no padding: some misaligned
    scale = float(np.float32(1.0 / ws))
    tab = K.table([srcs, views, None, None], SIZES, chunk_elems=2048, device=DEV if loc == "device" else None)
    sumsq = torch.empty(len(SIZES), dtype=torch.float32, device=DEV)
    K.pack(tab, TDT[sdt], TDT[ddt], scale, sumsq_out=sumsq, prefer_device=(loc == "device"))
    torch.cuda.synchronize()
    for s, v in zip(srcs, views):
        want = oracle.scale_cast(to_np(s), np.uint16 if ddt == "bf16" else np.float32, scale)
        assert bits_equal(v, want)
    assert torch.all(bucket[total:] == 7.0)
    # fused sum of squares of what was stored (fp32 accumulation vs the oracle's double)
    want_ss = np.array([oracle.sumsq(to_np(v)) for v in views])
    assert np.allclose(sumsq.cpu().numpy(), want_ss, rtol=2e-6, atol=0)
    # deterministic: same bits on a second launch
    sumsq2 = torch.empty_like(sumsq)
    K.pack(tab, TDT[sdt], TDT[ddt], scale, sumsq_out=sumsq2, prefer_device=(loc == "device"))
    assert torch.equal(sumsq, sumsq2)


def test_pack_matches_golden_torch_mul(K, golden_dir):
    z = np.load(os.path.join(golden_dir, "pack_scale.npz"))
    for dt in ("f32", "bf16"):
        x = from_np(z[f"{dt}/x"])
        for ws in (1, 2, 3, 4, 6, 8):
            out = torch.empty_like(x)
            tab = K.table([[x], [out], None, None], [x.numel()])
            K.pack(tab, x.dtype, x.dtype, float(np.float32(1.0 / ws)), prefer_device=False)
            assert bits_equal(out, z[f"{dt}/ws{ws}"]), (dt, ws)
            assert torch.equal(out, x.mul(1.0 / ws))        # torch's CUDA elementwise kernel, live
    x = from_np(z["f32/x"])
    out = torch.empty(x.numel(), dtype=torch.bfloat16, device=DEV)
    K.pack(K.table([[x], [out], None, None], [x.numel()]), torch.float32, torch.bfloat16, 1.0, prefer_device=False)
    assert bits_equal(out, z["f32_to_bf16"])


def test_pack_in_place_and_many_segments(K):
    # 1000 segments: exercises the by-value table slicing (448 per launch) and the device table
    sizes = [int(x) for x in np.random.default_rng(0).integers(0, 5000, size=1000)]
    for loc in ("host", "device"):
        src = [rand(n, torch.bfloat16, 0.02, i) for i, n in enumerate(sizes)]
        ref = [oracle.scale_cast(to_np(s), np.uint16, 0.25) for s in src]
        tab = K.table([src, src, None, None], sizes, device=DEV if loc == "device" else None)
        ss = torch.empty(len(sizes), dtype=torch.float32, device=DEV)
        K.pack(tab, torch.bfloat16, torch.bfloat16, 0.25, sumsq_out=ss, prefer_device=(loc == "device"))
        for s, r in zip(src, ref):
            assert bits_equal(s, r)
        want = np.array([oracle.sumsq(r) for r in ref])
        assert np.allclose(ss.cpu().numpy(), want, rtol=2e-6)


# ---- kernel (b): L2 norm + clip coefficient ------------------------------------------------------
@pytest.mark.parametrize("dt", ["f32", "bf16"])
def test_l2norm_against_oracle_and_torch(K, dt):
    grads = [rand(n, TDT[dt], 0.01 * (1 + i % 3), 100 + i) for i, n in enumerate(SIZES + [1 << 20])]
    sizes = [g.numel() for g in grads]
    tab = K.table([None, grads, None, None], sizes)
    ss = torch.empty(len(sizes), dtype=torch.float32, device=DEV)
    K.sumsq(tab, TDT[dt], ss, prefer_device=False)
    want = np.array([oracle.sumsq(to_np(g)) for g in grads])
    assert np.allclose(ss.cpu().numpy(), want, rtol=2e-6, atol=0)        # fp32 tree vs double
    for max_norm in (1e-3, 1.0, 1e6):
        for mode in (False, True):
            out = K.norm_finalize(ss, max_norm, torch_bf16=mode).cpu().numpy()
            t, c = oracle.total_norm_coef(ss.cpu().numpy().astype(np.float64), max_norm, mode)
            if mode:
                assert float(out[0]) == t and float(out[1]) == c            # bf16-rounded scalars: exact
            else:
                assert abs(out[0] - t) <= 2e-7 * t and abs(out[1] - c) <= 3e-7 * c
    # torch's own clip_grad_norm_ on the box: returned norm and the scaled grads
    ps = [torch.nn.Parameter(torch.zeros_like(g)) for g in grads]
    qs = [torch.nn.Parameter(torch.zeros_like(g)) for g in grads]
    for p, q, g in zip(ps, qs, grads):
        p.grad, q.grad = g.clone(), g.clone()
    want_n = torch.nn.utils.clip_grad_norm_(ps, 0.05)
    got_n = ddp_b200.clip_grad_norm_(qs, 0.05)
    assert got_n.dtype == want_n.dtype and got_n.shape == want_n.shape
    if dt == "bf16":
        assert abs(float(got_n) - float(want_n)) <= 2 ** -7 * float(want_n)      # <= 1 bf16 ulp
    else:
        assert abs(float(got_n) - float(want_n)) <= 2e-6 * float(want_n)
    for p, q in zip(ps, qs):
        tol = dict(rtol=2 ** -6, atol=1e-8) if dt == "bf16" else dict(rtol=3e-6, atol=0)
        assert torch.allclose(p.grad.float(), q.grad.float(), **tol)


def test_l2norm_edge_cases(K):
    # empty list of work, zero gradients, inf / nan propagate like torch
    z = [torch.zeros(n, device=DEV) for n in (0, 5, 4096)]
    ss = torch.full((3,), -1.0, device=DEV)
    K.sumsq(K.table([None, z, None, None], [0, 5, 4096]), torch.float32, ss, prefer_device=False)
    assert ss.tolist() == [0.0, 0.0, 0.0]
    out = K.norm_finalize(ss, 1.0, False).cpu()
    assert out[0] == 0.0 and out[1] == 1.0                                 # coef clamped to 1
    g = torch.ones(10, device=DEV)
    g[3] = float("inf")
    p = torch.nn.Parameter(torch.zeros(10, device=DEV))
    p.grad = g
    assert torch.isinf(ddp_b200.clip_grad_norm_([p], 1.0))
    with pytest.raises(RuntimeError, match="non-finite"):
        p.grad = g.clone()
        ddp_b200.clip_grad_norm_([p], 1.0, error_if_nonfinite=True)
    g2 = torch.ones(10, device=DEV)
    g2[0] = float("nan")
    p.grad = g2
    assert torch.isnan(ddp_b200.clip_grad_norm_([p], 1.0))
    assert float(ddp_b200.clip_grad_norm_([torch.nn.Parameter(torch.zeros(3, device=DEV))], 1.0)) == 0.0  # no grads


# ---- kernel (c): fused clip + AdamW --------------------------------------------------------------
HP = dict(lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2)
COMBOS = [("f32", "f32", "f32"), ("bf16", "bf16", "bf16"), ("bf16", "bf16", "f32"), ("f32", "bf16", "f32")]


@pytest.mark.parametrize("pdt,gdt,sdt", COMBOS)
@pytest.mark.parametrize("loc", ["host", "device"])
def test_fused_clip_adamw_bit_exact_vs_oracle(K, pdt, gdt, sdt, loc):
    sizes = SIZES
    p = [rand(n, TDT[pdt], 0.02, 200 + i) for i, n in enumerate(sizes)]
    m = [torch.zeros(n, dtype=TDT[sdt], device=DEV) for n in sizes]
    v = [torch.zeros(n, dtype=TDT[sdt], device=DEV) for n in sizes]
    pn, mn, vn = [to_np(x) for x in p], [to_np(x) for x in m], [to_np(x) for x in v]
    coef = torch.tensor([0.37], dtype=torch.float32, device=DEV)
    scale = torch.tensor([4.0], dtype=torch.float32, device=DEV)
    for step in range(1, 6):
        g = [rand(n, TDT[gdt], 0.01 * step, 1000 * step + i) for i, n in enumerate(sizes)]
        gn = [to_np(x) for x in g]
        use_clip, use_scale = step % 2 == 1, step == 3
        tab = K.table([p, g, m, v], sizes, chunk_elems=2048 if step < 4 else None,
                      device=DEV if loc == "device" else None)
        K.adamw(tab, TDT[pdt], TDT[gdt], TDT[sdt], step=step, clip_coef=coef if use_clip else None,
                grad_scale=scale if use_scale else None, store_grad=use_scale,
                prefer_device=(loc == "device"), **HP)
        for i in range(len(sizes)):
            oracle.fused_clip_adamw(pn[i], gn[i], mn[i], vn[i], step=step, clip_coef=0.37 if use_clip else None,
                                    grad_scale=4.0 if use_scale else None, store_grad=use_scale, **HP)
        torch.cuda.synchronize()
        for i in range(len(sizes)):
            assert bits_equal(p[i], pn[i]), (step, i, "p")
            assert bits_equal(m[i], mn[i]), (step, i, "m")
            assert bits_equal(v[i], vn[i]), (step, i, "v")
            if use_scale:
                assert bits_equal(g[i], gn[i]), (step, i, "g")


def test_fused_adamw_misaligned_segments_and_flags(K):
    """Bucket views at odd element offsets (scalar path), maximize, weight_decay=0, found_inf."""
    base = {k: torch.zeros(3 * 5000 + 8, dtype=torch.bfloat16, device=DEV) for k in "pgmv"}
    base["p"].copy_(rand(base["p"].numel(), torch.bfloat16, 0.02, 1))
    base["g"].copy_(rand(base["g"].numel(), torch.bfloat16, 0.01, 2))
    offs, n = [1, 5003, 10007], 4999
    views = {k: [t[o:o + n] for o in offs] for k, t in base.items()}
    ref = {k: [to_np(x) for x in views[k]] for k in "pgmv"}
    before = to_np(base["p"])
    hp = dict(HP, weight_decay=0.0)
    K.adamw(K.table([views["p"], views["g"], views["m"], views["v"]], [n] * 3), torch.bfloat16, torch.bfloat16,
            torch.bfloat16, step=1, maximize=True, prefer_device=False, **hp)
    for i in range(3):
        oracle.fused_clip_adamw(ref["p"][i], ref["g"][i], ref["m"][i], ref["v"][i], step=1, maximize=True, **hp)
        assert bits_equal(views["p"][i], ref["p"][i]) and bits_equal(views["v"][i], ref["v"][i])
    after = to_np(base["p"])
    gaps = np.ones(after.shape, bool)
    for o in offs:
        gaps[o:o + n] = False
    assert np.array_equal(after[gaps], before[gaps])            # nothing outside the segments touched
    # found_inf == 1 -> the launch is a no-op (GradScaler skip)
    snap = base["p"].clone()
    inf = torch.ones(1, dtype=torch.float32, device=DEV)
    K.adamw(K.table([views["p"], views["g"], views["m"], views["v"]], [n] * 3), torch.bfloat16, torch.bfloat16,
            torch.bfloat16, step=2, found_inf=inf, prefer_device=False, **hp)
    assert torch.equal(snap, base["p"])


@pytest.mark.parametrize("dt", ["bf16", "f32"])
def test_staged_adamw_alignment_and_ragged_cases(K, dt):
    """The TMA-staged variant (device table, fast modes; forced on for fp32 too): segments whose four streams
    are 16-byte aligned but ragged (the tail is read from global memory), segments where only the gradient — a
    bucket view behind an odd-sized tensor — is misaligned, segments with all streams at the same odd offset
    (neither can be staged: element by element), empty and tiny segments; bit-exact against the oracle and nothing
    outside the segments touched, over three steps."""
    import os
    tdt = TDT[dt]
    es = 2 if dt == "bf16" else 4
    lens = [0, 3, 8, 4099, 2048, 6151, 4999, 4999, 12289]
    #        p/m/v offset, g offset (elements) per segment: aligned ones are multiples of 16 bytes
    per = 16 // es
    poff = [0, per, 2 * per, 3 * per, 4 * per, 5 * per, 1, 6 * per, 7 * per]
    goff = [0, per, 2 * per, 3 * per, 4 * per, 5 * per, 1, 3, 7 * per + 1]
    stride = 16384
    base = {k: torch.zeros(stride * len(lens) + 64, dtype=tdt, device=DEV) for k in "pgmv"}
    base["p"].copy_(rand(base["p"].numel(), tdt, 0.02, 31))
    base["m"].copy_(rand(base["m"].numel(), tdt, 0.003, 32))
    base["v"].copy_(rand(base["v"].numel(), tdt, 0.01, 33).abs())
    views = {k: [base[k][i * stride + (goff[i] if k == "g" else poff[i]):][:n] for i, n in enumerate(lens)] for k in "pgmv"}
    ref = {k: [to_np(x) for x in views[k]] for k in "pmv"}
    before = {k: to_np(base[k]) for k in "pmv"}
    coef = torch.tensor([0.37], dtype=torch.float32, device=DEV)
    old = os.environ.get("DDPB200_ADAMW_BULK")
    os.environ["DDPB200_ADAMW_BULK"] = "2,2"
    try:
        for step in (1, 2, 3):
            base["g"].copy_(rand(base["g"].numel(), tdt, 0.01 * step, 40 + step))
            gn = [to_np(x) for x in views["g"]]
            tab = K.table([views["p"], views["g"], views["m"], views["v"]], lens, chunk_elems=4096, device=DEV)
            K.adamw(tab, tdt, tdt, tdt, step=step, clip_coef=coef if step != 2 else None, **HP)
            torch.cuda.synchronize()
            for i in range(len(lens)):
                oracle.fused_clip_adamw(ref["p"][i], gn[i], ref["m"][i], ref["v"][i], step=step,
                                        clip_coef=0.37 if step != 2 else None, **HP)
                for k in "pmv":
                    assert bits_equal(views[k][i], ref[k][i]), (dt, step, i, k)
    finally:
        if old is None:
            os.environ.pop("DDPB200_ADAMW_BULK", None)
        else:
            os.environ["DDPB200_ADAMW_BULK"] = old
    for k in "pmv":
        after = to_np(base[k])
        gaps = np.ones(after.shape, bool)
        for i, n in enumerate(lens):
            gaps[i * stride + poff[i]:i * stride + poff[i] + n] = False
        assert np.array_equal(after[gaps], before[k][gaps]), k      # nothing outside the segments touched


@pytest.mark.parametrize("dt", ["f32", "bf16"])
def test_clip_adamw_chain_matches_golden_torch(K, golden_dir, dt):
    """clip_grad_norm_ + AdamW through the public classes on the golden inputs vs torch's recorded
    results (default and fused CPU paths). Tolerances: fp32 2e-7 abs on |p|~0.02 (12 steps;
    torch's CPU paths multiply by float(1-lr*wd), the CUDA functor uses an fma); bf16 1e-3."""
    z = np.load(os.path.join(golden_dir, "adamw_steps.npz"))
    sizes = [int(x) for x in z["sizes"]]
    lr, b1, b2, eps, wd = [float(x) for x in z["hp"]]
    ps = [torch.nn.Parameter(from_np(z[f"{dt}/p0/{i}"])) for i in range(3)]
    opt = ddp_b200.AdamW(ps, lr=lr, betas=(b1, b2), eps=eps, weight_decay=wd)
    for s in range(int(z["steps"])):
        for i, p in enumerate(ps):
            p.grad = from_np(z[f"{dt}/g/{s}/{i}"])
        tn = ddp_b200.clip_grad_norm_(ps, float(z["max_norm"]), fused_optimizer=opt if s % 2 else None)
        want = z[f"{dt}/default/norms"][s]
        assert abs(float(tn) - want) <= (2 ** -7 if dt == "bf16" else 2e-6) * want
        opt.step()
        opt.zero_grad()
    tol_p, tol_m, tol_v = (2e-7, 5e-9, 2e-10) if dt == "f32" else (1e-3, 2.5e-4, 2e-7)
    for path in ("default", "fused"):
        for i, p in enumerate(ps):
            st = opt.state[p]
            f = lambda a: oracle.bf16_bits_to_f32(a) if a.dtype == np.uint16 else a
            assert np.max(np.abs(f(to_np(p)) - f(z[f"{dt}/{path}/p/{i}"]))) <= tol_p
            assert np.max(np.abs(f(to_np(st["exp_avg"])) - f(z[f"{dt}/{path}/m/{i}"]))) <= tol_m
            assert np.max(np.abs(f(to_np(st["exp_avg_sq"])) - f(z[f"{dt}/{path}/v/{i}"]))) <= tol_v


@pytest.mark.parametrize("dt", ["f32", "bf16"])
def test_adamw_vs_torch_fused_cuda(K, dt):
    """Against the incumbent on the same box: torch.optim.AdamW(fused=True), 20 steps.
    fp32: max-abs <= 1e-7 on |p|~0.02; bf16: <= 2 bf16 ulp (2^-7 relative)."""
    torch.manual_seed(0)
    sizes = [1 << 20, 1000, 4097]
    p0 = [rand(n, TDT[dt], 0.02, 300 + i) for i, n in enumerate(sizes)]
    ours = [torch.nn.Parameter(x.clone()) for x in p0]
    ref = [torch.nn.Parameter(x.clone()) for x in p0]
    o1 = ddp_b200.AdamW(ours, lr=1e-3, weight_decay=1e-2)
    o2 = torch.optim.AdamW(ref, lr=1e-3, weight_decay=1e-2, fused=True)
    for s in range(20):
        for i, (a, b) in enumerate(zip(ours, ref)):
            g = rand(sizes[i], TDT[dt], 0.01, 5000 + 10 * s + i)
            a.grad, b.grad = g.clone(), g.clone()
        torch.nn.utils.clip_grad_norm_(ref, 1.0)
        ddp_b200.clip_grad_norm_(ours, 1.0)
        o1.step()
        o2.step()
    for a, b in zip(ours, ref):
        d = (a.float() - b.float()).abs()
        if dt == "f32":
            assert float(d.max()) <= 1e-7
            assert float((d / b.float().abs().clamp_min(1e-3)).max()) <= 1e-4
        else:
            assert float((d / b.float().abs().clamp_min(1e-2)).max()) <= 2 ** -6
            assert float((d > 0).float().mean()) < 0.2
        sa, sb = o1.state[a], o2.state[b]
        assert torch.allclose(sa["exp_avg"].float(), sb["exp_avg"].float(), rtol=2 ** -6 if dt == "bf16" else 1e-5, atol=1e-9)
        assert torch.allclose(sa["exp_avg_sq"].float(), sb["exp_avg_sq"].float(), rtol=2 ** -6 if dt == "bf16" else 1e-5, atol=1e-12)


# ---- properties at full size -----------------------------------------------------------------
def test_full_size_properties_llama_1b_shapes(K, golden_dir):
    """BASELINE.json config 2 (Llama-3.2-1B shapes, bf16, 1.24e9 elements): no oracle pass at
    this size, so size-independent properties instead."""
    cases = json.load(open(os.path.join(golden_dir, "bucket_layouts.json")))
    numel = [c for c in cases if c["name"] == "llama-3.2-1b-bf16/init"][0]["numel"]
    assert sum(numel) == 1_235_814_400
    flat = torch.empty(sum(numel), dtype=torch.bfloat16, device=DEV)
    flat.normal_(0, 0.01, generator=torch.Generator(device=DEV).manual_seed(3))
    offs = np.cumsum([0] + numel[:-1])
    grads = [flat[o:o + n] for o, n in zip(offs, numel)]
    # (1) pack then unpack with scale 1 is the identity; pack with 1/2 twice == 1/4 once (exact in bf16)
    bucket = torch.empty_like(flat)
    bviews = [bucket[o:o + n] for o, n in zip(offs, numel)]
    tab = K.table([grads, bviews, None, None], numel, device=DEV)
    ss_pack = torch.empty(len(numel), dtype=torch.float32, device=DEV)
    K.pack(tab, torch.bfloat16, torch.bfloat16, 1.0, sumsq_out=ss_pack)
    assert torch.equal(bucket, flat)
    half = torch.empty_like(flat)
    hviews = [half[o:o + n] for o, n in zip(offs, numel)]
    K.pack(K.table([grads, hviews, None, None], numel, device=DEV), torch.bfloat16, torch.bfloat16, 0.5)
    K.pack(K.table([hviews, hviews, None, None], numel, device=DEV), torch.bfloat16, torch.bfloat16, 0.5)
    K.pack(K.table([grads, bviews, None, None], numel, device=DEV), torch.bfloat16, torch.bfloat16, 0.25)
    assert torch.equal(half, bucket)
    del half, hviews
    # (2) the norm from the fused pack pass == the stand-alone norm pass, bitwise; and the total
    #     agrees with torch's reduction over the flat buffer
    ss = torch.empty_like(ss_pack)
    K.sumsq(K.table([None, grads, None, None], numel, device=DEV), torch.bfloat16, ss)
    assert torch.equal(ss, ss_pack)
    total = float(K.norm_finalize(ss, 1.0, False)[0])
    want = float(torch.linalg.vector_norm(flat.float().view(-1, 1 << 10), dim=1).double().pow(2).sum().sqrt())
    assert abs(total - want) <= 1e-5 * want
    # (3) linearity of the norm: ||0.5 g|| == 0.5 ||g|| exactly for power-of-two scales
    K.sumsq(K.table([None, bviews, None, None], numel, device=DEV), torch.bfloat16, ss_pack)   # bucket = g/4
    assert torch.equal(ss_pack * 16.0, ss)
    # (4) AdamW on the flat tensor list == AdamW on one flat segment (segmentation-invariance), bitwise
    del bucket, bviews
    p1 = torch.empty_like(flat).normal_(0, 0.02, generator=torch.Generator(device=DEV).manual_seed(4))
    p2 = p1.clone()
    m1, v1, m2, v2 = (torch.zeros_like(flat) for _ in range(4))
    seg = lambda t: [t[o:o + n] for o, n in zip(offs, numel)]
    coef = torch.tensor([0.5], device=DEV)
    for step in (1, 2):
        K.adamw(K.table([seg(p1), grads, seg(m1), seg(v1)], numel, device=DEV), torch.bfloat16, torch.bfloat16,
                torch.bfloat16, step=step, clip_coef=coef, **HP)
        K.adamw(K.table([[p2], [flat], [m2], [v2]], [flat.numel()], device=DEV), torch.bfloat16, torch.bfloat16,
                torch.bfloat16, step=step, clip_coef=coef, **HP)
    assert torch.equal(p1, p2) and torch.equal(m1, m2) and torch.equal(v1, v2)
    # (5) spot-check a 1M-element window of that update against the oracle
    w = slice(123_456_789, 123_456_789 + (1 << 20))
    pn = to_np(torch.empty(1 << 20, dtype=torch.bfloat16).normal_(0, 1))  # placeholder, overwritten below
    p3 = torch.empty_like(flat).normal_(0, 0.02, generator=torch.Generator(device=DEV).manual_seed(4))
    pn, gn = to_np(p3[w]), to_np(flat[w])
    del p3
    mn, vn = np.zeros_like(pn), np.zeros_like(pn)
    for step in (1, 2):
        oracle.fused_clip_adamw(pn, gn, mn, vn, step=step, clip_coef=0.5, **HP)
    assert bits_equal(p1[w], pn) and bits_equal(m1[w], mn) and bits_equal(v1[w], vn)


# ---- the packed (f32x2) fast path of kernel (c): written-out sqrt / division vs the IEEE intrinsics -------
def _selfcheck(fn, *args):
    out = torch.zeros(3, dtype=torch.int64, device=DEV)
    ddp_b200._lib.check(fn(*args, out.data_ptr(), torch.cuda.current_stream().cuda_stream), fn.__name__)
    torch.cuda.synchronize()
    return [int(x) for x in out.tolist()]


def test_packed_sqrt_matches_ieee_over_the_whole_box(K):
    """Every float v in [0, 2^20) — zero and all denormals included — through the written-out square root
    (MUFU.RSQ + packed refinement on v*2^48), and through the whole denominator sqrt(v)/c + eps for the
    bias corrections of a few steps: bit-identical to __fsqrt_rn / div_by_const / __fadd_rn."""
    L = ddp_b200._lib.lib()
    box_end = 0x49800000                                    # 2^20
    bad = _selfcheck(L.ddpb200_selfcheck_sqrt, 0, box_end, -1.0, 0.0)
    assert bad[0] == 0, f"sqrt mismatches: {bad[0]}, first v bits {bad[1]:#x} -> {bad[2]:#x}"
    bc = np.zeros(2, dtype=np.float32)
    for beta2, step, eps in ((0.999, 1, 1e-8), (0.999, 7, 1e-8), (0.95, 3, 1e-6), (0.999, 100000, 1e-12), (0.99, 20, 1.0)):
        L.ddpb200_adamw_bias_corrections(0.9, beta2, step, bc.ctypes.data)
        for first, count in ((0, 1 << 26), (0x2f000000, 1 << 27), (0x45000000, box_end - 0x45000000)):
            bad = _selfcheck(L.ddpb200_selfcheck_sqrt, first, count, float(bc[1]), eps)
            assert bad[0] == 0, f"denominator mismatches for c={bc[1]} eps={eps}: {bad}"


def test_packed_division_matches_ieee_over_the_box(K):
    """2^32 pseudo-random (p, a, b) over the packed path's box (a: zero, denormal, up to 2^20; b in
    [2^-40, 2^21); |p| >= 2^-100): p - a/b as written out equals __fsub_rn(p, __fdiv_rn(a, b))."""
    L = ddp_b200._lib.lib()
    for seed in (1, 0x5eed5eed):
        bad = _selfcheck(L.ddpb200_selfcheck_div, seed, 1 << 31)
        assert bad[0] == 0, f"division mismatches: {bad[0]}, first (a,b) bits {bad[1]:#x}, p bits {bad[2]:#x}"


def test_div_by_const_matches_ieee_for_every_step(K):
    """div_by_const(x, c, RN(1/c)) == __fdiv_rn(x, c) for EVERY significand of x and every divisor
    c = sqrt(1 - beta2^t) the kernel can meet in t = 1..20000 for beta2 in {0.999, 0.99, 0.95}."""
    L = ddp_b200._lib.lib()
    cs = set()
    bc = np.zeros(2, dtype=np.float32)
    for beta2 in (0.999, 0.99, 0.95):
        for step in range(1, 20001):
            L.ddpb200_adamw_bias_corrections(0.9, beta2, step, bc.ctypes.data)
            cs.add(float(bc[1]))
            if bc[1] == 1.0:
                break
    cs = torch.tensor(sorted(cs), dtype=torch.float32, device=DEV)
    assert cs.numel() > 10000
    out = torch.zeros(3, dtype=torch.int64, device=DEV)
    ddp_b200._lib.check(L.ddpb200_selfcheck_div_by_const(cs.data_ptr(), cs.numel(), out.data_ptr(),
                                                         torch.cuda.current_stream().cuda_stream), "selfcheck")
    torch.cuda.synchronize()
    bad = [int(x) for x in out.tolist()]
    assert bad[0] == 0, f"{bad[0]} mismatches over {cs.numel()} constants; first (c,x) bits {bad[1]:#x}"


@pytest.mark.parametrize("dt", ["f32", "bf16"])
def test_fused_adamw_outside_the_packed_box_bit_exact(K, dt):
    """Inputs that leave the packed path's box take the IEEE tail instead; both must equal the oracle:
    zero and tiny parameters, zero / denormal / huge gradients and second moments, infinities, and
    per-launch constants outside the box (eps = 0, eps = 1e-20, beta2 so close to 1 that sqrt(bc2) < 2^-10)."""
    n = 8192 + 40
    rng = np.random.default_rng(11)
    tdt = TDT[dt]

    def mk(vals):
        return torch.tensor(vals, dtype=torch.float32).to(tdt).to(DEV)

    special_p = [0.0, -0.0, 1e-38, -3e-35, 1e-30, 0.5, -2.0, 1e20, 3e38]
    special_g = [0.0, 1e-40, -1e-38, 1e-25, 1e-12, 1e-3, -1.0, 1e4, 1e12, -3e19, 1e30, float("inf")]
    p0 = rng.choice(special_p, n)
    g0 = rng.choice(special_g, n)
    v0 = np.abs(rng.choice(special_g[:-1], n)) * np.abs(rng.choice(special_g[:-1], n))
    m0 = rng.choice(special_g[:-1], n) * 0.1
    # some whole groups stay inside the box so both tails run in one launch
    p0[:4096], g0[:4096] = rng.normal(0, 0.02, 4096), rng.normal(0, 0.01, 4096)
    v0[:4096], m0[:4096] = rng.normal(0, 0.01, 4096) ** 2, rng.normal(0, 0.005, 4096)
    for hp in (HP, dict(HP, eps=0.0), dict(HP, eps=1e-20), dict(HP, beta2=0.99999999), dict(HP, eps=1e-8, lr=10.0)):
        with np.errstate(all="ignore"):
            p, g, m, v = mk(p0), mk(g0), mk(m0), mk(np.float32(v0))
        pn, gn, mn, vn = to_np(p), to_np(g), to_np(m), to_np(v)
        for step, clip in ((1, None), (2, 0.5)):
            coef = None if clip is None else torch.tensor([clip], dtype=torch.float32, device=DEV)
            K.adamw(K.table([[p], [g], [m], [v]], [n]), tdt, tdt, tdt, step=step, clip_coef=coef,
                    prefer_device=False, **hp)
            with np.errstate(all="ignore"):
                oracle.fused_clip_adamw(pn, gn, mn, vn, step=step, clip_coef=clip, **hp)
            torch.cuda.synchronize()
            for name, got, want in (("p", p, pn), ("m", m, mn), ("v", v, vn)):
                a = to_np(got)
                a = a.view(np.uint16 if a.dtype == np.uint16 else np.uint32)
                b = want.view(np.uint16 if want.dtype == np.uint16 else np.uint32)
                gf, wf = got.float().cpu().numpy(), (from_np(want).float().cpu().numpy())
                same = (a == b) | (np.isnan(gf) & np.isnan(wf))          # NaN payloads are not part of the contract
                assert same.all(), (dt, hp, step, name, int((~same).sum()), np.flatnonzero(~same)[:5])
