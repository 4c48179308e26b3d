"""GPU tests of the whole step through the public classes: ours (DistributedDataParallel +
clip_grad_norm_ + AdamW on the sm_100a kernels) beside stock torch (DDP + clip_grad_norm_ +
AdamW(fused=True)) on the same model, weights and batches. Single GPU with a 1-rank NCCL group,
and a 2-rank NCCL run when the box has two GPUs."""
import copy
import importlib.util
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _golden_module():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _train_compare(rank, ws, dtype, gabv, steps=5, fold_clip=True, fused=False):
    import ddp_b200
    mg = _golden_module()
    dev = torch.device("cuda", rank if ws > 1 else 0)
    torch.manual_seed(0)
    base = mg.SmallNet().to(dev).to(dtype)
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a, device_ids=[dev.index], bucket_cap_mb=0.008,
                                                    gradient_as_bucket_view=gabv)
    ours = ddp_b200.DistributedDataParallel(b, device_ids=[dev.index], bucket_cap_mb=0.008,
                                            gradient_as_bucket_view=gabv, fused_allreduce=fused)
    o_ref = torch.optim.AdamW(a.parameters(), lr=1e-2, fused=True)
    o_our = ddp_b200.AdamW(b.parameters(), lr=1e-2)
    for s in range(steps):
        ids, tgt = mg.live_batch(s, rank)          # CPU tensors: DDP moves them (device_ids given)
        for m, o, mine in ((ref, o_ref, False), (ours, o_our, True)):
            loss = torch.nn.functional.cross_entropy(m(ids).flatten(0, 1).float(), tgt.to(dev).flatten())
            loss.backward()
            if s == 0:
                first_loss = float(loss.detach())
            if mine:
                tn = ours.clip_grad_norm_(0.5, fused_optimizer=o_our if fold_clip else None)
            else:
                tn_ref = torch.nn.utils.clip_grad_norm_(a.parameters(), 0.5)
            o.step()
            o.zero_grad()
        rel = 2 ** -6 if dtype == torch.bfloat16 else 1e-4
        assert abs(float(tn) - float(tn_ref)) <= rel * float(tn_ref) * (1 + s), (s, float(tn), float(tn_ref))
        if s >= 1:
            # layouts after the rebuild: same buckets as the stock reducer built
            want = str(ref._get_ddp_logging_data().get("rebuilt_bucket_sizes", ""))
            got = ours._get_ddp_logging_data()["rebuilt_bucket_sizes"]
            if want:        # the stock logger only samples runtime data on some iterations
                assert ", ".join(str(x) for x in got) == want
            # stock torch's planner on the grad-ready order recorded in the golden file
            ready = [7, 6, 5, 3, 4, 2, 1, 0]
            plist = list(b.parameters())
            cap = int(0.008 * 1024 * 1024)
            tb, _ = dist._compute_bucket_assignment_by_size([plist[i] for i in ready], [cap, cap], [False] * 8, ready)
            assert got == [sum(plist[i].numel() for i in bk) * plist[0].element_size() for bk in tb]
            assert [bk["parameters"] for bk in ours.bucket_layout()] == [list(x) for x in tb]
    for pa, pb in zip(a.parameters(), b.parameters()):
        d = (pa.float() - pb.float()).abs()
        if dtype == torch.float32:
            assert float(d.max()) <= 2e-5, float(d.max())         # 5 steps at lr 1e-2, |p| up to ~3
        else:
            assert float(d.max()) <= 0.1 and float(d.mean()) <= 4e-3, (float(d.max()), float(d.mean()))
    return first_loss


@pytest.fixture(scope="module")
def pg1():
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = "29661"
    torch.cuda.set_device(0)
    dist.init_process_group("nccl", rank=0, world_size=1)
    yield
    dist.destroy_process_group()


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("gabv", [False, True])
def test_single_gpu_step_matches_stock_torch(pg1, dtype, gabv):
    _train_compare(0, 1, dtype, gabv)


def test_single_gpu_unfolded_clip_scales_grads_in_place(pg1):
    _train_compare(0, 1, torch.float32, True, steps=3, fold_clip=False)
    _train_compare(0, 1, torch.float32, False, steps=3, fold_clip=False)


def test_reduce_gradients_entry_and_stats(pg1):
    """The manual gradient feed used by bench.py drives the same scheduler as a real backward."""
    import ddp_b200
    dev = torch.device("cuda", 0)
    torch.manual_seed(1)
    net = torch.nn.Sequential(torch.nn.Linear(64, 128), torch.nn.Linear(128, 64)).to(dev).to(torch.bfloat16)
    ddp = ddp_b200.DistributedDataParallel(net, bucket_cap_mb=0.01, gradient_as_bucket_view=True)
    params = list(net.parameters())
    grads = [torch.randn_like(p) for p in params]
    ddp.reduce_gradients([g.clone() for g in grads])
    ddp.reduce_gradients([g.clone() for g in grads])           # second call: rebuilt buckets
    assert ddp._get_ddp_logging_data()["has_rebuilt_buckets"]
    for p, g in zip(params, grads):
        assert torch.equal(p.grad, g)                           # ws=1: averaging is the identity
    want = torch.nn.utils.get_total_norm(grads)
    got = ddp.clip_grad_norm_(1e9)
    assert abs(float(got) - float(want)) <= 2 ** -7 * float(want)


def _two_rank_worker(rank, ws, port):
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", rank=rank, world_size=ws)
    for dtype in (torch.float32, torch.bfloat16):
        for gabv in (False, True):
            _train_compare(rank, ws, dtype, gabv)
    # the same steps with the bucket allreduce + norm done by the fused peer-memory kernel,
    # through NVLink peer loads/stores and through NVSwitch multicast
    for mode in ("peer", "mc"):
        os.environ["DDPB200_FUSED_MODE"] = mode
        for dtype in (torch.float32, torch.bfloat16):
            for gabv in (False, True):
                _train_compare(rank, ws, dtype, gabv, fused=True)
        _fused_vs_nccl(rank, ws, mode)
        _sharded_vs_replicated(rank, ws, mode)
    os.environ["DDPB200_FUSED_MODE"] = "auto"
    for fused in (False, True):
        _mixed_dtype_two_ranks(rank, fused)
        _bf16_buckets_two_ranks(rank, fused)
    dist.destroy_process_group()


def _bf16_buckets_two_ranks(rank, fused):
    """fp32 model, bf16 gradient buckets (pack casts down with one rounding, unpack widens back)."""
    import ddp_b200
    mg = _golden_module()
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    base = mg.SmallNet().to(dev)
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a, device_ids=[rank], bucket_cap_mb=0.008)
    ours = ddp_b200.DistributedDataParallel(b, device_ids=[rank], bucket_cap_mb=0.008, fused_allreduce=fused,
                                            grad_bucket_dtype=torch.bfloat16)
    for s in range(3):
        ids, tgt = mg.live_batch(70 + s, rank)
        for m in (ref, ours):
            m.zero_grad()
            torch.nn.functional.cross_entropy(m(ids).flatten(0, 1), tgt.to(dev).flatten()).backward()
        assert all(bk.dtype == torch.bfloat16 for bk in ours._buckets)
        for pa, pb in zip(a.parameters(), b.parameters()):
            scale = float(pa.grad.abs().max())
            assert pb.grad.dtype == torch.float32
            assert torch.allclose(pa.grad, pb.grad, rtol=0, atol=2 ** -7 * scale + 1e-9), (fused, s)
        n_ref = torch.nn.utils.clip_grad_norm_(a.parameters(), 1e9)
        n_our = ours.clip_grad_norm_(1e9)
        assert abs(float(n_ref) - float(n_our)) <= 2 ** -7 * float(n_ref)


def _mixed_dtype_two_ranks(rank, fused):
    """bf16 + fp32 parameters (two bucket dtypes, two arenas) across two ranks, NCCL and fused."""
    import ddp_b200
    dev = torch.device("cuda", rank)
    torch.manual_seed(3)
    base = _MixedNet().to(dev)
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a, device_ids=[rank], bucket_cap_mb=0.004)
    ours = ddp_b200.DistributedDataParallel(b, device_ids=[rank], bucket_cap_mb=0.004, gradient_as_bucket_view=True,
                                            fused_allreduce=fused)
    o_ref = torch.optim.AdamW(a.parameters(), lr=1e-2, fused=True)
    o_our = ddp_b200.AdamW(b.parameters(), lr=1e-2)
    for s in range(4):
        x = torch.randn(8, 32, device=dev, generator=torch.Generator(device=dev).manual_seed(10 * s + rank))
        ref(x).backward()
        ours(x).backward()
        for pa, pb in zip(a.parameters(), b.parameters()):
            if pa.numel() and s == 0:
                tol = dict(rtol=2 ** -6, atol=1e-3) if pa.dtype == torch.bfloat16 else dict(rtol=1e-5, atol=1e-6)
                assert torch.allclose(pa.grad.float(), pb.grad.float(), **tol), (fused, pa.dtype)
        n_ref = torch.nn.utils.clip_grad_norm_(a.parameters(), 0.1)
        n_our = ours.clip_grad_norm_(0.1, fused_optimizer=o_our)
        assert abs(float(n_ref) - float(n_our)) <= 3e-3 * float(n_ref) * (1 + s), (fused, s, float(n_ref), float(n_our))
        o_ref.step(); o_our.step()
        o_ref.zero_grad(); o_our.zero_grad()
    flat = torch.cat([p.detach().float().reshape(-1) for p in b.parameters()])
    both = [torch.empty_like(flat) for _ in range(2)]
    dist.all_gather(both, flat)
    assert torch.equal(both[0], both[1]), "ranks diverged"


def _sharded_vs_replicated(rank, ws, mode):
    """ZeRO-1-style sharded update (reduce-scatter -> norm -> AdamW on owned chunks -> all-gather of
    parameters) against the replicated path: parameters and norms BIT-IDENTICAL after every step, on
    every rank, although bucket order and chunk ownership differ between the two."""
    import ddp_b200
    mg = _golden_module()
    dev = torch.device("cuda", rank)
    for dtype in (torch.bfloat16, torch.float32):
        torch.manual_seed(0)
        base = mg.SmallNet().to(dev).to(dtype)
        a, b = copy.deepcopy(base), copy.deepcopy(base)
        rep = ddp_b200.DistributedDataParallel(a, device_ids=[rank], bucket_cap_mb=0.008, gradient_as_bucket_view=True,
                                               fused_allreduce=True)
        shd = ddp_b200.DistributedDataParallel(b, device_ids=[rank], bucket_cap_mb=0.008, gradient_as_bucket_view=True,
                                               fused_allreduce=True, shard_optimizer=True)
        o_rep = ddp_b200.AdamW(a.parameters(), lr=1e-2)
        o_shd = ddp_b200.ShardedAdamW(shd, lr=1e-2)
        for s in range(5):
            ids, tgt = mg.live_batch(40 + s, rank)
            norms = []
            for m, o, d in ((rep, o_rep, rep), (shd, o_shd, shd)):
                loss = torch.nn.functional.cross_entropy(m(ids).flatten(0, 1).float(), tgt.to(dev).flatten())
                loss.backward()
                norms.append(d.clip_grad_norm_(0.5, fused_optimizer=o))
                o.step()
                o.zero_grad()
            assert torch.equal(norms[0], norms[1]), (mode, dtype, s)
            for pa, pb in zip(a.parameters(), b.parameters()):
                assert torch.equal(pa, pb), (mode, dtype, s)
        flat = torch.cat([p.detach().reshape(-1) for p in b.parameters()])
        both = [torch.empty_like(flat) for _ in range(ws)]
        dist.all_gather(both, flat)
        assert all(torch.equal(both[0], x) for x in both[1:]), "ranks hold different parameters"
        sd = o_shd.state_dict()
        assert sd["sharded"] and sd["step"] == 5 and sd["exp_avg_shard"].abs().sum() > 0
        # the shards, gathered into torch's per-parameter format, equal the replicated optimizer's
        # state bit for bit, and load into a stock torch.optim.AdamW
        full = o_shd.consolidated_state_dict()
        rep_state = o_rep.state_dict()["state"]
        assert set(full["state"]) == set(rep_state)
        for i, st in full["state"].items():
            assert float(st["step"]) == float(rep_state[i]["step"]) == 5.0
            assert torch.equal(st["exp_avg"], rep_state[i]["exp_avg"]), (mode, dtype, i)
            assert torch.equal(st["exp_avg_sq"], rep_state[i]["exp_avg_sq"]), (mode, dtype, i)
        torch.optim.AdamW(list(b.parameters()), lr=1e-2).load_state_dict(full)
        # it is a torch Optimizer: LR schedulers drive it, and it loads torch-format state (its own
        # consolidated form, or the replicated optimizer's state_dict) back into the shards
        assert isinstance(o_shd, torch.optim.Optimizer)
        sched = torch.optim.lr_scheduler.StepLR(o_shd, step_size=1, gamma=0.5)
        sched.step()
        assert o_shd.param_groups[0]["lr"] == 5e-3
        for src in (full, o_rep.state_dict()):
            fresh = ddp_b200.ShardedAdamW(shd, lr=1e-2)
            fresh.load_state_dict(src)
            assert fresh.step_count == 5
            assert torch.equal(fresh.exp_avg, o_shd.exp_avg) and torch.equal(fresh.exp_avg_sq, o_shd.exp_avg_sq), (mode, dtype)
        fresh.load_state_dict(o_shd.state_dict())
        assert torch.equal(fresh.exp_avg, o_shd.exp_avg) and fresh.param_groups[0]["lr"] == 5e-3


def _fused_vs_nccl(rank, ws, mode):
    """Fused allreduce+norm on ragged segments against NCCL + a separate norm: sums within one
    rounding of the bucket dtype, identical bits on every rank, sums of squares of exactly what was
    stored."""
    import ddp_b200
    from ddp_b200.symm import SymmetricArena
    dev = torch.device("cuda", rank)
    K = ddp_b200.CudaKernels()
    sizes = [5, 4096, 70001, 2047, 1 << 20, 0, 33]
    starts, o = [], 0
    for n in sizes:                    # segments back to back (no padding): some land misaligned
        starts.append(o)
        o += n
    for dt in (torch.bfloat16, torch.float32):
        arena = SymmetricArena(o, dt, dev, dist.group.WORLD, max_chunks=o // 2048 + len(sizes) + 1)
        g = torch.Generator(device=dev).manual_seed(7 + rank)
        arena.tensor.normal_(0, 0.01, generator=g)
        ref = arena.tensor.clone()
        dist.all_reduce(ref)
        views = [arena.tensor[s:s + n] for s, n in zip(starts, sizes)]
        tab = K.table([None, views, None, None], sizes, chunk_elems=2048, device=dev)
        ss = torch.empty(len(sizes), dtype=torch.float32, device=dev)
        K.allreduce_sumsq(tab, dt, arena, ss)
        torch.cuda.synchronize()
        got = arena.tensor.clone()
        # NCCL rounds to the bucket dtype at every hop; tolerance in units of the TERMS' size (~0.03)
        tol = dict(rtol=2 ** -7, atol=5e-4) if dt == torch.bfloat16 else dict(rtol=1e-6, atol=1e-8)
        assert torch.allclose(got.float(), ref.float(), **tol), (mode, dt)
        want_ss = torch.stack([v.double().pow(2).sum() for v in views]).float()
        assert torch.allclose(ss, want_ss, rtol=3e-6, atol=0), (mode, dt, ss, want_ss)
        both = [torch.empty_like(got) for _ in range(ws)]
        dist.all_gather(both, got)
        assert torch.equal(both[0], both[1]), "ranks hold different averaged gradients"
        ss_all = [torch.empty_like(ss) for _ in range(ws)]
        dist.all_gather(ss_all, ss)
        assert torch.equal(ss_all[0], ss_all[1]), "ranks hold different norms"
        del arena


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs (gpurun --gpus 2)")
def test_two_gpu_nccl_step_matches_stock_torch():
    mp.spawn(_two_rank_worker, args=(2, 29671), nprocs=2, join=True)


def test_launch_script_single_gpu():
    """scripts/run_node0.sh (the reference's launch line collapsed onto one box) trains a few steps
    of the tiny model on the B200 path."""
    import subprocess
    env = dict(os.environ, NPROC="1", DDPB200_MODEL="tiny", DDPB200_MAX_STEPS="5", DDPB200_SEQ_LEN="64",
               MASTER_PORT="29741")
    env.pop("MASTER_ADDR", None)
    res = subprocess.run(["bash", os.path.join(ROOT, "scripts", "run_node0.sh")], cwd=ROOT, env=env,
                         capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-3000:]
    assert "done: 5 steps" in res.stderr and "Using device: cuda:0" in res.stderr


class _MixedNet(torch.nn.Module):
    """bf16 matrices with fp32 norm parameters and an EMPTY parameter: two bucket dtypes, two optimizer
    launches, a zero-length segment."""

    def __init__(self):
        super().__init__()
        self.fc1 = torch.nn.Linear(32, 64).to(torch.bfloat16)
        self.ln = torch.nn.LayerNorm(64)                      # fp32
        self.fc2 = torch.nn.Linear(64, 16).to(torch.bfloat16)
        self.empty = torch.nn.Parameter(torch.zeros(0))

    def forward(self, x):
        h = torch.relu(self.fc1(x.to(torch.bfloat16)))
        h = self.ln(h.float()).to(torch.bfloat16)
        return self.fc2(h).float().sum() + self.empty.sum()


@pytest.mark.parametrize("gabv", [False, True])
def test_mixed_dtype_and_empty_parameters(pg1, gabv):
    import ddp_b200
    dev = torch.device("cuda", 0)
    torch.manual_seed(3)
    base = _MixedNet().to(dev)
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a, device_ids=[0], bucket_cap_mb=0.004, gradient_as_bucket_view=gabv)
    ours = ddp_b200.DistributedDataParallel(b, device_ids=[0], bucket_cap_mb=0.004, gradient_as_bucket_view=gabv)
    o_ref = torch.optim.AdamW(a.parameters(), lr=1e-2, fused=True)
    o_our = ddp_b200.AdamW(b.parameters(), lr=1e-2)
    for s in range(4):
        x = torch.randn(8, 32, device=dev, generator=torch.Generator(device=dev).manual_seed(s))
        ref(x).backward()
        ours(x).backward()
        n_ref = torch.nn.utils.clip_grad_norm_(a.parameters(), 0.1)
        n_our = ours.clip_grad_norm_(0.1, fused_optimizer=o_our)
        assert abs(float(n_ref) - float(n_our)) <= 2e-3 * float(n_ref) * (1 + s)
        o_ref.step(); o_our.step()
        o_ref.zero_grad(); o_our.zero_grad()
    assert o_our.launches_last_step == 2                       # one per dtype group
    for pa, pb in zip(a.parameters(), b.parameters()):
        if pa.numel():
            # plumbing check: once the bf16 weights differ by an ulp the two runs see slightly different
            # gradients, and Adam turns a sign flip of a near-zero gradient into a full lr-sized step
            d = (pa.detach().float() - pb.detach().float()).abs()
            assert float(d.max()) <= 2.5e-2 and float(d.mean()) <= 2e-3, (pa.dtype, float(d.max()), float(d.mean()))
    # free function on a mixed list: norm in fp32 (not all grads are bf16), grads scaled in place
    x = torch.randn(8, 32, device=dev)
    ref(x).backward(); ours(x).backward()
    n_ref = torch.nn.utils.clip_grad_norm_(a.parameters(), 0.05)
    n_our = ddp_b200.clip_grad_norm_(b.parameters(), 0.05)
    assert n_our.dtype == n_ref.dtype
    assert abs(float(n_ref) - float(n_our)) <= 1e-2 * float(n_ref)


def test_grad_scaler_unscale_path(pg1):
    """torch.amp.GradScaler hands grad_scale / found_inf to optimizers that declare
    _step_supports_amp_scaling; the fused kernel then unscales inside the update (the "unscale" of the
    north star). Same flow on torch.optim.AdamW(fused=True): parameters agree, a step with an inf
    gradient is skipped by both, and the scaler backs off."""
    import ddp_b200
    dev = torch.device("cuda", 0)
    torch.manual_seed(5)
    base = torch.nn.Sequential(torch.nn.Linear(40, 64), torch.nn.Tanh(), torch.nn.Linear(64, 8)).to(dev)
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    o_ref = torch.optim.AdamW(a.parameters(), lr=1e-2, fused=True)
    o_our = ddp_b200.AdamW(b.parameters(), lr=1e-2)
    s_ref = torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=1000)
    s_our = torch.amp.GradScaler("cuda", init_scale=1024.0, growth_interval=1000)
    for s in range(5):
        x = torch.randn(16, 40, device=dev, generator=torch.Generator(device=dev).manual_seed(s))
        for m, o, sc in ((a, o_ref, s_ref), (b, o_our, s_our)):
            loss = m(x).pow(2).mean()
            sc.scale(loss).backward()
            if s == 3:                                   # poison one gradient: the step must be skipped
                next(m.parameters()).grad[0, 0] = float("inf")
            sc.step(o)
            sc.update()
            o.zero_grad()
        assert float(s_ref.get_scale()) == float(s_our.get_scale())
        for pa, pb in zip(a.parameters(), b.parameters()):
            assert torch.allclose(pa, pb, rtol=1e-5, atol=1e-6), (s, float((pa - pb).abs().max()))
    assert float(s_our.get_scale()) == 512.0             # backed off once
    # 5 calls, one skipped. Under a scaler the count lives on the device (no host sync inside a scaled step);
    # state_dict() brings the per-parameter CPU scalars up to date
    assert float(o_our.state_dict()["state"][0]["step"]) == 4.0
    assert float(o_our.state[next(b.parameters())]["step"]) == 4.0
