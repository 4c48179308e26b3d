"""world_size-2 gloo tests of the HOST logic on CPU: bucket scheduling, rebuild in grad-ready order,
comm hooks, no_sync, gradient_as_bucket_view, the optimizer's table handling. The kernels are
answered by the test-only oracle backend (tests/oracle_backend.py); what is checked here is the
plumbing around them, against (a) the golden record of a live stock-torch DDP run on the same
model and batches and (b) stock torch DDP running side by side in the same processes."""
import importlib.util
import json
import os
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


def _load_golden_module():
    spec = importlib.util.spec_from_file_location("make_golden", os.path.join(HERE, "golden", "make_golden.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def _setup(rank, ws, port):
    for p in (ROOT, HERE):
        if p not in sys.path:
            sys.path.insert(0, p)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    torch.set_num_threads(1)


def _worker_vs_golden(rank, ws, port, gabv):
    _setup(rank, ws, port)
    import ddp_b200
    from oracle_backend import OracleKernels
    mg = _load_golden_module()
    z = np.load(os.path.join(HERE, "golden", "ddp_live.npz"))
    cfg = json.loads(str(z["cfg"]))
    golden_layout = json.loads(str(z["layout"]))
    torch.manual_seed(0)
    net = mg.SmallNet()
    k = OracleKernels()
    ddp = ddp_b200.DistributedDataParallel(net, bucket_cap_mb=cfg["bucket_cap_mb"], gradient_as_bucket_view=gabv,
                                           kernels=k)
    params = list(net.parameters())
    opt = ddp_b200.AdamW(params, lr=cfg["lr"], kernels=k)
    for s in range(cfg["steps"]):
        ids, tgt = mg.live_batch(s, rank)
        loss = torch.nn.functional.cross_entropy(ddp(ids).flatten(0, 1), tgt.flatten())
        # bucket layout in force for this backward == what stock DDP used on that iteration
        want = sorted([e for e in golden_layout if e[0] == s], key=lambda e: e[1])
        got = ddp.bucket_layout()
        assert [b["parameters"] for b in got] == [e[2] for e in want], (s, got)
        assert [b["offsets"] for b in got] == [e[3] for e in want], s
        assert [b["numel"] for b in got] == [e[4] for e in want], s
        loss.backward()
        assert abs(float(loss.detach()) - float(z[f"loss/{s}/{rank}"])) < 1e-5
        for i, p in enumerate(params):
            # averaged gradients: bit-exact on the first step (ws=2: one addition, order-free);
            # later steps start from parameters a few ulp apart (fused vs default AdamW arithmetic)
            if s == 0:
                assert np.array_equal(p.grad.numpy(), z[f"avg_g/{s}/{i}"]), (s, i)
            else:
                assert np.allclose(p.grad.numpy(), z[f"avg_g/{s}/{i}"], rtol=1e-3, atol=1e-7), (s, i)
            if gabv:
                bk = ddp._bucket_of[i]
                assert p.grad.data_ptr() == bk.views[bk.spec.indices.index(i)].data_ptr()
        tn = ddp.clip_grad_norm_(cfg["max_norm"], fused_optimizer=opt)
        assert abs(float(tn) - float(z[f"norm/{s}"])) <= (1e-6 if s == 0 else 1e-4) * float(tn)
        opt.step()
        opt.zero_grad()
        for i, p in enumerate(params):
            assert np.allclose(p.detach().numpy(), z[f"p/{s}/{i}"], rtol=2e-6, atol=1e-7), (s, i)
    st = ddp._get_ddp_logging_data()
    assert st["has_rebuilt_buckets"] and st["iterations"] == cfg["steps"]
    assert st["rebuilt_bucket_sizes"] == [e[4] * 4 for e in sorted([e for e in golden_layout if e[0] == 1],
                                                                   key=lambda e: e[1])]
    # one pack per bucket per step: 1 + 3*3
    assert st["pack_launches"] == 1 + 3 * (cfg["steps"] - 1)
    dist.destroy_process_group()


@pytest.mark.parametrize("gabv", [False, True])
def test_ddp_scheduler_reproduces_live_torch_run(native_lib, gabv):
    mp.spawn(_worker_vs_golden, args=(2, 29641 + int(gabv), gabv), nprocs=2, join=True)


def _worker_side_by_side(rank, ws, port):
    """Stock torch DDP and ours in the same process on the same data: grads bit-exact, including
    gradient accumulation under no_sync() and a stock communication hook."""
    _setup(rank, ws, port)
    import copy
    import ddp_b200
    from oracle_backend import OracleKernels
    from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
    mg = _load_golden_module()
    torch.manual_seed(0)
    base = mg.SmallNet()
    a, b, c = copy.deepcopy(base), copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a, bucket_cap_mb=0.008)
    ours = ddp_b200.DistributedDataParallel(b, bucket_cap_mb=0.008, kernels=OracleKernels())
    hooked = ddp_b200.DistributedDataParallel(c, bucket_cap_mb=0.008, kernels=OracleKernels())
    hooked.register_comm_hook(None, default_hooks.allreduce_hook)
    with pytest.raises(RuntimeError):
        hooked.register_comm_hook(None, default_hooks.allreduce_hook)

    def lossf(m, s, r):
        ids, tgt = mg.live_batch(s, r)
        return torch.nn.functional.cross_entropy(m(ids).flatten(0, 1), tgt.flatten())

    for s in range(3):
        for m in (ref, ours, hooked):
            m.zero_grad()
            with m.no_sync():                      # micro-batch 1: local accumulation only
                lossf(m, 10 + s, rank).backward()
            lossf(m, 20 + s, rank).backward()      # micro-batch 2: synchronised
        for pr, po, ph in zip(a.parameters(), b.parameters(), c.parameters()):
            assert torch.equal(pr.grad, po.grad), s
            assert torch.equal(pr.grad, ph.grad), s          # ws=2: g/2 == g*0.5 exactly
    # state_dict keeps the `module.` prefix of the wrapper
    assert list(ours.state_dict().keys()) == list(ref.state_dict().keys())
    # parameters were broadcast from rank 0 at construction
    torch.manual_seed(100 + rank)
    d = mg.SmallNet()
    ddp_b200.DistributedDataParallel(d, kernels=OracleKernels())
    flat = torch.cat([p.detach().reshape(-1) for p in d.parameters()])
    gathered = [torch.empty_like(flat) for _ in range(ws)]
    dist.all_gather(gathered, flat)
    assert torch.equal(gathered[0], gathered[1])
    dist.destroy_process_group()


def test_ddp_side_by_side_with_stock_torch(native_lib):
    mp.spawn(_worker_side_by_side, args=(2, 29651), nprocs=2, join=True)


def test_adamw_state_dict_and_table_reuse(native_lib):
    """Single process, oracle backend: state_dict layout equals torch.optim.AdamW's, and the
    optimizer reuses its segment table when pointers repeat."""
    sys.path.insert(0, HERE)
    import ddp_b200
    from oracle_backend import OracleKernels
    torch.manual_seed(0)
    ps = [torch.nn.Parameter(torch.randn(n)) for n in (5, 300, 2049)]
    pt = [torch.nn.Parameter(p.detach().clone()) for p in ps]
    ours = ddp_b200.AdamW(ps, lr=1e-2, weight_decay=0.1, kernels=OracleKernels())
    ref = torch.optim.AdamW(pt, lr=1e-2, weight_decay=0.1)
    for s in range(3):
        gs = [torch.randn_like(p) for p in ps]
        for p, q, g in zip(ps, pt, gs):
            p.grad = g.clone()
            q.grad = g.clone()
        ours.step()
        ref.step()
    for p, q in zip(ps, pt):
        assert torch.allclose(p, q, rtol=2e-6, atol=1e-7)
    sd, rd = ours.state_dict(), ref.state_dict()
    assert sd["state"].keys() == rd["state"].keys()
    for kx in sd["state"]:
        assert set(sd["state"][kx]) == {"step", "exp_avg", "exp_avg_sq"} == set(rd["state"][kx])
        assert float(sd["state"][kx]["step"]) == float(rd["state"][kx]["step"]) == 3.0
    assert set(rd["param_groups"][0]) <= set(sd["param_groups"][0])
    assert sd["param_groups"][0]["decoupled_weight_decay"] is True
    # load into a fresh optimizer and continue: same trajectory
    ours2 = ddp_b200.AdamW(ps, lr=1e-2, weight_decay=0.1, kernels=OracleKernels())
    ours2.load_state_dict(sd)
    for p, q in zip(ps, pt):
        g = torch.ones_like(p)
        p.grad, q.grad = g.clone(), g.clone()
    ours2.step()
    ref.step()
    for p, q in zip(ps, pt):
        assert torch.allclose(p, q, rtol=3e-6, atol=1e-7)
    with pytest.raises(NotImplementedError):
        ddp_b200.AdamW(ps, amsgrad=True)


def _worker_coalesced_feed(rank, ws, port):
    """reduce_gradients with every gradient present: one pack / one allreduce / one norm launch for
    all buckets, same averaged gradients and norm as the bucket-by-bucket path."""
    _setup(rank, ws, port)
    import copy
    import ddp_b200
    from oracle_backend import OracleKernels
    mg = _load_golden_module()
    torch.manual_seed(0)
    base = mg.SmallNet()
    nets = [copy.deepcopy(base) for _ in range(3)]
    ka, kb, kc = OracleKernels(), OracleKernels(), OracleKernels()
    one = ddp_b200.DistributedDataParallel(nets[0], bucket_cap_mb=0.008, gradient_as_bucket_view=True, kernels=ka)
    per = ddp_b200.DistributedDataParallel(nets[1], bucket_cap_mb=0.008, gradient_as_bucket_view=True, kernels=kb)
    cop = ddp_b200.DistributedDataParallel(nets[2], bucket_cap_mb=0.008, gradient_as_bucket_view=False, kernels=kc)
    one._coalesce_groups = cop._coalesce_groups = 1      # one launch for everything (default splits in 4 to pipeline)
    order = [7, 6, 5, 3, 4, 2, 1, 0]
    for s in range(3):
        g = torch.Generator().manual_seed(50 + 10 * s + rank)
        grads = [torch.randn(p.shape, generator=g) for p in base.parameters()]
        for d, kw in ((one, {}), (per, dict(before_bucket=lambda b, idx: None)), (cop, {})):
            d.reduce_gradients([x.clone() for x in grads], order=order if s == 0 else None, **kw)
        for pa, pb, pc in zip(nets[0].parameters(), nets[1].parameters(), nets[2].parameters()):
            assert torch.equal(pa.grad, pb.grad) and torch.equal(pa.grad, pc.grad)
        na, nb, nc = one.clip_grad_norm_(1e9), per.clip_grad_norm_(1e9), cop.clip_grad_norm_(1e9)
        assert float(na) == float(nb) == float(nc)
    assert len(one.bucket_layout()) == 3
    # steps 1 and 2 ran on the 3 rebuilt buckets: coalesced = 1 pack each, bucket-by-bucket = 3 each
    assert ka.calls["pack"] == 1 + 2 * 1 and kb.calls["pack"] == 1 + 2 * 3
    assert one._get_ddp_logging_data()["allreduces"] == 3 and per._get_ddp_logging_data()["allreduces"] == 1 + 2 * 3
    dist.destroy_process_group()


def test_coalesced_gradient_feed(native_lib):
    mp.spawn(_worker_coalesced_feed, args=(2, 29661), nprocs=2, join=True)


def _worker_compress_hook(rank, ws, port):
    """A stock hook that returns a DIFFERENT tensor than the bucket (bf16_compress_hook: cast, divide,
    allreduce, decompress) goes through the same protocol as in torch: result copied back into the
    bucket, gradients equal to stock DDP running the same hook."""
    _setup(rank, ws, port)
    import copy
    import ddp_b200
    from oracle_backend import OracleKernels
    from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
    mg = _load_golden_module()
    torch.manual_seed(0)
    base = mg.SmallNet()
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a, bucket_cap_mb=0.008)
    ours = ddp_b200.DistributedDataParallel(b, bucket_cap_mb=0.008, kernels=OracleKernels())
    ref.register_comm_hook(None, default_hooks.bf16_compress_hook)
    ours.register_comm_hook(None, default_hooks.bf16_compress_hook)
    for s in range(3):
        for m in (ref, ours):
            m.zero_grad()
            ids, tgt = mg.live_batch(30 + s, rank)
            torch.nn.functional.cross_entropy(m(ids).flatten(0, 1), tgt.flatten()).backward()
        for pr, po in zip(a.parameters(), b.parameters()):
            assert torch.equal(pr.grad, po.grad), s
        n_ref = torch.nn.utils.clip_grad_norm_(a.parameters(), 1e9)
        n_our = ours.clip_grad_norm_(1e9)
        assert abs(float(n_ref) - float(n_our)) <= 1e-6 * float(n_ref)
    dist.destroy_process_group()


def test_compress_comm_hook(native_lib):
    mp.spawn(_worker_compress_hook, args=(2, 29671), nprocs=2, join=True)


def _worker_llama_layout(rank, ws, port):
    """Rebuilt bucket layout on a model whose gradient-ready order is NOT the reversed registration
    order (Llama: norm weights arrive out of line): ours equals what stock DDP's reducer builds,
    bucket by bucket, parameter by parameter, offset by offset."""
    _setup(rank, ws, port)
    import copy
    os.environ["HF_HUB_OFFLINE"] = "1"
    from transformers import LlamaConfig, LlamaForCausalLM
    import ddp_b200
    from oracle_backend import OracleKernels
    torch.manual_seed(0)
    cfg = LlamaConfig(hidden_size=32, num_hidden_layers=2, num_attention_heads=4, num_key_value_heads=2,
                      intermediate_size=64, vocab_size=128, tie_word_embeddings=True)
    base = LlamaForCausalLM(cfg)
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a, bucket_cap_mb=0.006)
    ours = ddp_b200.DistributedDataParallel(b, bucket_cap_mb=0.006, kernels=OracleKernels())
    pid = {id(p): i for i, p in enumerate(ref.module.parameters())}
    seen = {}

    def observe(state, bucket):
        offs = [int((g.data_ptr() - bucket.buffer().data_ptr()) // 4) for g in bucket.gradients()]
        seen.setdefault(state["it"], []).append((bucket.index(), [pid[id(p)] for p in bucket.parameters()], offs))
        fut = dist.all_reduce(bucket.buffer().div_(ws), async_op=True).get_future()
        return fut.then(lambda f: f.value()[0])

    state = {"it": 0}
    ref.register_comm_hook(state, observe)
    for it in range(3):
        state["it"] = it
        g = torch.Generator().manual_seed(100 * it + rank)
        ids = torch.randint(0, 128, (2, 8), generator=g)
        for m in (ref, ours):
            m.zero_grad()
            m(input_ids=ids, labels=ids).loss.backward()
        if it >= 1:
            want = sorted(seen[it])
            got = ours.bucket_layout()
            assert [x[1] for x in want] == [bk["parameters"] for bk in got], it
            assert [x[2] for x in want] == [bk["offsets"] for bk in got], it
        for pr, po in zip(a.parameters(), b.parameters()):
            assert torch.equal(pr.grad, po.grad), it
    assert len(ours.bucket_layout()) > 3
    ready = [pi for bk in ours.bucket_layout() for pi in bk["parameters"]]
    assert ready != list(range(len(ready) - 1, -1, -1)), "this model was chosen because its ready order is not the reversed one"
    dist.destroy_process_group()


def test_rebuilt_layout_on_llama_matches_stock_ddp(native_lib):
    mp.spawn(_worker_llama_layout, args=(2, 29681), nprocs=2, join=True)


def _worker_bf16_buckets(rank, ws, port):
    """grad_bucket_dtype=torch.bfloat16 on an fp32 model: the pack kernel scales and rounds once to
    bf16, the collective runs on bf16 buckets, the unpack kernel widens back into fp32 p.grad. Checked
    against the arithmetic written out with torch ops, and against stock DDP's bf16_compress_hook
    (same idea, rounding order reversed) within one bf16 ulp of the terms."""
    _setup(rank, ws, port)
    import copy
    import ddp_b200
    from oracle_backend import OracleKernels
    from torch.distributed.algorithms.ddp_comm_hooks import default_hooks
    mg = _load_golden_module()
    torch.manual_seed(0)
    base = mg.SmallNet()
    a, b, c = copy.deepcopy(base), copy.deepcopy(base), copy.deepcopy(base)
    ours = ddp_b200.DistributedDataParallel(a, bucket_cap_mb=0.008, kernels=OracleKernels(), grad_bucket_dtype=torch.bfloat16)
    ref = torch.nn.parallel.DistributedDataParallel(b, bucket_cap_mb=0.008)
    ref.register_comm_hook(None, default_hooks.bf16_compress_hook)
    plain = torch.nn.parallel.DistributedDataParallel(c, bucket_cap_mb=0.008)        # fp32 all the way
    with pytest.raises(ValueError):
        ddp_b200.DistributedDataParallel(copy.deepcopy(base), gradient_as_bucket_view=True, kernels=OracleKernels(),
                                         grad_bucket_dtype=torch.bfloat16)
    for s in range(2):
        ids, tgt = mg.live_batch(60 + s, rank)
        for m in (ours, ref, plain):
            m.zero_grad()
            torch.nn.functional.cross_entropy(m(ids).flatten(0, 1), tgt.flatten()).backward()
        # local gradients of both ranks, to write the expected arithmetic out
        for po, pr, pp in zip(a.parameters(), b.parameters(), c.parameters()):
            assert po.grad.dtype == torch.float32
            assert all(bk.dtype == torch.bfloat16 for bk in ours._buckets)
            scale = float(pp.grad.abs().max())      # roundings are relative to the TERMS (which may cancel)
            assert torch.allclose(po.grad, pp.grad, rtol=0, atol=2 ** -7 * scale + 1e-9)   # vs exact fp32 averaging
            assert torch.allclose(po.grad, pr.grad, rtol=0, atol=2 ** -6 * scale + 1e-9)   # vs torch's compress hook
            assert torch.equal(po.grad, po.grad.to(torch.bfloat16).float())            # values are bf16-representable
        n_o = ours.clip_grad_norm_(1e9)
        n_p = torch.nn.utils.clip_grad_norm_(c.parameters(), 1e9)
        assert n_o.dtype == torch.float32 and abs(float(n_o) - float(n_p)) <= 2 ** -7 * float(n_p)
    dist.destroy_process_group()


def test_bf16_gradient_buckets_for_fp32_model(native_lib):
    mp.spawn(_worker_bf16_buckets, args=(2, 29691), nprocs=2, join=True)


def _worker_buffers(rank, ws, port):
    """Buffers (BatchNorm running statistics) next to stock DDP through train / no_sync / eval forwards:
    torch gates the buffer broadcast on require_forward_param_sync, which the PREVIOUS forward set
    (TORCH/nn/parallel/distributed.py:1553-1660), and skips buffers named in
    _ddp_params_and_buffers_to_ignore; ours must land its broadcasts on the same iterations."""
    _setup(rank, ws, port)
    import copy
    import ddp_b200
    from oracle_backend import OracleKernels

    class Net(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.l1 = torch.nn.Linear(12, 16)
            self.bn = torch.nn.BatchNorm1d(16)
            self.l2 = torch.nn.Linear(16, 4)
            self.register_buffer("private", torch.zeros(3))

        def forward(self, x):
            self.private += x.mean() * (1 + dist.get_rank())          # rank-dependent, ignored by DDP below
            return self.l2(self.bn(torch.relu(self.l1(x))))

    torch.manual_seed(0)
    base = Net()
    base._ddp_params_and_buffers_to_ignore = ["private"]
    a, b = copy.deepcopy(base), copy.deepcopy(base)
    ref = torch.nn.parallel.DistributedDataParallel(a)
    ours = ddp_b200.DistributedDataParallel(b, kernels=OracleKernels())
    assert ours.require_forward_param_sync == ref.require_forward_param_sync

    def same_buffers(tag):
        for (n1, x), (n2, y) in zip(a.named_buffers(), b.named_buffers()):
            assert n1 == n2 and torch.equal(x, y), (tag, n1, rank)

    g = torch.Generator().manual_seed(100 + rank)
    plan = ["train", "nosync", "train", "eval", "train", "nosync", "nosync", "train"]
    for it, kind in enumerate(plan):
        x = torch.randn(8, 12, generator=g)
        for model in (ref, ours):
            if kind == "eval":
                model.eval()
                with torch.no_grad():
                    model(x)
                model.train()
            elif kind == "nosync":
                with model.no_sync():
                    model(x).sum().backward()
            else:
                model(x).sum().backward()
        assert ours.require_forward_param_sync == ref.require_forward_param_sync, (it, kind)
        same_buffers((it, kind))
        if kind == "train":
            for p, q in zip(a.parameters(), b.parameters()):
                assert torch.equal(p.grad, q.grad), (it, kind)
                p.grad = None
                q.grad = None
    # the ignored buffer was never broadcast: it still differs between the ranks
    mine = b.private.clone()
    other = [torch.zeros_like(mine) for _ in range(ws)]
    dist.all_gather(other, mine)
    assert not torch.equal(other[0], other[1])
    dist.destroy_process_group()


def test_buffer_broadcast_follows_torch(native_lib):
    mp.spawn(_worker_buffers, args=(2, 29671), nprocs=2, join=True)
