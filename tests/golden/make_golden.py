"""Generate the golden vectors under tests/golden/ by running the REAL torch code paths
(DistributedDataParallel on gloo, clip_grad_norm_, torch.optim.AdamW, and
dist._compute_bucket_assignment_by_size) on CPU in the build container.

The upstream repo has no implementation and no known-answer tests for this path
(SURVEY.md §0, §8c); its arithmetic lives in torch, so torch run here is the reference the
oracle (oracle/ddp_oracle.c) is pinned against. Re-run with:

    python tests/golden/make_golden.py

Outputs (committed):
    bucket_layouts.json   a-1  bucket assignment for the BASELINE.json model shapes + edge cases
    adamw_steps.npz       a-5/a-6  clip_grad_norm_ + AdamW, 12 steps, fp32 and bf16, default & fused CPU
    pack_scale.npz        a-2  g * (1/world_size) rounding, fp32 and bf16
    ddp_live.npz          a-1..a-6 live 2-rank gloo DDP + clip + AdamW on a small model, 4 steps
"""
from __future__ import annotations

import json
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
MiB = 1024 * 1024


def bits(t: torch.Tensor) -> np.ndarray:
    t = t.detach().contiguous()
    if t.dtype == torch.bfloat16:
        return t.view(torch.int16).numpy().view(np.uint16).copy()
    return t.numpy().copy()


# ---------------------------------------------------------------------------------------------
def model_param_shapes():
    """(name, numel list, elem_size) for the BASELINE.json configs, enumerated the way
    DDP._build_params_for_reducer does (TORCH/nn/parallel/distributed.py:1338-1382)."""
    from transformers import GPT2Config, GPT2LMHeadModel, LlamaConfig, LlamaForCausalLM

    def enum(model):
        seen, out = set(), []
        for _, mod in model.named_modules():
            for _, p in mod.named_parameters(recurse=False):
                if p.requires_grad and p not in seen:
                    seen.add(p)
                    out.append(p.numel())
        return out

    cfgs = {
        "gpt2-small-fp32": (lambda: GPT2LMHeadModel(GPT2Config()), 4),
        "llama-3.2-1b-bf16": (lambda: LlamaForCausalLM(LlamaConfig(
            hidden_size=2048, num_hidden_layers=16, num_attention_heads=32, num_key_value_heads=8,
            intermediate_size=8192, vocab_size=128256, tie_word_embeddings=True)), 2),
        "llama-3.1-8b-bf16": (lambda: LlamaForCausalLM(LlamaConfig(
            hidden_size=4096, num_hidden_layers=32, num_attention_heads=32, num_key_value_heads=8,
            intermediate_size=14336, vocab_size=128256, tie_word_embeddings=False)), 2),
        "llama-3.1-70b-4layer-bf16": (lambda: LlamaForCausalLM(LlamaConfig(
            hidden_size=8192, num_hidden_layers=4, num_attention_heads=64, num_key_value_heads=8,
            intermediate_size=28672, vocab_size=128256, tie_word_embeddings=False)), 2),
    }
    out = {}
    for name, (ctor, es) in cfgs.items():
        with torch.device("meta"):
            m = ctor()
        out[name] = (enum(m), es)
    return out


def torch_assign(numels, elem_sizes, limits, sparse=None, order=None):
    dts = {4: torch.float32, 2: torch.bfloat16, 8: torch.float64}
    tensors = [torch.empty(n, dtype=dts[e], device="meta") for n, e in zip(numels, elem_sizes)]
    sparse = sparse if sparse is not None else [False] * len(tensors)
    if order is not None:
        tensors = [tensors[i] for i in order]
        b, l = dist._compute_bucket_assignment_by_size(tensors, limits, sparse, order)
    else:
        b, l = dist._compute_bucket_assignment_by_size(tensors, limits, sparse)
    return [list(map(int, x)) for x in b], [int(x) for x in l]


def gen_bucket_layouts():
    cases = []
    rng = np.random.default_rng(0)

    def add(name, numels, es, limits, sparse=None, order=None):
        b, l = torch_assign(numels, es, limits, sparse, order)
        cases.append(dict(name=name, numel=list(map(int, numels)), elem_size=list(map(int, es)),
                          limits=[int(x) for x in limits], sparse=sparse, order=order,
                          buckets=b, bucket_limits=l))

    for name, (numels, es) in model_param_shapes().items():
        n = len(numels)
        # DDP construction: one bucket (limits=[sys.maxsize]), registration order
        add(f"{name}/init", numels, [es] * n, [sys.maxsize])
        # rebuild after iteration 0: reversed registration order as the grad-ready order
        order = list(range(n - 1, -1, -1))
        add(f"{name}/rebuild-1MiB-25MiB", numels, [es] * n, [MiB, 25 * MiB], order=order)
        add(f"{name}/rebuild-25MiB-25MiB", numels, [es] * n, [25 * MiB, 25 * MiB], order=order)
        add(f"{name}/find-unused-1MiB-25MiB", numels, [es] * n, [MiB, 25 * MiB])
    # edge cases
    add("four-256-fp32/1KiB", [256] * 4, [4] * 4, [1024])
    add("single-tensor", [7], [4], [MiB])
    add("zero-numel-inside", [10, 0, 300, 0, 5], [4] * 5, [1024])
    add("exactly-at-limit", [256, 256, 256], [4] * 3, [1024, 2048])
    numels = [int(x) for x in rng.integers(1, 5000, size=40)]
    es = [int(x) for x in rng.choice([2, 4], size=40)]
    add("mixed-dtype-random", numels, es, [4096, 16384])
    add("mixed-dtype-random/ready-order", numels, es, [4096, 16384],
        order=[int(x) for x in rng.permutation(40)])
    sparse = [bool(x) for x in (rng.random(40) < 0.15)]
    add("sparse-flags", numels, [4] * 40, [4096, 16384], sparse=sparse)
    add("sparse-flags/ready-order", numels, [4] * 40, [4096, 16384], sparse=sparse,
        order=[int(x) for x in rng.permutation(40)])
    add("three-limits", numels, [4] * 40, [2048, 8192, 32768])
    with open(os.path.join(HERE, "bucket_layouts.json"), "w") as f:
        json.dump(cases, f, separators=(",", ":"))
    print("bucket_layouts.json:", len(cases), "cases")


# ---------------------------------------------------------------------------------------------
def gen_adamw_steps():
    """clip_grad_norm_ + AdamW on three ragged tensors, 12 steps, four torch CPU paths."""
    out = {}
    sizes = [1031, 4096, 77]
    steps = 12
    hp = dict(lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2)
    max_norm = 0.05  # small enough that clipping is active on most steps
    out["sizes"] = np.array(sizes)
    out["steps"] = np.array(steps)
    out["max_norm"] = np.array(max_norm)
    out["hp"] = np.array([hp["lr"], hp["betas"][0], hp["betas"][1], hp["eps"], hp["weight_decay"]])
    for dt_name, dt in (("f32", torch.float32), ("bf16", torch.bfloat16)):
        g0 = torch.Generator().manual_seed(1234)
        p0 = [(torch.randn(n, generator=g0) * 0.02).to(dt) for n in sizes]
        grads = [[(torch.randn(n, generator=g0) * (0.01 * (1 + s % 3))).to(dt) for n in sizes]
                 for s in range(steps)]
        for i, p in enumerate(p0):
            out[f"{dt_name}/p0/{i}"] = bits(p)
        for s in range(steps):
            for i, g in enumerate(grads[s]):
                out[f"{dt_name}/g/{s}/{i}"] = bits(g)
        for path, kw in (("default", dict(foreach=False, fused=False)), ("foreach", dict(foreach=True)),
                         ("fused", dict(fused=True))):
            params = [torch.nn.Parameter(p.clone()) for p in p0]
            opt = torch.optim.AdamW(params, **hp, **kw)
            norms = []
            for s in range(steps):
                for p, g in zip(params, grads[s]):
                    p.grad = g.clone()
                tn = torch.nn.utils.clip_grad_norm_(params, max_norm)
                norms.append(float(tn))
                if s == 0:
                    for i, p in enumerate(params):
                        out[f"{dt_name}/{path}/clipped_g0/{i}"] = bits(p.grad)
                opt.step()
            out[f"{dt_name}/{path}/norms"] = np.array(norms, dtype=np.float64)
            for i, p in enumerate(params):
                st = opt.state[p]
                out[f"{dt_name}/{path}/p/{i}"] = bits(p)
                out[f"{dt_name}/{path}/m/{i}"] = bits(st["exp_avg"])
                out[f"{dt_name}/{path}/v/{i}"] = bits(st["exp_avg_sq"])
    np.savez_compressed(os.path.join(HERE, "adamw_steps.npz"), **out)
    print("adamw_steps.npz:", len(out), "arrays")


def gen_pack_scale():
    out = {}
    g0 = torch.Generator().manual_seed(7)
    x = torch.randn(5000, generator=g0) * 0.01
    x[:8] = torch.tensor([0.0, -0.0, 1e-38, -1e-38, 65504.0, 3.3895e38, float("inf"), 1.0])
    for dt_name, dt in (("f32", torch.float32), ("bf16", torch.bfloat16)):
        xd = x.to(dt)
        out[f"{dt_name}/x"] = bits(xd)
        for ws in (1, 2, 3, 4, 6, 8):
            # no-hook DDP: bucket_view = grad * (1/ws)  (Appendix A-4)
            out[f"{dt_name}/ws{ws}"] = bits(xd.mul(1.0 / ws))
    out["f32_to_bf16"] = bits(x.to(torch.bfloat16))
    np.savez_compressed(os.path.join(HERE, "pack_scale.npz"), **out)
    print("pack_scale.npz:", len(out), "arrays")


# ---------------------------------------------------------------------------------------------
class SmallNet(torch.nn.Module):
    """Small stand-in with the layer kinds that matter for bucketing: big and tiny tensors mixed."""

    def __init__(self):
        super().__init__()
        self.emb = torch.nn.Embedding(50, 24)
        self.fc1 = torch.nn.Linear(24, 48)
        self.ln = torch.nn.LayerNorm(48)
        self.fc2 = torch.nn.Linear(48, 48)
        self.head = torch.nn.Linear(48, 50, bias=False)

    def forward(self, ids):
        h = self.emb(ids)
        h = torch.relu(self.fc1(h))
        h = self.ln(h)
        h = h + torch.tanh(self.fc2(h))
        return self.head(h)


LIVE = dict(world_size=2, steps=4, bucket_cap_mb=0.008, lr=1e-2, max_norm=0.5, batch=4, seq=6)


def live_batch(step, rank):
    g = torch.Generator().manual_seed(1000 * step + rank)
    ids = torch.randint(0, 50, (LIVE["batch"], LIVE["seq"]), generator=g)
    tgt = torch.randint(0, 50, (LIVE["batch"], LIVE["seq"]), generator=g)
    return ids, tgt


def _live_worker(rank, ws, port, ret):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=ws)
    torch.manual_seed(0)
    torch.set_num_threads(1)
    net = SmallNet()
    ddp = torch.nn.parallel.DistributedDataParallel(net, bucket_cap_mb=LIVE["bucket_cap_mb"])
    params = list(net.parameters())
    pid = {id(p): i for i, p in enumerate(params)}
    layout_log = []

    def hook(state, bucket):
        offs = [int((g.data_ptr() - bucket.buffer().data_ptr()) // 4) for g in bucket.gradients()]
        layout_log.append((int(state["step"]), int(bucket.index()), [pid[id(p)] for p in bucket.parameters()],
                           offs, int(bucket.buffer().numel()), bool(bucket.is_last())))
        local = bucket.buffer().clone()
        state["local"].append((int(bucket.index()), local))
        fut = dist.all_reduce(bucket.buffer().div_(ws), async_op=True).get_future()
        return fut.then(lambda f: f.value()[0])

    state = {"step": 0, "local": []}
    ddp.register_comm_hook(state, hook)
    opt = torch.optim.AdamW(params, lr=LIVE["lr"])
    rec = {}
    for i, p in enumerate(params):
        rec[f"p0/{i}"] = bits(p)
    for s in range(LIVE["steps"]):
        state["step"] = s
        state["local"] = []
        ids, tgt = live_batch(s, rank)
        loss = torch.nn.functional.cross_entropy(ddp(ids).flatten(0, 1), tgt.flatten())
        loss.backward()
        # local (pre-averaging) grads of this rank, re-assembled per parameter from the hook's copies
        this_step = [e for e in layout_log if e[0] == s]
        for (bidx, local) in state["local"]:
            entry = [e for e in this_step if e[1] == bidx][0]
            for pi, off in zip(entry[2], entry[3]):
                n = params[pi].numel()
                rec[f"local_g/{s}/{rank}/{pi}"] = local[off:off + n].numpy().copy()
        for i, p in enumerate(params):
            rec[f"avg_g/{s}/{i}"] = bits(p.grad)
        tn = torch.nn.utils.clip_grad_norm_(params, LIVE["max_norm"])
        rec[f"norm/{s}"] = np.array(float(tn))
        opt.step()
        opt.zero_grad()
        for i, p in enumerate(params):
            rec[f"p/{s}/{i}"] = bits(p)
        rec[f"loss/{s}/{rank}"] = np.array(float(loss))
    rec["layout"] = np.array(json.dumps(layout_log))
    rec["rebuilt_bucket_sizes"] = np.array(str(ddp._get_ddp_logging_data().get("rebuilt_bucket_sizes", "")))
    ret[rank] = rec
    dist.destroy_process_group()


def gen_ddp_live():
    import torch.multiprocessing as mp
    ws = LIVE["world_size"]
    mgr = mp.Manager()
    ret = mgr.dict()
    mp.spawn(_live_worker, args=(ws, 29631, ret), nprocs=ws, join=True)
    out = {}
    r0, r1 = ret[0], ret[1]
    for k, v in r0.items():
        out[k] = v
    for k, v in r1.items():
        if k.startswith("local_g/") or k.startswith("loss/"):
            out[k] = v
        elif k.startswith("p/"):
            assert np.array_equal(v, r0[k]), f"ranks diverged at {k}"
    out["numel"] = np.array([int(p.numel()) for p in SmallNet().parameters()])
    out["cfg"] = np.array(json.dumps(LIVE))
    np.savez_compressed(os.path.join(HERE, "ddp_live.npz"), **out)
    print("ddp_live.npz:", len(out), "arrays; layout:", str(out["layout"])[:300])


if __name__ == "__main__":
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    gen_bucket_layouts()
    gen_adamw_steps()
    gen_pack_scale()
    gen_ddp_live()
