"""bench.py's synthetic workloads have exactly the parameter lists of the BASELINE.json model
configurations (checked against the numel lists recorded from transformers in the golden file)."""
import json
import os

import torch

import bench


def test_llama_shapes_match_transformers(golden_dir):
    cases = {c["name"]: c for c in json.load(open(os.path.join(golden_dir, "bucket_layouts.json")))}
    for wl, gold in (("llama-3.2-1b", "llama-3.2-1b-bf16/init"), ("llama-3.1-8b", "llama-3.1-8b-bf16/init"),
                     ("llama-3.1-70b-4layer", "llama-3.1-70b-4layer-bf16/init")):
        shapes = bench.llama_param_shapes(**bench.WORKLOADS[wl])
        numel = []
        for s in shapes:
            n = 1
            for d in s:
                n *= d
            numel.append(n)
        assert numel == cases[gold]["numel"], wl
    assert sum(cases["llama-3.2-1b-bf16/init"]["numel"]) == 1_235_814_400


def test_grad_ready_order_is_a_permutation():
    for wl, cfg in bench.WORKLOADS.items():
        n = len(bench.llama_param_shapes(**cfg))
        order = bench.grad_ready_order(n, cfg["layers"], cfg["tie"])
        assert sorted(order) == list(range(n)), wl
        assert order[-1] == 0          # the embedding's gradient is ready last


def test_grad_ready_order_and_shapes_match_real_autograd():
    """bench.py replays gradients in the order a real Llama backward makes them ready: checked here
    against autograd itself on a 3-layer LlamaForCausalLM (tied and untied), together with the
    parameter shape list in DDP registration order."""
    import torch
    from transformers import LlamaConfig, LlamaForCausalLM
    for tie in (True, False):
        cfg = LlamaConfig(hidden_size=64, num_hidden_layers=3, num_attention_heads=4, num_key_value_heads=2,
                          intermediate_size=128, vocab_size=256, tie_word_embeddings=tie)
        torch.manual_seed(0)
        model = LlamaForCausalLM(cfg)
        seen, params = set(), []
        for _, mod in model.named_modules():
            for _, p in mod.named_parameters(recurse=False):
                if p not in seen:
                    seen.add(p)
                    params.append(p)
        assert [tuple(p.shape) for p in params] == [tuple(s) for s in bench.llama_param_shapes(64, 3, 4, 2, 128, 256, tie)]
        index = {id(p): i for i, p in enumerate(params)}
        order = []
        for p in params:
            p.register_post_accumulate_grad_hook(lambda q: order.append(index[id(q)]))
        ids = torch.randint(0, 256, (2, 16))
        model(input_ids=ids, labels=ids).loss.backward()
        assert order == bench.grad_ready_order(len(params), 3, tie), tie


def test_reference_arm_under_torchrun_prints_one_line_from_rank_0():
    """`bench.py --impl reference` launched the way the driver launches it at N>1: rank 0 alone works (it drives
    its own N-rank gloo group on the host cores), the other rank exits 0 without output, and stdout is ONE JSON
    line with the contract's keys, on the same `config` as our arm would print."""
    import json
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(root, "bench.py"), "--impl", "reference", "--gpus", "2",
           "--workload", "tiny", "--steps", "3", "--warmup", "1"]
    res = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    lines = [ln for ln in res.stdout.splitlines() if ln.strip()]
    assert len(lines) == 1, lines
    d = json.loads(lines[0])
    assert d["impl"] == "reference" and d["n_gpus"] == 2 and d["steps"] == 3 and d["warmup"] == 1
    assert d["gpu_launches"] == 0 and d["vs_baseline"] is None and d["higher_is_better"] is True
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert d["cpu_baseline"]["kind"] == "reference" and d["cpu_baseline"]["value"] == d["value"]
    assert "gloo, 2 ranks" in d["implementation"]
    assert d["config"] == bench.workload_config("tiny", 20, 1377536, 2)


def test_parity_metric_of_the_bench_self_check():
    """bench.parity_check's error measure on CPU tensors: a round-to-nearest result of the fp32 sum is <= 0.5 ulp
    whatever order the reference was added in, one step off reads as 1.5 (2 across a power of two) at most, and a bf16 running sum — whose
    error comes from the inputs, not from rounding the result — is far outside; the plain ulp distance, kept for the
    record, explodes on cancelling inputs while the measure does not blame the result for it."""
    torch.manual_seed(0)
    ws, n = 8, 1 << 20
    xs = [(torch.randn(n) * 0.01 / ws).bfloat16() for _ in range(ws)]
    fwd, bwd, sumabs = torch.zeros(n), torch.zeros(n), torch.zeros(n)
    for x in xs:
        fwd += x.float()
        sumabs += x.float().abs()
    for x in reversed(xs):
        bwd += x.float()
    rn = fwd.bfloat16()
    assert bench._err_in_ulps(rn, bwd, sumabs, torch.bfloat16) <= 0.5
    up = (rn.float() + bench._ulp_of(rn.float(), torch.bfloat16)).bfloat16()
    # (a step of ulp(rn) taken across a power of two towards zero is two steps of the binade it lands in)
    assert 0.5 < bench._err_in_ulps(up, bwd, sumabs, torch.bfloat16) <= 2.0 + 1e-3
    seq = xs[0].clone()
    for x in xs[1:]:
        seq = (seq.float() + x.float()).bfloat16()
    assert bench._err_in_ulps(seq, bwd, sumabs, torch.bfloat16) > 8
    assert bench._max_ulp(rn, bwd.bfloat16()) <= 1
    assert bench._ulp_of(torch.tensor([1.0, 0.75, 0.0]), torch.bfloat16).tolist() == [2.0 ** -7, 2.0 ** -8, 2.0 ** -133]
    assert bench._ulp_of(torch.tensor([1.0]), torch.float32).tolist() == [2.0 ** -23]
