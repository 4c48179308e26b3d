"""bench.py's synthetic workloads have exactly the parameter lists of the BASELINE.json model
configurations (checked against the numel lists recorded from transformers in the golden file)."""
import json
import os

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
