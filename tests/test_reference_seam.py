"""The reference's own unit tests (REF/tests/test_distributed_finetuning.py:19-44), run against
this repo's src/ — which, unlike the reference, is importable under the module name those tests
use — plus a 2-rank launch of the script through its launch line on CPU/gloo."""
import os
import subprocess
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_custom_dataset():
    from src.distributed_finetuning import CustomDataset
    ds = CustomDataset(["text1", "text2", "text3"], [0, 1, 0])
    assert len(ds) == 3 and ds[1]["text"] == "text2" and ds[1]["label"] == 1


def test_process_batch_shape_and_type():
    from src.utils import process_batch
    loss = process_batch(torch.nn.Linear(1, 1), ["prompt1", "prompt2"], [0, 1], lambda p: "mocked response")
    assert isinstance(loss, torch.Tensor) and loss.shape == torch.Size([])


def test_gpu_tensor_operation_value():
    from src.utils import gpu_tensor_operation
    assert gpu_tensor_operation("test", torch.device("cpu")) == 112.0       # SURVEY.md §4 probe


def test_process_batch_runs_a_causal_lm():
    from src.distributed_finetuning import build_model
    from src.utils import process_batch
    model = build_model("tiny", torch.device("cpu"), torch.float32)
    loss = process_batch(model, ["a great movie", "boring plot"], [1, 0], seq_len=16)
    assert loss.shape == torch.Size([]) and loss.requires_grad and 4.0 < float(loss) < 9.0    # ~ln(512)


def test_script_runs_two_ranks_on_cpu():
    env = dict(os.environ, DDPB200_MODEL="tiny", DDPB200_MAX_STEPS="3", DDPB200_SEQ_LEN="32", CUDA_VISIBLE_DEVICES="",
               OMP_NUM_THREADS="1")
    env.pop("MASTER_ADDR", None)
    env.pop("MASTER_PORT", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29733", os.path.join(ROOT, "src", "distributed_finetuning.py")]
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=300)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "done: 3 steps" in res.stderr


def test_gpt2_small_two_ranks_cpu_gloo():
    """BASELINE.json configs[0]: GPT-2 small (124 M, random init) DDP fine-tune on the small synthetic
    dataset, CPU/gloo, world_size 2 — the plumbing case. Two steps through the launch line."""
    env = dict(os.environ, DDPB200_MODEL="gpt2", DDPB200_MAX_STEPS="2", DDPB200_SEQ_LEN="32", CUDA_VISIBLE_DEVICES="",
               OMP_NUM_THREADS="4")
    env.pop("MASTER_ADDR", None)
    env.pop("MASTER_PORT", None)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29735", os.path.join(ROOT, "src", "distributed_finetuning.py")]
    res = subprocess.run(cmd, cwd=ROOT, env=env, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stderr[-2000:]
    assert "done: 2 steps" in res.stderr
    # a random-init GPT-2 starts near ln(50257) = 10.8
    import re
    m = re.search(r"step 1 loss ([0-9.]+)", res.stderr)
    assert m and 9.5 < float(m.group(1)) < 12.0, res.stderr[-500:]
