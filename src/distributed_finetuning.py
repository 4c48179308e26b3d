"""DDP fine-tuning entry point: the file REF/scripts/run_node0.sh:16 and
REF/tests/test_distributed_finetuning.py:4 name (the reference ships it as
src/distributed_inference.py and never trains). Same public names — setup, cleanup, CustomDataset,
get_model_response, main — with the training step the reference's README promises put where its
loop only called a remote model (REF/src/distributed_inference.py:61-76).

On a CUDA device the step runs on the B200 path (ddp_b200: bucket pack / NCCL allreduce on a side
stream / overlapped norm / fused clip+AdamW). Without CUDA the script uses stock torch
(DistributedDataParallel on gloo + clip_grad_norm_ + AdamW): that IS the reference's CPU path, and
the B200 package itself has no CPU mode.
"""
import logging
import os
import sys

import torch
import torch.distributed as dist
from torch.utils.data import DataLoader, Dataset
from torch.utils.data.distributed import DistributedSampler

_HERE = os.path.dirname(os.path.abspath(__file__))
for _p in (_HERE, os.path.dirname(_HERE)):
    if _p not in sys.path:
        sys.path.insert(0, _p)

from config import CONFIG  # noqa: E402
from utils import gpu_tensor_operation, setup_logging, train_step  # noqa: E402


def setup(rank, world_size):
    """Join the process group. NCCL on CUDA (one process per GPU, device = LOCAL_RANK), gloo
    otherwise. torchrun's rendezvous address wins over CONFIG's (the reference overwrote it,
    REF/src/distributed_inference.py:15-16, which breaks single-box launches)."""
    os.environ.setdefault("MASTER_ADDR", CONFIG["MASTER_ADDR"])
    os.environ.setdefault("MASTER_PORT", CONFIG["MASTER_PORT"])
    if torch.cuda.is_available():
        torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", rank % torch.cuda.device_count())))
        dist.init_process_group("nccl", rank=rank, world_size=world_size)
    else:
        dist.init_process_group("gloo", rank=rank, world_size=world_size)
    dist.barrier()


def cleanup():
    dist.destroy_process_group()


class CustomDataset(Dataset):
    """{"text", "label"} rows backed by two lists (REF/src/distributed_inference.py:23-32)."""

    def __init__(self, texts, labels):
        self.texts, self.labels = texts, labels

    def __len__(self):
        return len(self.texts)

    def __getitem__(self, idx):
        return {"text": self.texts[idx], "label": self.labels[idx]}


def get_model_response(prompt):
    """The reference asked a remote LiteLLM proxy here (REF/src/distributed_inference.py:34-41).
    That needs a network and is outside the training path: stubbed."""
    return "Error: remote model disabled (no network in this build)"


def synthetic_imdb(rows: int = 250, seed: int = 0):
    """Stand-in for load_dataset("imdb", split="train[:1%]") (REF/src/distributed_inference.py:56),
    unreachable offline: 250 rows (1 % of 25,000) of {"text": str, "label": 0/1}."""
    g = torch.Generator().manual_seed(seed)
    words = ["movie", "film", "plot", "actor", "great", "awful", "scene", "story", "boring", "loved",
             "director", "ending", "music", "slow", "funny", "dark", "cast", "script", "worst", "best"]
    texts, labels = [], []
    for _ in range(rows):
        n = int(torch.randint(20, 120, (1,), generator=g))
        idx = torch.randint(0, len(words), (n,), generator=g).tolist()
        texts.append(" ".join(words[i] for i in idx))
        labels.append(int(torch.randint(0, 2, (1,), generator=g)))
    return texts, labels


def build_model(name: str, device, dtype):
    """Random-init causal LM of the BASELINE.json configurations (no checkpoints offline)."""
    os.environ.setdefault("HF_HUB_OFFLINE", "1")
    from transformers import GPT2Config, GPT2LMHeadModel, LlamaConfig, LlamaForCausalLM
    torch.manual_seed(0)
    if name == "tiny":
        model = GPT2LMHeadModel(GPT2Config(n_layer=2, n_head=2, n_embd=64, vocab_size=512, n_positions=256))
    elif name == "gpt2":
        model = GPT2LMHeadModel(GPT2Config())
    elif name == "llama-1b":
        model = LlamaForCausalLM(LlamaConfig(hidden_size=2048, num_hidden_layers=16, num_attention_heads=32,
                                             num_key_value_heads=8, intermediate_size=8192, vocab_size=128256,
                                             tie_word_embeddings=True))
    else:
        raise ValueError(f"unknown DDPB200_MODEL {name!r}")
    return model.to(device=device, dtype=dtype)


def main():
    setup_logging()
    try:
        rank, world_size = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        setup(rank, world_size)
        on_gpu = torch.cuda.is_available()
        device = torch.device("cuda", torch.cuda.current_device()) if on_gpu else torch.device("cpu")
        logging.info(f"Using device: {device}")

        texts, labels = synthetic_imdb()
        data = CustomDataset(texts, labels)
        sampler = DistributedSampler(data, num_replicas=world_size, rank=rank)
        loader = DataLoader(data, batch_size=4, sampler=sampler, pin_memory=torch.cuda.is_available())

        model = build_model(CONFIG["TRAIN_MODEL"], device, torch.bfloat16 if on_gpu else torch.float32)
        if on_gpu:
            import ddp_b200
            ddp = ddp_b200.DistributedDataParallel(model, device_ids=[device.index], gradient_as_bucket_view=True)
            opt = ddp_b200.AdamW(model.parameters(), lr=CONFIG["LR"])
            clip = lambda mx: ddp.clip_grad_norm_(mx, fused_optimizer=opt)      # noqa: E731
        else:
            ddp = torch.nn.parallel.DistributedDataParallel(model)
            opt = torch.optim.AdamW(model.parameters(), lr=CONFIG["LR"])
            clip = lambda mx: torch.nn.utils.clip_grad_norm_(model.parameters(), mx)   # noqa: E731

        step = 0
        for epoch in range(CONFIG["EPOCHS"]):
            logging.info(f"Starting epoch {epoch}")
            sampler.set_epoch(epoch)
            for batch in loader:
                loss, total = train_step(ddp, opt, batch["text"], batch["label"], clip=clip,
                                         max_norm=CONFIG["MAX_NORM"], seq_len=CONFIG["SEQ_LEN"])
                step += 1
                if rank == 0 and (step == 1 or step % 10 == 0):
                    probe = gpu_tensor_operation(batch["text"][0], device)
                    logging.info(f"step {step} loss {float(loss):.4f} grad_norm {float(total):.4f} probe {probe:.2f}")
                if CONFIG["MAX_STEPS"] and step >= CONFIG["MAX_STEPS"]:
                    break
            if CONFIG["MAX_STEPS"] and step >= CONFIG["MAX_STEPS"]:
                break
        if rank == 0:
            logging.info(f"done: {step} steps, last loss {float(loss):.4f}")
        cleanup()
    except Exception as e:
        logging.error(f"An error occurred in the main function: {e}")
        sys.exit(1)


if __name__ == "__main__":
    main()
