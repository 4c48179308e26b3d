"""Helpers of the fine-tuning script. Same names and signatures as REF/src/utils.py:5-28
(`setup_logging`, `process_batch`, `gpu_tensor_operation`); `process_batch` is the seam where the
reference only simulated a loss (REF/src/utils.py:12-23) and where the real step now sits."""
import logging
from typing import Callable, List, Optional, Sequence

import torch


def setup_logging(level: int = logging.INFO) -> None:
    logging.basicConfig(level=level, format="%(asctime)s - %(levelname)s - %(message)s",
                        datefmt="%Y-%m-%d %H:%M:%S")


def gpu_tensor_operation(text: str, device) -> float:
    """Mean code point of `text`, computed on `device` (kept for the reference's test,
    REF/tests/test_distributed_finetuning.py:38-44)."""
    codes = torch.tensor([ord(ch) for ch in text], dtype=torch.float32, device=device)
    return codes.mean().item()


def encode_texts(texts: Sequence[str], seq_len: int, vocab_size: int, device) -> torch.Tensor:
    """Byte-level stand-in tokenizer (no tokenizer files exist offline): UTF-8 bytes, offset into
    the model's vocabulary, right-padded / truncated to seq_len. Returns int64 [batch, seq_len]."""
    ids = torch.zeros(len(texts), seq_len, dtype=torch.int64)
    for r, t in enumerate(texts):
        b = list(t.encode("utf-8"))[:seq_len]
        if b:
            ids[r, :len(b)] = torch.tensor(b, dtype=torch.int64) % vocab_size
    return ids.to(device, non_blocking=True)


def lm_loss(model, ids: torch.Tensor) -> torch.Tensor:
    """Next-token cross entropy of a causal LM on `ids` (HF models return it as `.loss`)."""
    out = model(input_ids=ids, labels=ids)
    return out.loss if hasattr(out, "loss") else out


def process_batch(model, prompts: List[str], labels, get_model_response: Optional[Callable] = None,
                  *, seq_len: int = 128, vocab_size: Optional[int] = None) -> torch.Tensor:
    """Loss of one batch as a 0-d tensor (shape and type as the reference's placeholder). If `model`
    is a causal LM (has a `config.vocab_size`, or vocab_size is given) the prompts are tokenised and
    run through it; otherwise — the reference's own unit test passes an unrelated nn.Linear — the
    loss is a binary cross entropy of random logits against `labels`, exactly what the reference
    simulated. `get_model_response` (the remote LiteLLM call) is not needed for training and is only
    invoked when given, for parity with the reference's call pattern."""
    if get_model_response is not None:
        for p in prompts:
            get_model_response(p)
    cfg = getattr(getattr(model, "module", model), "config", None)
    vs = vocab_size or getattr(cfg, "vocab_size", None)
    if vs is None:
        tgt = torch.as_tensor(list(labels), dtype=torch.int64)
        return torch.nn.functional.cross_entropy(torch.randn(len(tgt), 2), tgt)
    device = next(model.parameters()).device
    return lm_loss(model, encode_texts(prompts, seq_len, vs, device))


def train_step(ddp_model, optimizer, prompts: List[str], labels, *, clip, max_norm: float, seq_len: int = 128):
    """forward -> backward (bucket allreduce overlapped) -> clip -> AdamW -> zero_grad.
    `clip(max_norm)` is a callable so the caller decides between torch's clip_grad_norm_ over the
    parameters and the DDP object's overlapped one. Returns (loss, total_norm) as tensors."""
    loss = process_batch(ddp_model, prompts, labels, seq_len=seq_len)
    loss.backward()
    total = clip(max_norm)
    optimizer.step()
    optimizer.zero_grad()
    return loss.detach(), total
