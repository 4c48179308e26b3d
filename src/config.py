"""Settings read by src/distributed_finetuning.py. The reference imports a `config.CONFIG` that it
does not ship (REF/src/distributed_inference.py:12, REF/.gitignore:29); this is the minimal one.
Every value can be overridden from the environment; torchrun's MASTER_ADDR / MASTER_PORT win."""
import os

CONFIG = {
    "MASTER_ADDR": os.environ.get("MASTER_ADDR", "127.0.0.1"),
    "MASTER_PORT": os.environ.get("MASTER_PORT", "29500"),
    # remote LiteLLM proxy (REF/src/distributed_inference.py:34-41,53-54): needs a network, out of scope
    "MODEL_NAME": os.environ.get("DDPB200_REMOTE_MODEL", "stub/none"),
    "API_KEY": os.environ.get("OPENAI_API_KEY", ""),
    "API_BASE": os.environ.get("DDPB200_API_BASE", ""),
    # the local model that is fine-tuned: "tiny" | "gpt2" | "llama-1b"
    "TRAIN_MODEL": os.environ.get("DDPB200_MODEL", "gpt2"),
    "SEQ_LEN": int(os.environ.get("DDPB200_SEQ_LEN", "128")),
    "EPOCHS": int(os.environ.get("DDPB200_EPOCHS", "3")),
    "MAX_STEPS": int(os.environ.get("DDPB200_MAX_STEPS", "0")),     # 0 = whole epochs
    "LR": float(os.environ.get("DDPB200_LR", "5e-5")),
    "MAX_NORM": float(os.environ.get("DDPB200_MAX_NORM", "1.0")),
}
