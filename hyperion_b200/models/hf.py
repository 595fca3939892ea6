"""Hugging Face checkpoint directories <-> the Llama / GPT-2 models of this package.

The reference fine-tunes *pretrained* Llama-2 (``AutoModelForCausalLM.from_pretrained(model_id, token=hf_token)``,
02_development/distributed_utils.py:462-467).  This image has no network, so the trainers default to random weights of the same
architecture; on a machine that holds the checkpoint (``config.json`` + ``*.safetensors`` / ``pytorch_model*.bin``), point
``--model_id`` (``train_llama_fsdp(model_id=...)``) at the directory and the weights are read from it.

* ``LlamaForCausalLM`` here has HF's parameter names, so the tensors are copied name by name (shard by shard, tensor by tensor: the
  7B checkpoint never exists twice in host memory).
* ``GPT2LMHeadModel`` keeps HF's names without the ``transformer.`` prefix and stores ``nn.Linear`` ``[out, in]`` weights where HF's
  ``Conv1D`` stores ``[in, out]``; ``lm_head.weight`` is the tied ``wte.weight``.
* ``save_hf`` writes a directory that ``transformers`` loads back (``config.json`` + ``model.safetensors``).

Features of newer checkpoints that the kernels here do not implement (RoPE scaling, attention / MLP biases, untied GPT-2 heads) are
refused with a message instead of being loaded wrongly.
"""
from __future__ import annotations

import json
from pathlib import Path
from typing import Iterator

import torch

from .gpt2 import GPT2Config, GPT2LMHeadModel
from .llama import LlamaConfig, LlamaForCausalLM

_GPT2_TRANSPOSED = ("attn.c_attn.weight", "attn.c_proj.weight", "mlp.c_fc.weight", "mlp.c_proj.weight")


def is_checkpoint_dir(path) -> bool:
    return path is not None and Path(str(path)).is_dir() and (Path(str(path)) / "config.json").exists()


def read_config(path) -> dict:
    return json.loads((Path(path) / "config.json").read_text())


def iter_tensors(path) -> Iterator[tuple[str, torch.Tensor]]:
    """(name, CPU tensor) for every tensor of the checkpoint, one at a time; safetensors shards first, ``.bin`` files otherwise."""
    path = Path(path)
    st = sorted(path.glob("*.safetensors"))
    if st:
        from safetensors import safe_open

        for f in st:
            with safe_open(str(f), framework="pt", device="cpu") as fh:
                for k in fh.keys():
                    yield k, fh.get_tensor(k)
        return
    bins = sorted(path.glob("pytorch_model*.bin"))
    if not bins:
        raise FileNotFoundError(f"no *.safetensors or pytorch_model*.bin under {path}")
    for f in bins:
        sd = torch.load(f, map_location="cpu", weights_only=True)
        yield from sd.items()
        del sd


def _build(ctor, dtype, device):
    prev = torch.get_default_dtype()
    if dtype is not None:
        torch.set_default_dtype(dtype)
    try:
        if device is not None:
            with torch.device(device):
                return ctor()
        return ctor()
    finally:
        torch.set_default_dtype(prev)


def _copy_in(model: torch.nn.Module, tensors, rename=lambda k, v: (k, v)) -> None:
    own = dict(model.named_parameters())
    own.update(dict(model.named_buffers()))
    seen = set()
    for k, v in tensors:
        r = rename(k, v)
        if r is None:
            continue
        k, v = r
        if k not in own:
            raise KeyError(f"checkpoint tensor {k!r} has no counterpart in {type(model).__name__}")
        if own[k].shape != v.shape:
            raise ValueError(f"{k}: checkpoint shape {tuple(v.shape)} != model shape {tuple(own[k].shape)}")
        with torch.no_grad():
            own[k].copy_(v)               # converts dtype and moves to the parameter's device
        seen.add(k)
    missing = [k for k, p in model.named_parameters() if k not in seen]
    if missing:
        raise KeyError(f"checkpoint has no tensor for {missing[:5]}{' ...' if len(missing) > 5 else ''}")


# ------------------------------------------------------------------------------------------------------------------- Llama
def llama_config_from_hf(cfg: dict) -> LlamaConfig:
    if cfg.get("rope_scaling"):
        raise NotImplementedError(f"rope_scaling={cfg['rope_scaling']} (Llama-3.1 style) is not implemented by the RoPE kernel here")
    if cfg.get("attention_bias") or cfg.get("mlp_bias"):
        raise NotImplementedError("attention_bias / mlp_bias checkpoints are not supported (the Llama-2 projections have no bias)")
    if cfg.get("tie_word_embeddings"):
        raise NotImplementedError("tie_word_embeddings=True: this LlamaForCausalLM keeps a separate lm_head like Llama-2")
    heads = cfg["num_attention_heads"]
    pad = cfg.get("pad_token_id")
    return LlamaConfig(vocab_size=cfg["vocab_size"], hidden_size=cfg["hidden_size"], intermediate_size=cfg["intermediate_size"],
                       num_hidden_layers=cfg["num_hidden_layers"], num_attention_heads=heads,
                       num_key_value_heads=cfg.get("num_key_value_heads") or heads, rms_norm_eps=cfg.get("rms_norm_eps", 1e-5),
                       rope_theta=float(cfg.get("rope_theta", 10000.0)),
                       pad_token_id=pad if pad is not None else cfg.get("eos_token_id", 2))   # the reference pads with eos


def llama_from_pretrained(path, dtype: torch.dtype | None = None, device=None) -> LlamaForCausalLM:
    model = _build(lambda: LlamaForCausalLM(llama_config_from_hf(read_config(path))), dtype, device)
    _copy_in(model, iter_tensors(path), lambda k, v: None if k.endswith("rotary_emb.inv_freq") else (k, v))
    return model


# ------------------------------------------------------------------------------------------------------------------- GPT-2
def gpt2_config_from_hf(cfg: dict) -> GPT2Config:
    act = cfg.get("activation_function", "gelu_new")
    if act != "gelu_new":
        raise NotImplementedError(f"activation_function={act!r}: the fused MLP here implements GPT-2's tanh GELU (gelu_new)")
    if cfg.get("tie_word_embeddings", True) is False:
        raise NotImplementedError("untied GPT-2 head: GPT2LMHeadModel here ties lm_head to wte")
    return GPT2Config(vocab_size=cfg["vocab_size"], n_positions=cfg.get("n_positions", 1024), n_embd=cfg["n_embd"],
                      n_layer=cfg["n_layer"], n_head=cfg["n_head"], layer_norm_eps=cfg.get("layer_norm_epsilon", 1e-5),
                      dropout=float(cfg.get("resid_pdrop", 0.0)))


def _gpt2_rename(k: str, v: torch.Tensor):
    k = k[len("transformer."):] if k.startswith("transformer.") else k
    if k == "lm_head.weight" or k.endswith(".attn.bias") or k.endswith(".attn.masked_bias"):
        return None                        # tied head; causal-mask buffers of older checkpoints
    if k.endswith(_GPT2_TRANSPOSED):
        v = v.t()
    return k, v


def gpt2_from_pretrained(path, dtype: torch.dtype | None = None, device=None) -> GPT2LMHeadModel:
    model = _build(lambda: GPT2LMHeadModel(gpt2_config_from_hf(read_config(path))), dtype, device)
    _copy_in(model, iter_tensors(path), _gpt2_rename)
    return model


# ------------------------------------------------------------------------------------------------------------------ export
def save_hf(model, out_dir, dtype: torch.dtype | None = None) -> Path:
    """Write ``config.json`` + ``model.safetensors`` in the layout ``transformers.AutoModelForCausalLM.from_pretrained`` reads.
    ``model``: a ``LlamaForCausalLM`` / ``GPT2LMHeadModel`` of this package (unwrap DDP / FSDP first, or pass a gathered state_dict
    through ``model.load_state_dict``)."""
    from safetensors.torch import save_file

    model = getattr(model, "module", model)
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    sd = {k: v.detach().to("cpu", dtype or v.dtype) for k, v in model.state_dict().items()}
    if isinstance(model, LlamaForCausalLM):
        c = model.config
        cfg = {"architectures": ["LlamaForCausalLM"], "model_type": "llama", "vocab_size": c.vocab_size, "hidden_size": c.hidden_size,
               "intermediate_size": c.intermediate_size, "num_hidden_layers": c.num_hidden_layers,
               "num_attention_heads": c.num_attention_heads, "num_key_value_heads": c.num_key_value_heads, "hidden_act": "silu",
               "rms_norm_eps": c.rms_norm_eps, "rope_theta": c.rope_theta, "pad_token_id": c.pad_token_id, "tie_word_embeddings": False,
               "attention_bias": False, "mlp_bias": False, "max_position_embeddings": 4096}
        tensors = {k: v.contiguous() for k, v in sd.items()}
    elif isinstance(model, GPT2LMHeadModel):
        c = model.config
        cfg = {"architectures": ["GPT2LMHeadModel"], "model_type": "gpt2", "vocab_size": c.vocab_size, "n_positions": c.n_positions,
               "n_ctx": c.n_positions, "n_embd": c.n_embd, "n_layer": c.n_layer, "n_head": c.n_head,
               "layer_norm_epsilon": c.layer_norm_eps, "activation_function": "gelu_new", "resid_pdrop": c.dropout,
               "embd_pdrop": c.dropout, "attn_pdrop": c.dropout, "tie_word_embeddings": True}
        tensors = {"transformer." + k: (v.t() if k.endswith(_GPT2_TRANSPOSED) else v).contiguous() for k, v in sd.items()}
    else:
        raise TypeError(f"save_hf: unsupported model type {type(model).__name__}")
    cfg["torch_dtype"] = str(next(iter(tensors.values())).dtype).replace("torch.", "")
    (out / "config.json").write_text(json.dumps(cfg, indent=2))
    save_file(tensors, str(out / "model.safetensors"), metadata={"format": "pt"})
    return out
