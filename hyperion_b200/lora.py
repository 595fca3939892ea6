"""LoRA: low-rank adapters injected into ``nn.Linear`` projections, peft-compatible export.

The reference uses ``peft.LoraConfig(r=16, lora_alpha=32, lora_dropout=0.05, target_modules=
[q_proj, k_proj, v_proj, o_proj], bias="none", task_type=CAUSAL_LM)`` + ``get_peft_model``
(02_development/distributed_utils.py:469-474); peft is not installed here, so the injection,
the forward ``base(x) + B(A(dropout(x))) * alpha/r`` and the ``save_pretrained`` layout
(``adapter_config.json`` + ``adapter_model.safetensors`` with keys
``base_model.model.<path>.lora_A.weight``) are implemented directly.

On CUDA the rank-r update is fused into the base GEMM: ``t = x A^T`` is a skinny GEMM and
``t B^T * scale`` is accumulated into the *same* TMEM accumulator as ``x W^T`` by the dual-K
mode of gemm_tcgen05 (``ops.lora_linear``), so the adapter costs no extra pass over the output.
"""
from __future__ import annotations

import json
import math
from pathlib import Path

import torch
import torch.nn as nn

from . import ops

DEFAULT_TARGETS = ("q_proj", "k_proj", "v_proj", "o_proj")


class LoRALinear(nn.Module):
    """Wraps a frozen ``nn.Linear``; attribute names follow peft (``base_layer``, ``lora_A``, ``lora_B``)."""

    def __init__(self, base: nn.Linear, r: int = 16, alpha: int = 32, dropout: float = 0.05):
        super().__init__()
        self.base_layer = base
        self.r, self.alpha, self.scaling, self.p = r, alpha, alpha / r, dropout
        self.lora_A = nn.Linear(base.in_features, r, bias=False)
        self.lora_B = nn.Linear(r, base.out_features, bias=False)
        nn.init.kaiming_uniform_(self.lora_A.weight, a=math.sqrt(5))
        nn.init.zeros_(self.lora_B.weight)
        self.lora_A.weight.data = self.lora_A.weight.data.to(base.weight.dtype).to(base.weight.device)
        self.lora_B.weight.data = self.lora_B.weight.data.to(base.weight.dtype).to(base.weight.device)
        for p in base.parameters():
            p.requires_grad_(False)

    @property
    def weight(self):
        return self.base_layer.weight

    @property
    def bias(self):
        return self.base_layer.bias

    def lora_forward(self, x, residual=None):
        xd = ops.dropout(x, self.p, True) if (self.training and self.p > 0) else x
        return ops.lora_linear(x, xd, self.base_layer.weight, self.base_layer.bias, self.lora_A.weight,
                               self.lora_B.weight, self.scaling, residual=residual)

    def forward(self, x):
        return self.lora_forward(x)

    def merged_weight(self) -> torch.Tensor:
        return self.base_layer.weight + self.scaling * (self.lora_B.weight @ self.lora_A.weight)


def inject_lora(model: nn.Module, r: int = 16, alpha: int = 32, dropout: float = 0.05,
                target_modules=DEFAULT_TARGETS) -> nn.Module:
    """Freeze every parameter, wrap the target projections, return the same model (peft's ``get_peft_model``)."""
    for p in model.parameters():
        p.requires_grad_(False)
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if name in target_modules and isinstance(child, nn.Linear):
                setattr(parent, name, LoRALinear(child, r, alpha, dropout))
    model._lora_config = dict(r=r, lora_alpha=alpha, lora_dropout=dropout, target_modules=list(target_modules),
                              bias="none", task_type="CAUSAL_LM", peft_type="LORA")
    return model


def lora_state_dict(model: nn.Module) -> dict:
    """Adapter tensors under peft's key convention."""
    out = {}
    for name, mod in model.named_modules():
        if isinstance(mod, LoRALinear):
            out[f"base_model.model.{name}.lora_A.weight"] = mod.lora_A.weight.detach().cpu().contiguous()
            out[f"base_model.model.{name}.lora_B.weight"] = mod.lora_B.weight.detach().cpu().contiguous()
    return out


def trainable_parameters(model: nn.Module) -> int:
    return sum(p.numel() for p in model.parameters() if p.requires_grad)


def save_pretrained(model: nn.Module, out_dir: str | Path, base_model_name: str = "") -> Path:
    """peft ``save_pretrained`` layout.  Accepts a DDP/FSDP wrapper (the reference calls it on
    the DDP wrapper and crashes): ``.module`` is unwrapped first."""
    model = getattr(model, "module", model)
    out = Path(out_dir)
    out.mkdir(parents=True, exist_ok=True)
    cfg = dict(getattr(model, "_lora_config", {}))
    cfg.setdefault("peft_type", "LORA")
    cfg["base_model_name_or_path"] = base_model_name
    cfg["inference_mode"] = True
    (out / "adapter_config.json").write_text(json.dumps(cfg, indent=2))
    sd = lora_state_dict(model)
    try:
        from safetensors.torch import save_file

        save_file(sd, str(out / "adapter_model.safetensors"))
    except ImportError:  # pragma: no cover
        torch.save(sd, out / "adapter_model.bin")
    return out


def load_adapter(model: nn.Module, adapter_dir: str | Path) -> None:
    adapter_dir = Path(adapter_dir)
    f = adapter_dir / "adapter_model.safetensors"
    if f.exists():
        from safetensors.torch import load_file

        sd = load_file(str(f))
    else:
        sd = torch.load(adapter_dir / "adapter_model.bin", map_location="cpu")
    mods = dict(model.named_modules())
    for k, v in sd.items():
        path, which, _ = k[len("base_model.model."):].rsplit(".", 2)
        lin = getattr(mods[path], which)
        lin.weight.data.copy_(v.to(lin.weight.dtype))
