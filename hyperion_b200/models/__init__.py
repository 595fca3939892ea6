"""Model zoo: every architecture the reference trains or benchmarks,
plus GPT-2 and the real ViT-B/16 that BASELINE.json adds.  state_dict keys match the reference's
classes (torch.nn.Transformer*, torchvision ResNet / ViT, HF Llama)."""
from .cnn import FallbackCNN
from .custom_transformer import CustomTransformer, create_custom_transformer
from .gpt2 import GPT2Config, GPT2LMHeadModel
from .llama import LlamaConfig, LlamaForCausalLM
from .resnet import create_resnet50, resnet18, resnet50
from .simple_lm import CompileBenchLM, NotebookTransformerLM, SimpleTransformerLM
from .vit import create_vit_model, vit_b_16

__all__ = ["FallbackCNN", "CustomTransformer", "create_custom_transformer", "GPT2Config", "GPT2LMHeadModel",
           "LlamaConfig", "LlamaForCausalLM", "create_resnet50", "resnet18", "resnet50", "CompileBenchLM",
           "NotebookTransformerLM", "SimpleTransformerLM", "create_vit_model", "vit_b_16"]
