"""Model zoo.  Every capitalised export is a valid ``--model`` choice, as in the
reference (/root/reference/CommEfficient/utils.py:114-118)."""
from .resnet9 import ResNet9
from .resnet18 import ResNet18, FixupResNet18
from .fixup import FixupResNet9, FixupResNet50
from .resnets import (ResNet, resnet18, resnet34, resnet50, resnet101, resnet152, resnext50_32x4d,
                      resnext101_32x8d, wide_resnet50_2, wide_resnet101_2)
from .resnet101ln import ResNet101LN
from .configs import ModelConfig, FixupResNet50Config
from .gpt2 import GPT2Config, GPT2DoubleHeadsModel, OpenAIGPTDoubleHeadsModel


def model_names():
    return sorted(n for n in globals() if n[0].isupper() and not n.startswith("__"))
