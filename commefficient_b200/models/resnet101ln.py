"""ResNet-101 with LayerNorm for FEMNIST (62 classes, 1x28x28 input).
Reference: /root/reference/CommEfficient/models/resnet101ln.py:7-13."""
import torch.nn as nn

from .resnets import resnet101

__all__ = ["ResNet101LN"]


class ResNet101LN(nn.Module):
    def __init__(self, *args, num_classes=62, **kwargs):
        super().__init__()
        self.model = resnet101(num_classes=num_classes, norm_layer=nn.LayerNorm)

    def forward(self, x):
        return self.model(x)
