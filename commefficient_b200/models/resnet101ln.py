"""FEMNIST classifier of the reference's model zoo: a 101-layer bottleneck ResNet whose normalisation
layers are LayerNorms over (C, H, W) -- there is no batch statistic, so clients of any batch size can
be batched together freely.  62 classes, 1x28x28 input.  Parameter names are ``model.<resnet key>``
as in /root/reference/CommEfficient/models/resnet101ln.py:7-13, so checkpoints interchange."""
from torch import nn

from .resnets import Bottleneck, ResNet

__all__ = ["ResNet101LN"]

_DEPTHS_101 = (3, 4, 23, 3)


class ResNet101LN(nn.Module):
    FEMNIST_CLASSES = 62

    def __init__(self, *_positional, num_classes: int = FEMNIST_CLASSES, **_ignored):
        super().__init__()
        self.model = ResNet(Bottleneck, list(_DEPTHS_101), num_classes=num_classes, norm_layer=nn.LayerNorm)

    def forward(self, images):
        return self.model(images)
