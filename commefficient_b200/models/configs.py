"""Per-model training presets.  Reference:
/root/reference/CommEfficient/models/configs.py:1-37 (there ``PiecewiseLinear`` is
used without being imported, so the class cannot be instantiated; fixed here)."""
from ..utils.schedules import PiecewiseLinear

__all__ = ["ModelConfig", "FixupResNet50Config"]


class ModelConfig:
    """Attribute bag whose fields are copied onto the argparse namespace."""

    def set_args(self, args):
        for name, val in vars(self).items():
            setattr(args, name, val)
        return args


class FixupResNet50Config(ModelConfig):
    """Step schedule 0.1 -> 0.01 -> 0.001 -> 0.0001 at epochs 30 / 60 / 90."""

    def __init__(self):
        self.model_config = {}
        self.lr_scale = 0.1
        self.lr_schedule = PiecewiseLinear([0, 30, 30, 60, 60, 90, 90, 100],
                                           [0.1, 0.1, 0.01, 0.01, 0.001, 0.001, 0.0001, 0.0001])
