"""Federated datasets, the round sampler, transforms and synthetic-data writers
(reference package: /root/reference/CommEfficient/data_utils/)."""
from .fed_dataset import FedDataset
from .fed_sampler import FedSampler, ShardedEvalSampler
from .fed_vision import FedCIFAR10, FedCIFAR100, FedEMNIST, FedImageNet
from .fed_persona import (FedPERSONA, personachat_collate_fn, SimpleTokenizer, BPETokenizer, load_tokenizer, SPECIAL_TOKENS,
                          MODEL_INPUTS, ATTR_TO_SPECIAL_TOKEN)
from .synthetic import (write_synthetic_cifar, write_synthetic_emnist, write_synthetic_persona, write_synthetic_imagenet,
                        synthetic_image_batch)
from . import transforms

_TRANSFORM_NAMES = set(transforms.__all__) - {"NORMALIZATION"}


def __getattr__(name):
    if name in _TRANSFORM_NAMES:
        return getattr(transforms, name)
    raise AttributeError(name)
