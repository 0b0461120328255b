"""Augmentation pipelines (torchvision), same recipes as
/root/reference/CommEfficient/data_utils/transforms.py:13-75: CIFAR reflect-pad-4
random crop + horizontal flip + normalise; FEMNIST pad-2 crop, resized crop, +-5 deg
rotation; ImageNet 224 standard.  Built lazily so importing the package does not
require torchvision."""
from __future__ import annotations

__all__ = ["cifar10_train_transforms", "cifar10_test_transforms", "cifar100_train_transforms",
           "cifar100_test_transforms", "femnist_train_transforms", "femnist_test_transforms",
           "imagenet_train_transforms", "imagenet_val_transforms", "NORMALIZATION"]

NORMALIZATION = {
    "CIFAR10": ((0.4914, 0.4822, 0.4465), (0.2471, 0.2435, 0.2616)),
    "CIFAR100": ((0.5071, 0.4867, 0.4408), (0.2675, 0.2565, 0.2761)),
    "EMNIST": ((0.9637,), (0.1597,)),
    "ImageNet": ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225)),
}


def _build():
    from torchvision import transforms as T

    def cifar(name, train):
        mean, std = NORMALIZATION[name]
        aug = [T.RandomCrop(32, padding=4, padding_mode="reflect"), T.RandomHorizontalFlip()] if train else []
        return T.Compose(aug + [T.ToTensor(), T.Normalize(mean, std)])

    mean, std = NORMALIZATION["EMNIST"]
    femnist_train = T.Compose([
        T.RandomCrop(28, padding=2, padding_mode="constant", fill=1.0),
        T.RandomResizedCrop(28, scale=(0.8, 1.2), ratio=(4. / 5., 5. / 4.)),
        T.RandomRotation(5, fill=1.0),
        T.ToTensor(), T.Normalize(mean, std)])
    femnist_test = T.Compose([T.ToTensor(), T.Normalize(mean, std)])
    mean, std = NORMALIZATION["ImageNet"]
    tail = [T.ToTensor(), T.Normalize(mean, std)]
    imagenet_train = T.Compose([T.RandomResizedCrop(224), T.RandomHorizontalFlip()] + tail)
    imagenet_val = T.Compose([T.Resize(int(224 * 1.14)), T.CenterCrop(224)] + tail)
    return {
        "cifar10_train_transforms": cifar("CIFAR10", True), "cifar10_test_transforms": cifar("CIFAR10", False),
        "cifar100_train_transforms": cifar("CIFAR100", True), "cifar100_test_transforms": cifar("CIFAR100", False),
        "femnist_train_transforms": femnist_train, "femnist_test_transforms": femnist_test,
        "imagenet_train_transforms": imagenet_train, "imagenet_val_transforms": imagenet_val,
    }


_cache = None


def __getattr__(name):
    global _cache
    if name in __all__ and name != "NORMALIZATION":
        if _cache is None:
            _cache = _build()
        return _cache[name]
    raise AttributeError(name)
