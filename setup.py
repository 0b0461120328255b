"""Packaging for commefficient_b200 (reference counterpart: /root/reference/setup.py:1-16, a pure-Python
package).  The Python package is pure Python; the native library ``commefficient_b200/lib/libceb200.so``
is built in-tree by ``build_native.sh`` (nvcc, sm_100a) -- ``python setup.py build_native`` or
``pip install .`` run it when nvcc is available and skip it (CPU-only install) otherwise."""
import os
import shutil
import subprocess

from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


def _build_native():
    if shutil.which("nvcc") is None and not os.path.exists("/usr/local/cuda/bin/nvcc"):
        print("nvcc not found: skipping the native library (CPU-only install)")
        return
    subprocess.check_call(["bash", os.path.join(HERE, "build_native.sh")], cwd=HERE)


class BuildNative(Command):
    description = "compile csrc/*.cu for sm_100a into commefficient_b200/lib/libceb200.so"
    user_options = []

    def initialize_options(self):
        pass

    def finalize_options(self):
        pass

    def run(self):
        _build_native()


class BuildPy(build_py):
    def run(self):
        _build_native()
        super().run()


setup(
    name="commefficient_b200",
    version="0.1.0",
    description="Blackwell-native communication-efficient federated learning (FetchSGD sketching, top-k, FedAvg)",
    packages=find_packages(include=["commefficient_b200", "commefficient_b200.*", "CommEfficient"]),
    package_data={"commefficient_b200": ["lib/*.so"]},
    py_modules=["cv_train", "gpt2_train"],
    python_requires=">=3.10",
    install_requires=["torch>=2.6", "numpy"],
    extras_require={"data": ["torchvision"], "logging": ["tensorboard"], "test": ["pytest", "hypothesis"]},
    cmdclass={"build_native": BuildNative, "build_py": BuildPy},
)
