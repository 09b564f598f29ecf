"""Packaging. The native libraries (sm_100a kernels, CPU index builders) are built in-tree by
``python -m epfl_megatron_b200.ops.build`` (nvcc + g++); ``pip install -e .`` runs that step when nvcc is
on PATH and otherwise installs the Python packages only (CPU / Gloo paths keep working)."""
import os
import shutil
import subprocess
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py


class build_with_native(build_py):
    def run(self):
        if shutil.which("nvcc") or os.path.exists("/usr/local/cuda/bin/nvcc"):
            subprocess.check_call([sys.executable, "-m", "epfl_megatron_b200.ops.build"])
        else:
            print("nvcc not found: skipping the CUDA kernels (PyTorch fallbacks only)")
        super().run()


setup(
    name="epfl-megatron-b200",
    version="0.1.0",
    description="Blackwell-native 3D-parallel LLM training with the Megatron-LLM surface",
    packages=find_packages(include=("epfl_megatron_b200*", "megatron*", "tasks*", "tools*", "weights2megatron*")),
    py_modules=["finetune", "pretrain_gpt", "pretrain_bert", "pretrain_t5", "pretrain_ict", "verify_correctness"],
    package_data={"epfl_megatron_b200": ["ops/_lib/*.so", "csrc/*", "static/*"]},
    python_requires=">=3.10",
    install_requires=["torch", "numpy", "regex", "sentencepiece", "einops"],
    cmdclass={"build_py": build_with_native},
)
