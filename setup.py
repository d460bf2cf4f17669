"""Packaging (reference Makefile:4,7,29,39 expects `pip install .` and the distribution name pytorch-distbelief; the
reference ships no setup.py).  The CUDA library is built by build_native.py during build_py."""
import subprocess
import sys
from pathlib import Path

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

ROOT = Path(__file__).resolve().parent


class BuildWithNative(build_py):
    def run(self):
        subprocess.check_call([sys.executable, str(ROOT / "build_native.py")])
        super().run()


setup(
    name="pytorch-distbelief",
    version="0.1.0",
    license="MIT",
    description="Blackwell-native asynchronous/synchronous SGD (parameter server over NVLink) with the API of distributed_ML_pytorch",
    packages=find_packages(include=["distributed_ml_pytorch_b200*", "asgd*"]),
    package_data={"distributed_ml_pytorch_b200._native": ["libdmlb.so"]},
    include_package_data=True,
    data_files=[("include", ["include/dmlb.h"])],       # the C ABI of libdmlb.so
    python_requires=">=3.10",
    install_requires=["torch>=2.4", "numpy"],
    cmdclass={"build_py": BuildWithNative},
)
