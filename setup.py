"""Packaging for collective_communication_mpi_b200 (the reference has none: it is run from its checkout).

    python -m pip install --no-build-isolation --no-deps .        # builds the native libraries first

The two shared libraries are compiled by the package's own build module (nvcc for sm_100a + g++), exactly as
``python -m collective_communication_mpi_b200.build`` does in a checkout, and shipped as package data together
with the sources they were built from (the loader rebuilds them if they ever go stale). The top-level shims
that give the reference's import paths -- ``mpi_wrapper``, ``model``, ``data``, ``mpi4py`` (defers to a real
mpi4py when one is installed) -- are installed alongside.
"""
import os
import sys

from setuptools import find_namespace_packages, find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


class BuildWithNative(build_py):
    def run(self):
        sys.path.insert(0, ROOT)
        from collective_communication_mpi_b200 import build as native
        native.build(verbose=True)
        super().run()


setup(
    name="collective_communication_mpi_b200",
    version="0.1.0",
    description="B200-native 2-D (data x tensor model parallel) communication framework: fused GEMM + collective "
                "kernels for sm_100a, an mpi4py-compatible front end, tensor-parallel layers",
    python_requires=">=3.10",
    # `data` has no __init__.py, like the reference's (tests import data.data_parallel_preprocess from a namespace package)
    packages=find_packages(include=["collective_communication_mpi_b200*", "mpi_wrapper", "model", "mpi4py"])
    + find_namespace_packages(include=["data"]),
    py_modules=[],
    package_data={"collective_communication_mpi_b200": ["lib/*.so"]},
    install_requires=["numpy", "torch"],
    cmdclass={"build_py": BuildWithNative},
    entry_points={"console_scripts": ["ccmpi-launch=collective_communication_mpi_b200.launch:main",
                                      "ccmpi-info=collective_communication_mpi_b200.info:main"],
                  "pytest11": ["ccmpi_mpi=collective_communication_mpi_b200.pytest_plugin"]},
)
