"""Packaging for the B200-native framework.

`pip install -e .` (or `python setup.py build_ext --inplace`) compiles `csrc/*.cu` for sm_100a into
`distributedvolunteercomputing_b200/ops/libdvc_b200.so` with nvcc (see `ops/build.py`; the library is a plain C ABI
loaded through ctypes, so there is no Python extension module to link) and installs the package plus the two
reference-compatible entry points.
"""
import os
import sys

from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


def _build_native():
    sys.path.insert(0, ROOT)
    from distributedvolunteercomputing_b200.ops import build as native_build
    native_build.build()


class BuildNative(Command):
    description = "compile the sm_100a native library in-tree (nvcc)"
    user_options = []

    def initialize_options(self):
        pass

    def finalize_options(self):
        pass

    def run(self):
        _build_native()


class BuildPyWithNative(build_py):
    def run(self):
        try:
            _build_native()
        except Exception as e:   # a CPU-only install still gets the pure-Python parts (CpuEngine, protocol, scheduler)
            print(f"warning: native library not built ({e}); GPU engine unavailable until `python -m "
                  "distributedvolunteercomputing_b200.ops.build` succeeds", file=sys.stderr)
        super().run()


setup(
    name="distributedvolunteercomputing_b200",
    version="0.1.0",
    description="Blackwell-native distributed video inference with volunteer-computing semantics",
    packages=find_packages(include=["distributedvolunteercomputing_b200", "distributedvolunteercomputing_b200.*"]),
    package_data={"distributedvolunteercomputing_b200.ops": ["libdvc_b200.so"]},
    python_requires=">=3.10",
    install_requires=["numpy", "torch"],
    extras_require={"net": ["pyzmq", "opencv-python-headless"]},
    entry_points={"console_scripts": ["dvc-server=distributedvolunteercomputing_b200.coordinator:main",
                                      "dvc-worker=distributedvolunteercomputing_b200.client:main"]},
    cmdclass={"build_ext": BuildNative, "build_native": BuildNative, "build_py": BuildPyWithNative},
)
