"""`pip install -e .` (or `python setup.py build`): builds the native library in-tree with the Makefile first
(nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -> cuda_mpi_reductions_b200/lib/libb2r.so), then packages the
Python face around it. The library is loaded with ctypes, so there is no CPython extension module to compile."""
import os
import subprocess

from setuptools import setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


class BuildNativeFirst(build_py):
    def run(self):
        jobs = str(min(16, os.cpu_count() or 4))
        subprocess.check_call(["make", "-C", HERE, "-j", jobs, "lib"])
        super().run()


setup(cmdclass={"build_py": BuildNativeFirst})
