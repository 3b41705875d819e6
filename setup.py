"""Packaging: ``pip install .`` (or ``pip install -e .``) builds the native libraries in-tree and installs the
``hetseq_9cme_b200`` package with its console entry points.

The reference builds one Cython extension from its setup.py (setup.py:30-37 there); here the native code is two
C-ABI shared objects — ``libb200seq.so`` (every csrc/*.cu, nvcc, sm_100a only) and ``libb200seq_host.so``
(csrc/host/*.cpp) — produced by ``tools/build_native.py`` and shipped as package data, loaded with ctypes.
nvcc cross-compiles, so no GPU is needed to build; without nvcc only the host library is built and the package
still works on CPUs (the CUDA ops then fail loudly on a GPU machine instead of falling back).
"""
import os
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


class BuildWithNative(build_py):
    def run(self):
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import build_native
        build_native.build_host()
        try:
            build_native.build_cuda()
        except RuntimeError as e:             # nvcc missing: CPU-only install
            print("warning: CUDA library not built (%s)" % str(e).splitlines()[0])
        super().run()


setup(
    name="hetseq_9cme_b200",
    version="2.0.0",
    description="Blackwell-native (sm_100a) data-parallel training framework with HetSeq's CLI, tasks and checkpoint format",
    packages=find_packages(include=["hetseq_9cme_b200", "hetseq_9cme_b200.*"]),
    package_data={"hetseq_9cme_b200": ["lib/*.so"]},
    python_requires=">=3.10",
    install_requires=["numpy", "torch>=2.4"],
    extras_require={"ner": ["transformers", "datasets"], "bench": ["nvidia-ml-py"]},
    entry_points={"console_scripts": [
        "hetseq-b200-train=hetseq_9cme_b200.train:cli_main",
        "hetseq-b200-eval-mnist=hetseq_9cme_b200.eval_mnist:main",
        "hetseq-b200-eval-ner=hetseq_9cme_b200.eval_bert_fine_tuning_ner:main",
    ]},
    cmdclass={"build_py": BuildWithNative},
    zip_safe=False,
)
