"""pip packaging (reference: pyspark/setup.py + make-dist.sh, which ship the python API together with the jar and the
native .so files). ``pip install .`` / ``python setup.py bdist_wheel`` compile the sm_100a CUDA library and the C++
host runtime in-tree (``python -m bigdl_b200.build``: nvcc + g++, no GPU needed) and package them as data of
``bigdl_b200``; the ``bigdl`` compatibility package is installed alongside."""
import os
import re
import subprocess
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py

HERE = os.path.dirname(os.path.abspath(__file__))


def version():
    with open(os.path.join(HERE, "bigdl_b200", "__init__.py")) as f:
        return re.search(r'__version__\s*=\s*"([^"]+)"', f.read()).group(1)


class BuildWithNative(build_py):
    def run(self):
        if os.environ.get("BIGDL_B200_SKIP_NATIVE") != "1":
            subprocess.check_call([sys.executable, "-m", "bigdl_b200.build"], cwd=HERE)
        super().run()


setup(
    name="bigdl-b200",
    version=version(),
    description="Blackwell-native distributed deep-learning framework with the BigDL API surface",
    long_description=open(os.path.join(HERE, "README.md")).read(),
    long_description_content_type="text/markdown",
    packages=find_packages(include=["bigdl_b200", "bigdl_b200.*", "bigdl", "bigdl.*"]),
    package_data={"bigdl_b200": ["lib/*.so", "../include/*.h", "../conf/*.conf"]},
    include_package_data=True,
    python_requires=">=3.10",
    install_requires=["torch>=2.4", "numpy", "protobuf"],
    extras_require={"vision": ["opencv-python-headless"], "frames": ["pandas", "pyarrow"], "test": ["pytest"]},
    entry_points={"console_scripts": [
        "bigdl-b200-convert-model=bigdl_b200.cli.convert_model:main",
        "bigdl-b200-distri-perf=bigdl_b200.cli.distri_optimizer_perf:main",
        "bigdl-b200-local-perf=bigdl_b200.cli.local_optimizer_perf:main",
        "bigdl-b200-perf=bigdl_b200.cli.perf:main",
        "bigdl-b200-imagenet-seqfile=bigdl_b200.cli.imagenet_seqfile_generator:main",
        "bigdl-b200-coco-seqfile=bigdl_b200.cli.coco_seqfile_generator:main",
    ]},
    cmdclass={"build_py": BuildWithNative},
)
