"""Packaging for jax_llama_b200.  `pip install -e .` / `python setup.py build_ext --inplace` compile the CUDA
library in-tree with nvcc for sm_100a (see jax_llama_b200/ops/build.py); nothing is built by torch."""
from setuptools import Command, find_packages, setup
from setuptools.command.build_py import build_py


class build_cuda(Command):
    description = "compile csrc/*.cu into jax_llama_b200/ops/_C/libjllama_b200.so (nvcc, sm_100a)"
    user_options = [("force", "f", "rebuild even if the source hash matches")]

    def initialize_options(self):
        self.force = False

    def finalize_options(self):
        pass

    def run(self):
        from jax_llama_b200.ops.build import build_library
        build_library(force=bool(self.force), verbose=True)


class build_py_with_cuda(build_py):
    def run(self):
        self.run_command("build_cuda")
        super().run()


setup(
    name="jax_llama_b200",
    version="0.1.0",
    description="B200-native (sm_100a) LLaMA-1/2/3 inference with the API and checkpoint format of JAX_llama",
    packages=find_packages(include=["jax_llama_b200", "jax_llama_b200.*"]),
    package_data={"jax_llama_b200.ops": ["_C/*.so"]},
    python_requires=">=3.10",
    install_requires=["torch", "numpy", "safetensors", "sentencepiece", "tiktoken"],
    cmdclass={"build_cuda": build_cuda, "build_ext": build_cuda, "build_py": build_py_with_cuda},
)
