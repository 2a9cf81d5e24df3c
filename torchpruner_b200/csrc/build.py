"""Build libtorchpruner_b200.so in-tree with nvcc for sm_100a (B200).

    python -m torchpruner_b200.csrc.build [--force] [--verbose]

The library is a plain C-ABI shared object (see include/torchpruner_b200.h): it does not link
against torch or Python, so it is compiled directly with nvcc rather than through
torch.utils.cpp_extension.  nvcc cross-compiles without a GPU present.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
LIB_NAME = "libtorchpruner_b200.so"
LIB_PATH = os.path.join(HERE, LIB_NAME)

SOURCES = ["api.cu", "reduce_rows.cu", "shapley.cu", "gather.cu", "nan_trick.cu", "conv_tf32x3.cu", "incremental.cu"]
HEADERS = ["common.cuh", os.path.join(ROOT, "include", "torchpruner_b200.h")]

NVCC_FLAGS = [
    "-O3",
    "-std=c++17",
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr",
]


def _nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found; the CUDA toolkit is required to build torchpruner_b200")
    return exe


def _stale():
    if not os.path.exists(LIB_PATH):
        return True
    t = os.path.getmtime(LIB_PATH)
    deps = [os.path.join(HERE, s) for s in SOURCES] + [
        h if os.path.isabs(h) else os.path.join(HERE, h) for h in HEADERS
    ] + [os.path.abspath(__file__)]
    return any(os.path.getmtime(d) > t for d in deps)


def build(force=False, verbose=False):
    """Compile every .cu into one shared library next to the sources. Returns its path."""
    if not force and not _stale():
        return LIB_PATH
    nvcc = _nvcc()
    objs = []
    for src in SOURCES:
        obj = os.path.join(HERE, src.replace(".cu", ".o"))
        cmd = [nvcc, *NVCC_FLAGS, "-c", os.path.join(HERE, src), "-o", obj]
        if verbose:
            cmd.insert(1, "-Xptxas=-v")
            print(" ".join(cmd), flush=True)
        subprocess.run(cmd, check=True, cwd=HERE)
        objs.append(obj)
    tmp = LIB_PATH + ".tmp"
    cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", tmp, *objs]
    if verbose:
        print(" ".join(cmd), flush=True)
    subprocess.run(cmd, check=True, cwd=HERE)
    os.replace(tmp, LIB_PATH)
    return LIB_PATH


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="--verbose" in sys.argv)
    print(path)
