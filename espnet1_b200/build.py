"""Build libb200ctc.so (the C-ABI CUDA library) in-tree with nvcc for sm_100a.

``python -m espnet1_b200.build`` or ``build_library()``.  The .so is written
next to the sources so that it travels with the repository snapshot to the GPU
box; nothing is installed into site-packages and no JIT cache is used.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libb200ctc.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-O3",
    "--expt-relaxed-constexpr",
    "-Xptxas", "-v",
]


def _extra_flags():
    """tuning builds only: B200CTC_NVCC_EXTRA="-DB200CTC_REF_OWN=2" python -m espnet1_b200.build --force"""
    return os.environ.get("B200CTC_NVCC_EXTRA", "").split()


def _nvcc() -> str:
    cand = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(cand):
        raise RuntimeError("nvcc not found: libb200ctc.so cannot be built")
    return cand


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


FLAGS_FILE = LIB + ".flags"  # the extra flags the library on disk was built with


def _built_flags() -> str:
    try:
        return open(FLAGS_FILE).read().strip()
    except OSError:
        return ""


def _stale() -> bool:
    if not os.path.exists(LIB):
        return True
    # a tuning build (B200CTC_NVCC_EXTRA=-DB200CTC_TUNING ...) left in the tree must not be
    # taken for the shipped library, and vice versa
    if _built_flags() != " ".join(_extra_flags()):
        return True
    t = os.path.getmtime(LIB)
    deps = sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + [
        os.path.join(os.path.dirname(HERE), "include", "b200ctc.h")]
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force: bool = False, verbose: bool = False) -> str:
    """Compile every .cu under csrc/ into one shared library; returns its path."""
    if not force and not _stale():
        return LIB
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    nvcc = _nvcc()
    procs = []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        cmd = [nvcc, *NVCC_FLAGS, *_extra_flags(), "-c", src, "-o", obj]
        procs.append((src, obj, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    objs = []
    log = []
    for src, obj, p in procs:
        out = p.communicate()[0].decode()
        log.append(f"== {os.path.basename(src)}\n{out}")
        if p.returncode != 0:
            raise RuntimeError(f"nvcc failed for {src}:\n{out}")
        objs.append(obj)
    with open(os.path.join(objdir, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    cmd = [nvcc, "-shared", "-o", LIB, *objs, "-gencode", "arch=compute_100a,code=sm_100a"]
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)
    if r.returncode != 0:
        raise RuntimeError("link failed:\n" + r.stdout.decode())
    with open(FLAGS_FILE, "w") as f:
        f.write(" ".join(_extra_flags()))
    if verbose:
        print("\n".join(log))
    return LIB


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
