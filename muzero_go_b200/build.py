"""Build muzero_go_b200/libmzgo.so with nvcc for sm_100a (in-tree, so the .so travels with gpurun).

    python -m muzero_go_b200.build [--force]
"""
import glob
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
SO = os.path.join(HERE, "libmzgo.so")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
              "-Xcompiler", "-fPIC", "-Wno-deprecated-gpu-targets"]


def sources():
    return sorted(glob.glob(os.path.join(CSRC, "*.cu")))


def _deps():
    return sources() + glob.glob(os.path.join(CSRC, "*.cuh")) + [os.path.join(os.path.dirname(HERE), "include", "mzgo.h")]


def _digest():
    import hashlib
    h = hashlib.sha256()
    for f in sorted(_deps()):
        h.update(os.path.basename(f).encode())
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


CHECKED_FLAG = "-DMZGO_CHECKED"     # device-side bounds assertions (csrc/common.cuh, MZGO_DCHECK)


def up_to_date():
    """Content hash of the sources (not mtimes: the tree is copied to the GPU box by gpurun)."""
    if not (os.path.exists(SO) and os.path.exists(SO + ".hash")):
        return False
    with open(SO + ".hash") as fh:
        return fh.read().strip() == _digest()


def build(force=False, verbose=False, checked=False, defines=()):
    """checked=True: the same library with MZGO_DCHECK assertions compiled in (always rebuilt, and the recorded hash is
    left stale on purpose so that the next plain build() rebuilds the normal library)."""
    checked = checked or bool(defines)
    if not force and not checked and up_to_date():
        return SO
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    objdir = os.path.join(HERE, "build")
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for src in sources():
        obj = os.path.join(objdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + ([CHECKED_FLAG] if checked and not defines else []) + ["-D" + d for d in defines] + (["-Xptxas", "-v"] if verbose else []) + ["-c", src, "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd)))
    for cmd, p in procs:
        if p.wait() != 0:
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
    subprocess.check_call([nvcc, "-shared", "-o", SO] + objs)
    with open(SO + ".hash", "w") as fh:
        fh.write(_digest() + ("-checked" if checked else ""))
    return SO


if __name__ == "__main__":
    _defs = [sys.argv[i + 1] for i, a in enumerate(sys.argv[:-1]) if a == "--define"]
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, checked="--checked" in sys.argv, defines=_defs))
