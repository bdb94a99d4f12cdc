"""Build librac_b200.so (sm_100a only) in-tree with nvcc.

Usage: python build.py [--force] [--verbose]
The shared library lands next to this file so it travels with the repo snapshot to the
GPU box; it is git-ignored.  There is no other backend and no CPU fallback.
"""
import hashlib
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "librac_b200.so")
STAMP = os.path.join(HERE, ".librac_b200.stamp")
SOURCES = ["rac_api.cu", "raster.cu", "pose_update.cu", "add_loss.cu", "aux_kernels.cu", "host_pack.cu", "pose_loss_step.cu"]
HEADERS = ["rac_common.cuh", "pose_update_dev.cuh", "add_loss_dev.cuh", "raster.cuh", "raster_setup.inl", "raster_shade.inl", "raster_tile.inl", "raster_finalize.inl",
           os.path.join("..", "..", "include", "rac.h")]

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-lineinfo", "-std=c++17",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden", "-Xcompiler", "-pthread",
    "--shared",
]


def _nvcc():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; librac_b200 cannot be built (there is no CPU fallback)")


def _digest():
    h = hashlib.sha256()
    for rel in SOURCES + HEADERS:
        with open(os.path.join(CSRC, rel), "rb") as f:
            h.update(f.read())
    h.update(" ".join(NVCC_FLAGS).encode())
    return h.hexdigest()


def build_variant(out_path, extra_flags, verbose=False):
    """Diagnostic builds (e.g. -DRAC_PHASE_TIMING for profiles/phase_timing.py); never the product library."""
    cmd = [_nvcc()] + NVCC_FLAGS + list(extra_flags) + (["-Xptxas", "-v"] if verbose else [])
    cmd += [os.path.join(CSRC, s) for s in SOURCES] + ["-o", out_path]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building " + out_path)
    return out_path


def build(force=False, verbose=False):
    digest = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as f:
            if f.read().strip() == digest:
                return LIB
    cmd = [_nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else [])
    cmd += [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if verbose or res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed building librac_b200.so")
    with open(STAMP, "w") as f:
        f.write(digest)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
