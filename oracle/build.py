"""Builds the CPU oracle (oracle/raster_ref.c -> oracle/liboracle_raster.so) with gcc.

TEST INFRASTRUCTURE ONLY (see oracle/README.md).  `oracle/_ref/` — a build of the
reference's own sources — does not exist for this project: the reference is pure Python
(no C/C++ to compile) and its render path needs pyrender + OpenGL, which are unavailable.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "raster_ref.c")
LIB = os.path.join(HERE, "liboracle_raster.so")
STAMP = os.path.join(HERE, ".liboracle_raster.stamp")


def _cpu_has(flag):
    try:
        with open("/proc/cpuinfo") as f:
            for line in f:
                if line.startswith("flags"):
                    return flag in line.split()
    except OSError:
        pass
    return False


def _flags():
    # -ffp-contract=off: the spec says where an fma is used (explicit fmaf); nothing else may fuse.
    flags = ["-O2", "-std=gnu11", "-fPIC", "-shared", "-fopenmp", "-ffp-contract=off", "-fno-fast-math",
             "-fvisibility=hidden"]
    if _cpu_has("fma") and _cpu_has("avx2"):
        flags += ["-mavx2", "-mfma"]   # fmaf() becomes one instruction; results are identical
    return flags


def build(force=False):
    flags = _flags()
    with open(SRC, "rb") as f:
        digest = hashlib.sha256(f.read() + " ".join(flags).encode()).hexdigest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP):
        with open(STAMP) as f:
            if f.read().strip() == digest:
                return LIB
    cmd = ["gcc"] + flags + [SRC, "-o", LIB, "-lm"]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        sys.stderr.write(res.stdout + res.stderr)
        raise RuntimeError("gcc failed building the raster oracle")
    with open(STAMP, "w") as f:
        f.write(digest)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
