"""Build libexabrick_b200.so in-tree with nvcc for sm_100a (cross-compiles without a GPU).

    python -m owlexabrick_b200.build
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libexabrick_b200.so")
OBJ = os.path.join(HERE, "build")

NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC,-fopenmp", "-DEXA_BUILD"]
if os.environ.get("EXA_DEBUG"):
    COMMON = COMMON + ["-DEXA_DEBUG"]
if os.environ.get("EXA_VARIANT"):
    # kernel-variant build for A/B timing: EXA_VARIANT=name EXA_FLAGS="-DX=1 ..." -> libexabrick_b200.<name>.so
    COMMON = COMMON + os.environ.get("EXA_FLAGS", "").split()
    LIB = os.path.join(HERE, "libexabrick_b200.%s.so" % os.environ["EXA_VARIANT"])
    OBJ = os.path.join(HERE, "build", os.environ["EXA_VARIANT"])
SOURCES = ["host.cpp", "colormap.cpp", "builder.cpp", "gpu_builder.cu", "abr.cu", "render.cu", "capi.cu"]
HEADERS = ["exa_internal.h", "builder_common.h", "render.cuh", "surfaces.cuh", "capi_render.inl", "capi_group.inl", "colormaps_data.inc", os.path.join("..", "..", "include", "exabrick_b200.h")]


def _stale(out, deps):
    if not os.path.exists(out):
        return True
    t = os.path.getmtime(out)
    return any(os.path.getmtime(d) > t for d in deps)


def build_lib(force=False, verbose=False):
    """Compile what is stale and link.  Safe to call from several processes at once (the ranks of a torchrun job all
    import the package): an exclusive file lock serialises the builders, and every output is written under a temporary
    name and renamed into place, so a process that loads the library meanwhile sees a complete file."""
    import fcntl
    os.makedirs(OBJ, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]

    def run(cmd, out):
        tmp = "%s.tmp%d" % (out, os.getpid())
        cmd = [tmp if a == out else a for a in cmd]
        if verbose:
            print(" ".join(cmd))
        try:
            subprocess.check_call(cmd)
            os.replace(tmp, out)
        finally:
            if os.path.exists(tmp):
                os.unlink(tmp)

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        if force or _stale(o, [s] + hdrs):
            cmd = [NVCC] + ARCH + COMMON + ["-c", s, "-o", o]
            if src.endswith(".cpp"):
                cmd = [NVCC] + COMMON + ["-x", "cu", "-c", s, "-o", o] + ARCH
            run(cmd, o)
        return o

    with open(os.path.join(OBJ, ".lock"), "w") as lock:
        fcntl.flock(lock, fcntl.LOCK_EX)
        try:
            with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
                objs = list(ex.map(compile_one, SOURCES))
            if force or _stale(LIB, objs):
                run([NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-Xcompiler", "-fopenmp", "-lgomp"], LIB)
        finally:
            fcntl.flock(lock, fcntl.LOCK_UN)
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose=True))
