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
    os.makedirs(OBJ, exist_ok=True)
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]

    def compile_one(src):
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src + ".o")
        if force or _stale(o, [s] + hdrs):
            cmd = [NVCC] + ARCH + COMMON + ["-c", s, "-o", o]
            if src.endswith(".cpp"):
                cmd = [NVCC] + COMMON + ["-x", "cu", "-c", s, "-o", o] + ARCH
            if verbose:
                print(" ".join(cmd))
            subprocess.check_call(cmd)
        return o

    with ThreadPoolExecutor(max_workers=len(SOURCES)) as ex:
        objs = list(ex.map(compile_one, SOURCES))
    if force or _stale(LIB, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs + ["-Xcompiler", "-fopenmp", "-lgomp"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return LIB


if __name__ == "__main__":
    print(build_lib(force="--force" in sys.argv, verbose=True))
