"""In-tree build of libmcpt.so (CUDA kernels for sm_100a + the C++ host shell) and the mcpt_cli binary.

nvcc cross-compiles without a GPU.  lbvh.cu is built with --fmad=false: the Morton kernel's float
pipeline must be IEEE-identical to the reference's host code (MTPC/morton code.cpp:23-32).
"""
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
HOST = os.path.join(HERE, "host")
TAG = os.environ.get("MCPT_BUILD_TAG", "")   # developer knob: A/B builds live beside the product as libmcpt_<tag>.so
OBJ = os.path.join(HERE, "build" + ("_" + TAG if TAG else ""))
LIB = os.path.join(HERE, "libmcpt%s.so" % ("_" + TAG if TAG else ""))
CLI = os.path.join(HERE, "mcpt_cli")
NVCC = os.environ.get("MCPT_NVCC", "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("MCPT_CXX", "/usr/bin/g++")  # not $CXX: the image exports a wrapper unusable for linking

ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ARCH + ["-lineinfo", "-O3", "-std=c++17", "-ccbin", CXX, "-Xcompiler", "-fPIC,-fvisibility=hidden"]
NVCC_FLAGS += os.environ.get("MCPT_NVCC_EXTRA", "").split()   # developer knob for A/B builds (e.g. -DMCPT_SHADE_MIN_CTAS=5)
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-fvisibility=hidden", "-Wall"]

CU_SOURCES = {"lbvh.cu": ["--fmad=false"], "ploc.cu": [], "wide.cu": [], "render.cu": [], "wavefront.cu": [], "png_device.cu": [], "obj_ingest.cu": [], "api.cu": [], "multigpu.cu": [], "microbench.cu": []}
CPP_SOURCES = ["scene_loader.cpp", "png_writer.cpp", "render_scene.cpp", "mcpt_error.cpp"]


def _newer(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _headers():
    out = [os.path.join(HERE, "..", "include", "mcpt.h"), os.path.abspath(__file__)]
    for d in (CSRC, HOST):
        out += [os.path.join(d, f) for f in os.listdir(d) if f.endswith((".h", ".cuh"))]
    return out


def _run(cmd):
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if r.returncode != 0:
        raise RuntimeError("build step failed: %s\n%s" % (" ".join(cmd), r.stdout))
    return r.stdout


def build(force=False, verbose=False):
    """Compile everything that is out of date; returns the path of libmcpt.so."""
    os.makedirs(OBJ, exist_ok=True)
    hdrs = _headers()
    jobs, objs = [], []
    for src, extra in CU_SOURCES.items():
        s, o = os.path.join(CSRC, src), os.path.join(OBJ, src + ".o")
        objs.append(o)
        if force or _newer(o, [s] + hdrs):
            jobs.append([NVCC] + NVCC_FLAGS + extra + ["-c", s, "-o", o])
    for src in CPP_SOURCES:
        s, o = os.path.join(HOST, src), os.path.join(OBJ, src + ".o")
        objs.append(o)
        if force or _newer(o, [s] + hdrs):
            jobs.append([CXX] + CXX_FLAGS + ["-c", s, "-o", o])
    with ThreadPoolExecutor(max_workers=8) as ex:
        for out in ex.map(_run, jobs):
            if verbose and out.strip():
                print(out)
    if force or jobs or not os.path.exists(LIB):
        _run([NVCC] + ARCH + ["-shared", "-ccbin", CXX, "-o", LIB] + objs + ["-cudart", "static", "-ldl", "-lpthread"])
    cli_src = os.path.join(HOST, "mcpt_cli.cpp")
    if not TAG and os.path.exists(cli_src) and (force or _newer(CLI, [cli_src, LIB])):
        _run([CXX] + CXX_FLAGS + [cli_src, "-o", CLI, "-L" + HERE, "-lmcpt", "-Wl,-rpath,$ORIGIN"])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
