"""Builds libpimpt.so (the CUDA library behind include/pim_pt.h) in-tree for sm_100a."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpimpt.so")
SOURCES = ["api.cu", "bvh_build.cu", "scene.cu", "wavefront.cu", "pathkernels.cu", "bakes.cu", "lmpack.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-std=c++17", "-O3", "-lineinfo",
    # The reference is built -O3 -ffast-math with FTZ/DAZ (CMakeLists.txt:168, src/threading/task.c:73-74).
    # --use_fast_math (ftz, approximate div/sqrt, __sinf/__cosf/__logf/__expf) keeps >= 99 % of paths
    # within 1e-3 of the reference path for path and is 14 % faster in the shading kernels; the
    # triangle test, flatten and light-grid ray setup use explicit IEEE intrinsics (bvh.h divr/sqrtr).
    "--use_fast_math",
    "-Xcompiler", "-fPIC", "-Xptxas", "-v",
]


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_library(force=False, verbose=False, defines=(), lib=None, tag=""):
    """Compile csrc/*.cu and link libpimpt.so. `defines`/`lib`/`tag` build tuning variants
    (experiments only; the shipped library is the default build)."""
    nvcc = os.environ.get("NVCC", "nvcc")
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    headers.append(os.path.join(os.path.dirname(HERE), "include", "pim_pt.h"))
    objdir = os.path.join(HERE, "build" + tag)
    out_lib = lib or LIB
    os.makedirs(objdir, exist_ok=True)
    objs = []
    procs = []
    for src in SOURCES:
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            continue
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [path] + headers):
            cmd = [nvcc] + NVCC_FLAGS + ["-D" + d for d in defines] + ["-c", path, "-o", obj]
            log = open(obj + ".log", "w")
            procs.append((src, subprocess.Popen(cmd, stdout=log, stderr=subprocess.STDOUT), log, obj))
    failed = False
    for src, p, log, obj in procs:
        rc = p.wait()
        log.close()
        if rc != 0:
            failed = True
            sys.stderr.write(open(obj + ".log").read())
        elif verbose:
            sys.stdout.write(open(obj + ".log").read())
    if failed:
        raise RuntimeError("nvcc failed")
    if force or procs or _stale(out_lib, objs):
        cmd = [nvcc, "-shared", "-o", out_lib] + objs + ["-lcudart"]
        subprocess.check_call(cmd)
    return out_lib


if __name__ == "__main__":
    print(build_library(force="--force" in sys.argv, verbose="-v" in sys.argv))
