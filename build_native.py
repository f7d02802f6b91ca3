#!/usr/bin/env python
"""Build the in-tree native libraries (no pip, no JIT cache):

  batchai_retinanet_horovod_coco_b200/_native/libb200ops.so   sm_100a CUDA kernels (csrc/*.cu)
  batchai_retinanet_horovod_coco_b200/_native/libb200rt.so    C++ host runtime    (csrc/rt/*.cpp)

nvcc cross-compiles for sm_100a without a GPU.  Objects are cached by source mtime.
"""
from __future__ import annotations

import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

ROOT = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(ROOT, "csrc")
OUT = os.path.join(ROOT, "batchai_retinanet_horovod_coco_b200", "_native")
OBJ = os.path.join(ROOT, "build", "obj")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC", "--use_fast_math", "-Xptxas", "-v", "-I", os.path.join(ROOT, "include")]
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-pthread", "-I", os.path.join(ROOT, "include")]


def _stale(src: str, obj: str, deps) -> bool:
    if not os.path.exists(obj):
        return True
    t = os.path.getmtime(obj)
    return any(os.path.getmtime(d) > t for d in [src] + deps)


def _run(cmd, log):
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        sys.stderr.write(r.stdout + r.stderr)
        raise RuntimeError("build failed: " + " ".join(cmd))
    return r.stdout + r.stderr


def build(verbose: bool = False) -> dict:
    os.makedirs(OUT, exist_ok=True)
    os.makedirs(OBJ, exist_ok=True)
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    headers += [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
    cu = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    jobs = []
    objs = []
    trace = os.environ.get("B200_CONV_TRACE", "0") == "1"      # in-kernel timeline of the conv kernel (scripts/conv_timeline.py)
    for f in cu:
        src = os.path.join(CSRC, f)
        traced = trace and f == "conv_sm100.cu"
        obj = os.path.join(OBJ, f[:-3] + (".trace.o" if traced else ".o"))
        objs.append(obj)
        if _stale(src, obj, headers):
            jobs.append(([NVCC] + NVCC_FLAGS + (["-DB200_CONV_TRACE"] if traced else []) + ["-c", src, "-o", obj], obj + ".log"))
    rt_dir = os.path.join(CSRC, "rt")
    rt_objs = []
    if os.path.isdir(rt_dir):
        rt_headers = headers + [os.path.join(rt_dir, f) for f in os.listdir(rt_dir) if f.endswith((".h", ".hpp", ".inc"))]
        for f in sorted(os.listdir(rt_dir)):
            if f.endswith(".cpp"):
                src = os.path.join(rt_dir, f)
                obj = os.path.join(OBJ, "rt_" + f[:-4] + ".o")
                rt_objs.append(obj)
                if _stale(src, obj, rt_headers):
                    jobs.append((["g++"] + CXX_FLAGS + ["-c", src, "-o", obj], obj + ".log"))
    with ThreadPoolExecutor(max_workers=8) as ex:
        outs = list(ex.map(lambda j: _run(*j), jobs))
    if verbose:
        for o in outs:
            print(o)
    res = {}
    if objs:
        lib = os.path.join(OUT, "libb200ops.so")
        if jobs or not os.path.exists(lib) or any(os.path.getmtime(o) > os.path.getmtime(lib) for o in objs):
            _run([NVCC, "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a", "-lcudart_static",
                                                        "-lpthread", "-ldl", "-lrt"], os.path.join(OBJ, "link_ops.log"))
        res["ops"] = lib
    if rt_objs:
        lib = os.path.join(OUT, "libb200rt.so")
        if jobs or not os.path.exists(lib) or any(os.path.getmtime(o) > os.path.getmtime(lib) for o in rt_objs):
            _run(["g++", "-shared", "-o", lib] + rt_objs + ["-pthread", "-lrt"], os.path.join(OBJ, "link_rt.log"))
        res["rt"] = lib
    return res


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv))
