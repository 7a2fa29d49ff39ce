#!/usr/bin/env python
"""Build the two in-tree native libraries.

  csrc/kernels/*.cu  -> tensorflow-distributed-learning_b200/_native/libtfdl_kernels.so   (nvcc, sm_100a only)
  csrc/runtime/*.cc  -> tensorflow-distributed-learning_b200/_native/libtfdl_runtime.so   (g++, host only)

nvcc cross-compiles for sm_100a without a GPU present, so this runs on the CPU build box; the
resulting .so files travel to the GPU box with the repo snapshot. Objects are cached under
build/ keyed by source mtime + flags, so rebuilds only recompile what changed.
"""
from __future__ import annotations

import concurrent.futures as cf
import hashlib
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "tensorflow-distributed-learning_b200")
OUT = os.path.join(PKG, "_native")
BUILD = os.path.join(ROOT, "build")

NVCC = os.environ.get("NVCC", shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc")
CXX = os.environ.get("CXX", "g++")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC",
    "-Xptxas", "-v",
    "--expt-relaxed-constexpr",
] + os.environ.get("TFDL_NVCC_EXTRA", "").split()
CXX_FLAGS = ["-O2", "-std=c++17", "-fPIC", "-Wall", "-pthread"]


def _sig(path: str, flags: list[str], deps: list[str]) -> str:
    h = hashlib.sha1()
    h.update(" ".join(flags).encode())
    for p in [path] + deps:
        st = os.stat(p)
        h.update(f"{p}:{st.st_mtime_ns}:{st.st_size}".encode())
    return h.hexdigest()[:16]


def _compile(cmd: list[str], obj: str, log: str) -> tuple[str, int, str]:
    r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    with open(log, "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout)
    return obj, r.returncode, r.stdout


def build_lib(srcdir: str, exts: tuple[str, ...], compiler: str, flags: list[str], link: list[str],
              libname: str, verbose: bool) -> str:
    os.makedirs(BUILD, exist_ok=True)
    os.makedirs(OUT, exist_ok=True)
    srcs = sorted(os.path.join(srcdir, f) for f in os.listdir(srcdir) if f.endswith(exts))
    hdrs = sorted(os.path.join(srcdir, f) for f in os.listdir(srcdir) if f.endswith((".cuh", ".h", ".hpp")))
    hdrs += sorted(os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include")))
    jobs, objs = [], []
    for s in srcs:
        sig = _sig(s, flags, hdrs)
        obj = os.path.join(BUILD, f"{os.path.basename(s)}.{sig}.o")
        objs.append(obj)
        if not os.path.exists(obj):
            jobs.append(([compiler, *flags, "-I", os.path.join(ROOT, "include"), "-c", s, "-o", obj], obj,
                         os.path.join(BUILD, os.path.basename(s) + ".log")))
    if jobs:
        with cf.ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for obj, rc, out in ex.map(lambda j: _compile(*j), jobs):
                if rc != 0:
                    sys.stderr.write(out)
                    raise RuntimeError(f"compile failed: {obj}")
                if verbose:
                    sys.stdout.write(out)
    lib = os.path.join(OUT, libname)
    newest = max(os.stat(o).st_mtime_ns for o in objs)
    if not os.path.exists(lib) or os.stat(lib).st_mtime_ns < newest or jobs:
        cmd = [compiler, "-shared", "-o", lib, *objs, *link]
        r = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout)
            raise RuntimeError(f"link failed: {lib}")
    return lib


def build_kernels(verbose: bool = False) -> str:
    return build_lib(os.path.join(ROOT, "csrc", "kernels"), (".cu",), NVCC, NVCC_FLAGS,
                     ["-gencode", "arch=compute_100a,code=sm_100a", "-Xcompiler", "-fPIC", "-lcudart"],
                     "libtfdl_kernels.so", verbose)


def build_runtime(verbose: bool = False) -> str:
    return build_lib(os.path.join(ROOT, "csrc", "runtime"), (".cc",), CXX, CXX_FLAGS, ["-pthread"],
                     "libtfdl_runtime.so", verbose)


def main() -> None:
    verbose = "-v" in sys.argv
    which = [a for a in sys.argv[1:] if not a.startswith("-")] or ["kernels", "runtime"]
    if "runtime" in which and os.path.isdir(os.path.join(ROOT, "csrc", "runtime")) and any(
            f.endswith(".cc") for f in os.listdir(os.path.join(ROOT, "csrc", "runtime"))):
        print("built", build_runtime(verbose))
    if "kernels" in which:
        print("built", build_kernels(verbose))


if __name__ == "__main__":
    main()
