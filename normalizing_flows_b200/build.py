"""Build libflowcuda.so in-tree with nvcc for sm_100a (B200).

    python -m normalizing_flows_b200.build [--force]

The library is a plain C-ABI shared object (include/flowcuda.h); no torch headers are
involved, so the build is a handful of `nvcc -dc`-free compiles and one link, a few seconds
per file.  Objects are cached under normalizing_flows_b200/csrc/_obj keyed by mtime.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(PKG, "libflowcuda.so")
# diagnostics of the tcgen05 primitives (include/flowcuda_selftest.h): a separate TEST library, not part of the product .so
SELFTEST_LIB = os.path.join(PKG, "libflowcuda_selftest.so")
SELFTEST_SOURCES = ("tc_selftest.cu",)
INCLUDE = os.path.join(os.path.dirname(PKG), "include")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC",
    "--expt-relaxed-constexpr", "--extended-lambda",
    "-I", INCLUDE,
]


def _nvcc() -> str:
    for cand in (shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found: libflowcuda cannot be built")


def sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _deps_mtime() -> float:
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(INCLUDE, "flowcuda.h"))
    hdrs.append(os.path.join(INCLUDE, "flowcuda_selftest.h"))
    hdrs.append(os.path.abspath(__file__))
    return max(os.path.getmtime(h) for h in hdrs)


def _xla_ffi_include():
    """Directory holding xla/ffi/api/ffi.h, or None (the usual case here: JAX is not installable on either box)."""
    cand = [os.environ.get("XLA_FFI_INCLUDE")]
    try:
        import importlib

        cand.append(importlib.import_module("jax.ffi").include_dir())
    except Exception:
        pass
    for c in cand:
        if c and os.path.exists(os.path.join(c, "xla", "ffi", "api", "ffi.h")):
            return c
    return None


def build_xla_ffi_adapter(verbose: bool = False):
    """integration/flowcuda_xla_ffi.cc -> integration/libflowcuda_xla_ffi.so when the XLA FFI headers exist (UNVERIFIED
    path: never exercised on this project's machines); returns the library path or None."""
    inc = _xla_ffi_include()
    if inc is None:
        return None
    root = os.path.dirname(PKG)
    src = os.path.join(root, "integration", "flowcuda_xla_ffi.cc")
    out = os.path.join(root, "integration", "libflowcuda_xla_ffi.so")
    cuda_inc = os.path.join(os.path.dirname(os.path.dirname(_nvcc())), "include")
    cmd = ["g++", "-std=c++17", "-O2", "-shared", "-fPIC", "-I", inc, "-I", INCLUDE, "-I", cuda_inc, src, "-L", PKG, "-lflowcuda",
           "-Wl,-rpath," + PKG, "-o", out]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("XLA FFI adapter failed to build: %s\n%s" % (" ".join(cmd), r.stderr))
    return out


def build(force: bool = False, verbose: bool = False) -> str:
    os.makedirs(OBJ, exist_ok=True)
    nvcc = _nvcc()
    flags = list(NVCC_FLAGS)
    if os.environ.get("FLOWCUDA_BUILD_TRACE") == "1":  # diagnostics build (tools/dt_trace.py); never the shipped library
        flags.append("-DFLOWCUDA_TRACE")
        force = True
    # objects built with other flags (a diagnostics build left behind) must not be linked into this library
    stamp = os.path.join(OBJ, ".flags")
    want = " ".join(flags)
    if not os.path.exists(stamp) or open(stamp).read() != want:
        force = True
    hdr_m = _deps_mtime()
    jobs = []
    objs = []
    for src in sources():
        s = os.path.join(CSRC, src)
        o = os.path.join(OBJ, src[:-3] + ".o")
        objs.append(o)
        if force or not os.path.exists(o) or os.path.getmtime(o) < max(os.path.getmtime(s), hdr_m):
            cmd = [nvcc, *flags, "-c", s, "-o", o]
            if verbose:
                cmd.insert(1, "-Xptxas=-v")
            jobs.append(cmd)

    def run(cmd):
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("nvcc failed: %s\n%s\n%s" % (" ".join(cmd), r.stdout, r.stderr))
        return r.stderr

    if jobs:
        with ThreadPoolExecutor(max_workers=min(8, len(jobs))) as ex:
            for out in ex.map(run, jobs):
                if verbose and out:
                    print(out, file=sys.stderr)
    with open(stamp, "w") as f:
        f.write(want)
    test_objs = [o for o in objs if os.path.basename(o)[:-2] + ".cu" in SELFTEST_SOURCES]
    prod_objs = [o for o in objs if o not in test_objs]
    core_obj = os.path.join(OBJ, "core.o")
    for lib, members in ((LIB, prod_objs), (SELFTEST_LIB, test_objs + [core_obj])):
        need_link = bool(jobs) or not os.path.exists(lib) or any(os.path.getmtime(o) > os.path.getmtime(lib) for o in members)
        if need_link:
            run([nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", lib, *members])
    return LIB


if __name__ == "__main__":
    path = build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(path)
    adapter = build_xla_ffi_adapter()
    if adapter:
        print(adapter)
