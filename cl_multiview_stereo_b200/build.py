"""Builds the native artefacts in-tree.

  libmvs_b200.so      CUDA kernels + C ABI (include/mvs_capi.h), sm_100a only
  libmvs_host.so      C++ host classes with the reference's API (include/*.h)
  oracle/liboracle.so CPU oracle (TEST INFRASTRUCTURE, built by oracle/Makefile)
"""
import os
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "cl_multiview_stereo_b200")
CSRC = os.path.join(PKG, "csrc")
HOST = os.path.join(PKG, "host")
LIB = os.path.join(PKG, "libmvs_b200.so")
HOSTLIB = os.path.join(PKG, "libmvs_host.so")

NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    # parity: no implicit contraction; IEEE division and square root; no flush-to-zero
    "-fmad=false", "-prec-div=true", "-prec-sqrt=true", "-ftz=false",
    "-Xcompiler", "-fPIC",
]


def _nvcc():
    return shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"


def _newer(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources)


def build_native(force=False, verbose=False):
    cu = [os.path.join(CSRC, f) for f in ("slic.cu", "photo.cu", "refine.cu", "propagate.cu", "fuse.cu", "io.cu", "capi.cu")]
    deps = cu + [os.path.join(CSRC, f) for f in ("common.cuh", "mvs_math.cuh")] + [os.path.join(ROOT, "include", "mvs_capi.h")]
    if force or _newer(LIB, deps):
        cmd = [_nvcc()] + NVCC_FLAGS + ["-shared", "-o", LIB] + cu + ["-ldl"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    host_src = sorted(os.path.join(HOST, f) for f in os.listdir(HOST) if f.endswith(".cpp")) if os.path.isdir(HOST) else []
    if host_src:
        hdeps = host_src + [os.path.join(ROOT, "include", f) for f in os.listdir(os.path.join(ROOT, "include"))]
        if force or _newer(HOSTLIB, hdeps + [LIB]):
            cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-I", os.path.join(ROOT, "include"),
                   "-o", HOSTLIB] + host_src + ["-L", PKG, "-lmvs_b200", "-Wl,-rpath,$ORIGIN"]
            if verbose:
                print(" ".join(cmd))
            subprocess.check_call(cmd)
    # the reference-shaped entry point (examples/clMVDE.cpp ~ reference clMVDE.cpp:12-43)
    ex_src = os.path.join(ROOT, "examples", "clMVDE.cpp")
    ex_bin = os.path.join(ROOT, "examples", "clMVDE")
    if os.path.exists(ex_src) and os.path.exists(HOSTLIB) and (force or _newer(ex_bin, [ex_src, HOSTLIB, LIB])):
        cmd = ["/usr/bin/g++", "-O2", "-std=c++17", "-I", os.path.join(ROOT, "include"), "-o", ex_bin, ex_src,
               "-L", PKG, "-lmvs_host", "-lmvs_b200", "-Wl,-rpath,$ORIGIN/../cl_multiview_stereo_b200"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
    return LIB


def build_oracle(force=False):
    odir = os.path.join(ROOT, "oracle")
    if force:
        subprocess.check_call(["make", "-C", odir, "clean"])
    subprocess.check_call(["make", "-C", odir, "-s"])
    ref_mk = os.path.join(odir, "ref", "Makefile")
    if os.path.isdir("/root/reference/clMVDE") and os.path.exists(ref_mk):
        subprocess.check_call(["make", "-C", os.path.join(odir, "ref"), "-s"])
    return os.path.join(odir, "liboracle.so")


if __name__ == "__main__":
    build_native(force=True, verbose=True)
    build_oracle()
