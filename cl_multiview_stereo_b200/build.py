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


def _source_hash(paths):
    """sha256 over the named sources (name + contents): recorded beside the library and checked when it is loaded"""
    import hashlib
    h = hashlib.sha256()
    for q in sorted(paths):
        h.update(os.path.basename(q).encode() + b"\0")
        with open(q, "rb") as fh:
            h.update(fh.read())
        h.update(b"\0")
    return h.hexdigest()


def native_sources():
    cu = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cu")]
    deps = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC)) if f.endswith(".cuh")]
    deps.append(os.path.join(ROOT, "include", "mvs_capi.h"))
    return cu, deps


def build_native(force=False, verbose=False):
    """every .cu -> build/<name>.o (in parallel, only what changed), linked into libmvs_b200.so; the hash of the
    sources the library was built from is written to libmvs_b200.so.srchash (_native.lib() refuses a stale one).
    MVS_NVCC_DEFS adds -D switches (kernel experiments on the GPU box)."""
    cu, deps = native_sources()
    extra = os.environ.get("MVS_NVCC_DEFS", "").split()
    want = _source_hash(cu + deps) + (" " + " ".join(extra) if extra else "")
    stamp = LIB + ".srchash"
    have = open(stamp).read().strip() if os.path.exists(stamp) else ""
    if force or have != want or not os.path.exists(LIB):
        odir = os.path.join(ROOT, "build")
        os.makedirs(odir, exist_ok=True)
        procs = []
        objs = []
        for src in cu:
            obj = os.path.join(odir, os.path.basename(src)[:-3] + ".o")
            objs.append(obj)
            tag = obj + ".srchash"
            w = _source_hash([src] + deps) + " " + " ".join(extra)
            if not force and os.path.exists(obj) and os.path.exists(tag) and open(tag).read() == w:
                continue
            cmd = [_nvcc()] + NVCC_FLAGS + extra + ["-c", "-o", obj, src]
            if verbose:
                print(" ".join(cmd))
            procs.append((subprocess.Popen(cmd), tag, w, cmd))
        for pr, tag, w, cmd in procs:
            if pr.wait() != 0:
                raise subprocess.CalledProcessError(pr.returncode, cmd)
            with open(tag, "w") as fh:
                fh.write(w)
        cmd = [_nvcc(), "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", LIB] + objs + ["-ldl"]
        if verbose:
            print(" ".join(cmd))
        subprocess.check_call(cmd)
        with open(stamp, "w") as fh:
            fh.write(want + "\n")
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
        # the reference's HOST sources (CPU twins) over inert OpenCL / OpenCV headers: a timed baseline, not an oracle
        subprocess.check_call(["make", "-C", os.path.join(odir, "ref_host"), "-s"])
    return os.path.join(odir, "liboracle.so")


if __name__ == "__main__":
    build_native(force=True, verbose=True)
    build_oracle()
