"""Build libkgcnn_b200.so (sm_100a) in-tree with nvcc.  Used by __graft_entry__.build() and by hand:

    python -m gcnn_keras_b200.build [--force] [--verbose]

The library has no torch / python dependency: it is the framework-neutral C-ABI of include/kgcnn_b200.h.
"""
import concurrent.futures
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "csrc", "build")
LIB = os.path.join(HERE, "libkgcnn_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v"]


def _sources():
    return sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest(path):
    h = hashlib.sha256()
    for f in [path] + sorted(os.path.join(CSRC, x) for x in os.listdir(CSRC) if x.endswith((".cuh", ".h"))) + \
            [os.path.join(HERE, "..", "include", "kgcnn_b200.h")]:
        with open(f, "rb") as fh:
            h.update(fh.read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def _compile(src, force, verbose):
    path = os.path.join(CSRC, src)
    obj = os.path.join(OBJ, src[:-3] + ".o")
    stamp = obj + ".sha"
    dig = _digest(path)
    if not force and os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj, ""
    cmd = [NVCC] + FLAGS + ["-c", path, "-o", obj]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, res.stdout, res.stderr))
    with open(stamp, "w") as fh:
        fh.write(dig)
    with open(obj + ".ptxas.log", "w") as fh:
        fh.write(res.stderr)
    return obj, res.stderr if verbose else ""


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    srcs = _sources()
    with concurrent.futures.ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        results = list(ex.map(lambda s: _compile(s, force, verbose), srcs))
    objs = [r[0] for r in results]
    for _, log in results:
        if log:
            print(log)
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB) or os.path.getmtime(LIB) < newest:
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
        res = subprocess.run(cmd, capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (res.stdout, res.stderr))
    return LIB


LIB_PRECISE = os.path.join(HERE, "libkgcnn_b200_precise.so")


def build_precise(force=False):
    """The same library with csrc/cfconv.cu compiled under -DKGCNN_CF_PRECISE (libm expf / log1pf / division in the
    fused cfconv's softplus chain instead of the fast intrinsics) -> libkgcnn_b200_precise.so.  Only that one source is
    recompiled; every other object is the regular build's.  Used by tests/test_gpu_cfconv.py so that the precise path
    keeps compiling and keeps its parity."""
    build(force=force)
    src = os.path.join(CSRC, "cfconv.cu")
    obj = os.path.join(OBJ, "cfconv_precise.o")
    stamp = obj + ".sha"
    dig = _digest(src) + "+precise"
    if force or not (os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig):
        res = subprocess.run([NVCC] + FLAGS + ["-DKGCNN_CF_PRECISE", "-c", src, "-o", obj], capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("nvcc failed for cfconv.cu (precise):\n%s\n%s" % (res.stdout, res.stderr))
        with open(stamp, "w") as fh:
            fh.write(dig)
    objs = [os.path.join(OBJ, s[:-3] + ".o") for s in _sources() if s != "cfconv.cu"] + [obj]
    newest = max(os.path.getmtime(o) for o in objs)
    if force or not os.path.exists(LIB_PRECISE) or os.path.getmtime(LIB_PRECISE) < newest:
        res = subprocess.run([NVCC, "-shared", "-o", LIB_PRECISE] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"],
                             capture_output=True, text=True)
        if res.returncode != 0:
            raise RuntimeError("link failed (precise):\n%s\n%s" % (res.stdout, res.stderr))
    return LIB_PRECISE


if __name__ == "__main__":
    if "--precise" in sys.argv:
        print(build_precise(force="--force" in sys.argv))
        sys.exit(0)
    print(build(force="--force" in sys.argv, verbose="--verbose" in sys.argv))
