"""Builds nerf_b200/libnerf_b200.so in-tree with nvcc for sm_100a (no torch linkage:
the library is a plain CUDA-runtime C ABI, see include/nerf_b200.h)."""
import glob
import hashlib
import os
import shutil
import subprocess
import sys

PKG = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(PKG, "csrc")
OUT = os.path.join(PKG, "libnerf_b200.so")
OBJ = os.path.join(PKG, "build")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
NVCC_FLAGS = ["-O3", "-lineinfo", "-std=c++17", "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr",
              "-Xptxas", "-v"]


def nvcc():
    for c in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if c and os.path.exists(c):
            return c
    raise RuntimeError("nvcc not found")


def _digest(path, deps):
    h = hashlib.sha1()
    for p in [path] + deps:
        h.update(open(p, "rb").read())
    h.update(" ".join(ARCH + NVCC_FLAGS).encode())
    return h.hexdigest()


def build(verbose=False, force=False):
    os.makedirs(OBJ, exist_ok=True)
    headers = sorted(glob.glob(os.path.join(CSRC, "*.cuh"))) + [os.path.join(os.path.dirname(PKG), "include", "nerf_b200.h")]
    objs, rebuilt = [], False
    for src in sorted(glob.glob(os.path.join(CSRC, "*.cu"))):
        obj = os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")
        stamp = obj + ".sha1"
        dg = _digest(src, headers)
        if force or not os.path.exists(obj) or not os.path.exists(stamp) or open(stamp).read() != dg:
            cmd = [nvcc()] + ARCH + NVCC_FLAGS + ["-c", src, "-o", obj]
            r = subprocess.run(cmd, capture_output=True, text=True)
            log = r.stdout + r.stderr
            open(obj + ".log", "w").write(log)
            if r.returncode != 0:
                sys.stderr.write(log)
                raise RuntimeError(f"nvcc failed on {src}")
            if verbose:
                print(log)
            open(stamp, "w").write(dg)
            rebuilt = True
        objs.append(obj)
    if rebuilt or not os.path.exists(OUT):
        cmd = [nvcc()] + ARCH + ["-shared", "-o", OUT] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            sys.stderr.write(r.stdout + r.stderr)
            raise RuntimeError("link failed")
    return OUT


if __name__ == "__main__":
    print(build(verbose="-v" in sys.argv, force="-f" in sys.argv))
