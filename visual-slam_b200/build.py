"""Builds visual-slam_b200/csrc/libvisnav_b200.so (the C-ABI library) with nvcc for sm_100a.

In-tree build: the .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import os
import shutil
import subprocess

_HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(_HERE, "csrc")
LIB = os.path.join(CSRC, "libvisnav_b200.so")
SOURCES = ["api.cu", "detect.cu", "describe.cu", "match.cu", "match_mma.cu", "landmarks.cu", "pyramid.cu", "geometry.cu", "bow.cu", "orb.cu", "noalloc.cu"]
HEADERS = ["common.cuh", "rbrief_pattern.inc", os.path.join("..", "..", "include", "visnav_b200.h")]
# geometry.cu spells its arithmetic with plain * and +: without contraction it evaluates every expression exactly as the
# oracle (gcc -ffp-contract=off) does, so inlier / visibility decisions do not depend on where nvcc would fuse.  The other
# files pin their rounding with explicit __f*_rn / __d*_rn / __fma_rn intrinsics and are unaffected by the flag.
PER_FILE_FLAGS = {"geometry.cu": ["-fmad=false"]}
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
              "-Xcompiler", "-fPIC"]


def _nvcc():
    nvcc = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(nvcc):
        raise RuntimeError("nvcc not found: the CUDA extension cannot be built")
    return nvcc


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def build_native(force=False, verbose=False):
    """Compile every .cu to an object and link the shared library.  Returns the .so path."""
    nvcc = _nvcc()
    hdrs = [os.path.join(CSRC, h) for h in HEADERS]
    objs = []
    procs = []
    for src in SOURCES:
        s = os.path.join(CSRC, src)
        o = os.path.join(CSRC, src[:-3] + ".o")
        objs.append(o)
        if force or _stale(o, [s] + hdrs):
            extra = os.environ.get("VSLAM_NVCC_EXTRA", "").split()   # e.g. -DVSLAM_SEL_PROFILE for latency work
            cmd = [nvcc] + NVCC_FLAGS + PER_FILE_FLAGS.get(src, []) + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", s, "-o", o]
            procs.append((src, subprocess.Popen(cmd, cwd=CSRC, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for src, p in procs:
        out, _ = p.communicate()
        if verbose or p.returncode:
            print(out.decode(errors="replace"))
        if p.returncode:
            raise RuntimeError("nvcc failed on %s" % src)
    if force or procs or _stale(LIB, objs):
        cmd = [nvcc, "-shared", "-gencode", "arch=compute_100a,code=sm_100a", "-o", LIB] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"]
        subprocess.run(cmd, cwd=CSRC, check=True)
    return LIB


if __name__ == "__main__":
    import sys
    print(build_native(force="--force" in sys.argv, verbose="-v" in sys.argv))
