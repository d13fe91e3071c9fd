"""Build libmatas_b200.so in-tree with nvcc for sm_100a (no torch extension machinery: the
library is a plain C-ABI shared object, see include/matas_b200.h).

    python -m mat_dcml_b200.build [--force]
"""
import concurrent.futures
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "lib", "libmatas_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC", "-Xptxas", "-v"]
# env.cu / gae.cu restate float arithmetic whose branches / bits depend on unfused rounding
EXTRA_FLAGS = {"env.cu": ["-fmad=false", "-maxrregcount=64"], "gae.cu": ["-fmad=false"]}


def _sources():
    return {s: EXTRA_FLAGS.get(s, []) for s in sorted(os.listdir(CSRC)) if s.endswith(".cu")}


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps)


def _headers():
    hs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".h", ".cuh"))]
    hs.append(os.path.join(os.path.dirname(HERE), "include", "matas_b200.h"))
    return hs


def _compile(src, flags):
    obj = os.path.join(OBJ, src.replace(".cu", ".o"))
    cmd = [NVCC] + ARCH + COMMON + flags + ["-c", os.path.join(CSRC, src), "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    with open(obj + ".log", "w") as f:
        f.write(" ".join(cmd) + "\n" + r.stdout + r.stderr)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s" % (src, r.stdout + r.stderr))
    return obj


def build(force=False, verbose=False):
    os.makedirs(OBJ, exist_ok=True)
    os.makedirs(os.path.dirname(LIB), exist_ok=True)
    srcs = _sources()
    hdrs = _headers()
    todo = {s: f for s, f in srcs.items()
            if force or _stale(os.path.join(OBJ, s.replace(".cu", ".o")), [os.path.join(CSRC, s), __file__] + hdrs)}
    if todo:
        with concurrent.futures.ThreadPoolExecutor(max_workers=8) as ex:
            list(ex.map(lambda kv: _compile(*kv), todo.items()))
    objs = [os.path.join(OBJ, s.replace(".cu", ".o")) for s in srcs]
    if force or todo or _stale(LIB, objs):
        cmd = [NVCC] + ARCH + ["-shared", "-o", LIB] + objs
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n" + r.stdout + r.stderr)
        if verbose:
            print("built", LIB)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose=True))
