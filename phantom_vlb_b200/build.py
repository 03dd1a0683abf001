"""In-tree build of libvlb_b200.so (sm_100a only) with nvcc.  `python -m phantom_vlb_b200.build`.

The shared library is the product's only compute path: there is no JIT, no Triton and no CPU fallback.
Objects are compiled in parallel (one nvcc per .cu) and linked into phantom_vlb_b200/libvlb_b200.so, which
is git-ignored but travels to the GPU box with the repo snapshot.
"""
import hashlib
import os
import subprocess
import sys
from concurrent.futures import ThreadPoolExecutor

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(HERE, "build")
LIB = os.path.join(HERE, "libvlb_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
         "--use_fast_math", "-Xcompiler", "-fPIC", "-I", os.path.join(os.path.dirname(HERE), "include")]


def _digest(path):
    h = hashlib.sha1()
    for p in [path] + sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))):
        with open(p, "rb") as f:
            h.update(f.read())
    inc = os.path.join(os.path.dirname(HERE), "include")
    if os.path.isdir(inc):
        for f in sorted(os.listdir(inc)):
            with open(os.path.join(inc, f), "rb") as fh:
                h.update(fh.read())
    h.update(" ".join(FLAGS).encode())
    return h.hexdigest()


def _compile(src):
    name = os.path.splitext(os.path.basename(src))[0]
    obj = os.path.join(OBJ, name + ".o")
    stamp = obj + ".sha1"
    dig = _digest(src)
    if os.path.exists(obj) and os.path.exists(stamp) and open(stamp).read() == dig:
        return obj, False
    cmd = [NVCC] + FLAGS + ["-c", src, "-o", obj]
    r = subprocess.run(cmd, capture_output=True, text=True)
    if r.returncode != 0:
        raise RuntimeError("nvcc failed for %s:\n%s\n%s" % (src, r.stdout, r.stderr))
    with open(stamp, "w") as f:
        f.write(dig)
    return obj, True


def build_lib(verbose=True):
    os.makedirs(OBJ, exist_ok=True)
    srcs = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))
    with ThreadPoolExecutor(max_workers=min(8, len(srcs))) as ex:
        res = list(ex.map(_compile, srcs))
    objs = [o for o, _ in res]
    changed = any(c for _, c in res) or not os.path.exists(LIB)
    if changed:
        cmd = [NVCC, "-shared", "-o", LIB] + objs + ["-lcuda", "-gencode", "arch=compute_100a,code=sm_100a"]
        r = subprocess.run(cmd, capture_output=True, text=True)
        if r.returncode != 0:
            raise RuntimeError("link failed:\n%s\n%s" % (r.stdout, r.stderr))
    if verbose:
        print("libvlb_b200: %d sources, %s -> %s" % (len(srcs), "rebuilt" if changed else "up to date", LIB))
    return LIB


if __name__ == "__main__":
    build_lib()
    sys.exit(0)
