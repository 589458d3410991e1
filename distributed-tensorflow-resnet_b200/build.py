"""Build librn_b200.so in-tree with nvcc for sm_100a (no torch, no cmake).

    python distributed-tensorflow-resnet_b200/build.py [--force]

The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "librn_b200.so")
STAMP = os.path.join(HERE, "build", "sources.sha256")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-lineinfo", "-std=c++17",
         "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr"]


def sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def digest():
    h = hashlib.sha256()
    files = sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + \
        [os.path.join(os.path.dirname(HERE), "include", "rn_b200.h"), os.path.abspath(__file__)]
    for f in files:
        if os.path.isdir(f):
            continue
        h.update(os.path.basename(f).encode())     # names, not absolute paths: the GPU box unpacks the tree elsewhere
        h.update(open(f, "rb").read())
    return h.hexdigest()


def build(force=False, verbose=True):
    d = digest()
    if not force and os.path.exists(OUT) and os.path.exists(STAMP) and open(STAMP).read().strip() == d:
        return OUT
    if not os.path.exists(NVCC):
        if os.path.exists(OUT):
            return OUT   # GPU box without a toolchain change: use the shipped binary
        raise RuntimeError("nvcc not found at %s and no prebuilt %s" % (NVCC, OUT))
    os.makedirs(os.path.join(HERE, "build"), exist_ok=True)
    objs = []
    procs = []
    for src in sources():
        obj = os.path.join(HERE, "build", os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [NVCC] + FLAGS + ["-c", src, "-o", obj]
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out.decode())
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
        if verbose and out.strip():
            sys.stderr.write(out.decode())
    cmd = [NVCC, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-o", OUT] + objs + ["-lcudart_static", "-ldl", "-lrt", "-lpthread"]
    subprocess.check_call(cmd)
    open(STAMP, "w").write(d)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
