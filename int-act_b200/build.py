"""Builds libpi0b200.so (the C-ABI CUDA library of the Pi0 hot path) in-tree with nvcc for sm_100a.

Usage: python int-act_b200/build.py [--force]
The .so is git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libpi0b200.so")
STAMP = os.path.join(HERE, "build", "stamp.txt")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17",
    "-Xcompiler", "-fPIC", "--expt-relaxed-constexpr", "-Xptxas", "-v",
] + os.environ.get("PI0_EXTRA_NVCC_FLAGS", "").split()


def _sources():
    return sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith(".cu"))


def _digest():
    h = hashlib.sha256()
    inc = os.path.join(os.path.dirname(HERE), "include", "pi0b200.h")
    h.update(os.environ.get("PI0_EXTRA_NVCC_FLAGS", "").encode())
    for p in sorted(os.path.join(CSRC, f) for f in os.listdir(CSRC)) + [inc, __file__]:
        with open(p, "rb") as f:
            h.update(p.encode())
            h.update(f.read())
    return h.hexdigest()


def build(force=False, verbose=False, variant=""):
    """Compile every .cu under csrc/ to an object (in parallel) and link libpi0b200.so. `variant` (with
    PI0_EXTRA_NVCC_FLAGS) builds an instrumented copy libpi0b200_<variant>.so next to it (development aid, see
    tools/loop_timeline.py; selected at run time with PI0_LIB)."""
    global LIB, STAMP
    bdir = "build" + ("_" + variant if variant else "")
    if variant:
        LIB = os.path.join(HERE, f"libpi0b200_{variant}.so")
        STAMP = os.path.join(HERE, bdir, "stamp.txt")
    dig = _digest()
    if not force and os.path.exists(LIB) and os.path.exists(STAMP) and open(STAMP).read().strip() == dig:
        return LIB
    os.makedirs(os.path.join(HERE, bdir), exist_ok=True)
    procs, objs = [], []
    for src in _sources():
        obj = os.path.join(HERE, bdir, os.path.basename(src)[:-3] + ".o")
        objs.append(obj)
        cmd = [NVCC] + FLAGS + ["-c", src, "-o", obj]
        procs.append((src, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    log = []
    for src, p in procs:
        out, _ = p.communicate()
        log.append(f"==== {os.path.basename(src)} ====\n{out}")
        if p.returncode != 0:
            sys.stderr.write(out)
            raise RuntimeError(f"nvcc failed on {src}")
    with open(os.path.join(HERE, bdir, "ptxas.log"), "w") as f:
        f.write("\n".join(log))
    if verbose:
        print("\n".join(log))
    subprocess.check_call([NVCC, "-shared", "-o", LIB] + objs + ["-lcudart"])
    with open(STAMP, "w") as f:
        f.write(dig)
    return LIB


if __name__ == "__main__":
    _variant = [a.split("=", 1)[1] for a in sys.argv if a.startswith("--variant=")]
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv, variant=_variant[0] if _variant else ""))
