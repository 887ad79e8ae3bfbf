"""Builds the sm_100a shared library (crocoddyl_b200/lib/libcrocoddyl_b200.so) in-tree with nvcc.

nvcc cross-compiles without a GPU, so this also runs on the CPU-only driver box. The .so is
git-ignored but travels to the GPU box with the gpurun snapshot.
"""
import hashlib
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIBDIR = os.path.join(HERE, "lib")
LIB = os.path.join(LIBDIR, "libcrocoddyl_b200.so")
# per-source extra flags (none: the register budget of the thread-per-problem solve kernel — 128, so that 65,536 problems
# are resident in ONE wave — is set by its __launch_bounds__, and its tail instantiation deliberately has none)
EXTRA = {}
SOURCES = ["cta_arm7.cu", "cta_arm4.cu", "cta_arm2.cu", "cta_backend.cu", "tpp_backend.cu", "capi.cu", "noalloc.cu"]
NVCC_FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
              "-Xcompiler", "-fvisibility=hidden"]


def _digest():
    h = hashlib.sha256()
    h.update(" ".join(NVCC_FLAGS).encode())
    root = os.path.dirname(HERE)
    files = [os.path.join(CSRC, f) for f in sorted(os.listdir(CSRC))] + [os.path.join(root, "include", "crocoddyl_b200.h")]
    for f in files:
        with open(f, "rb") as fh:
            h.update(os.path.basename(f).encode())  # not the full path: the tree may be checked out anywhere
            h.update(fh.read())
    return h.hexdigest()


def build(force=False, quiet=False, variant=None, extra=None):
    """Compile every CUDA source for sm_100a and link the C-ABI library. Returns the library path.
    variant/extra: development knob — builds lib/libcrocoddyl_b200_<variant>.so with other per-source flags
    (select it at run time with CROCODDYL_B200_LIB)."""
    os.makedirs(LIBDIR, exist_ok=True)
    LIB = os.path.join(LIBDIR, "libcrocoddyl_b200%s.so" % ("_" + variant if variant else ""))
    EXTRA = dict(globals()["EXTRA"]) if extra is None else extra
    stamp = os.path.join(LIBDIR, "build%s.stamp" % ("_" + variant if variant else ""))
    dig = _digest() + repr(sorted(EXTRA.items()))
    if not force and os.path.exists(LIB) and os.path.exists(stamp) and open(stamp).read() == dig:
        return LIB
    nvcc = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
    if not os.path.exists(nvcc):
        nvcc = "nvcc"
    objs = []
    procs = []
    for src in SOURCES:
        obj = os.path.join(LIBDIR, src.replace(".cu", "%s.o" % ("_" + variant if variant else "")))
        objs.append(obj)
        cmd = [nvcc] + NVCC_FLAGS + EXTRA.get(src, []) + ["-c", os.path.join(CSRC, src), "-o", obj]
        if not quiet:
            print(" ".join(cmd), flush=True)
        procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            sys.stderr.write(out.decode())
            raise RuntimeError("nvcc failed: " + " ".join(cmd))
        if not quiet and out:
            sys.stdout.write(out.decode())
    cmd = [nvcc, "-shared", "-o", LIB] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"]
    if not quiet:
        print(" ".join(cmd), flush=True)
    subprocess.check_call(cmd)
    with open(stamp, "w") as fh:
        fh.write(dig)
    return LIB


PARITY_FLAGS = ["-fmad=false", "-DCGX_PARITY=1"]


def build_parity(force=False, quiet=False):
    """lib/libcrocoddyl_b200_parity.so: the same sources without FMA contraction (-fmad=false, as the oracle's
    -ffp-contract=off) and with IEEE divisions by the LLT pivots instead of reciprocal multiplies (-DCGX_PARITY).
    Used by the parity tests that demand the reference's exact iteration count / accept sequence on every problem
    (select it at run time with CROCODDYL_B200_LIB); the default build is the fast one."""
    extra = {src: globals()["EXTRA"].get(src, []) + PARITY_FLAGS for src in SOURCES}
    return build(force=force, quiet=quiet, variant="parity", extra=extra)


if __name__ == "__main__":
    print(build(force="--force" in sys.argv))
    if "--parity" in sys.argv:
        print(build_parity(force="--force" in sys.argv))
