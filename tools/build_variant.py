"""Development aid: build lib/libcrocoddyl_b200_<variant>.so from the default objects with SOME sources recompiled under
extra flags (select it at run time with CROCODDYL_B200_LIB).
  python tools/build_variant.py <variant> <src.cu>[,<src.cu>...] [-DFLAG ...]"""
import os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from crocoddyl_b200 import build as B

variant, srcs, flags = sys.argv[1], sys.argv[2].split(","), sys.argv[3:]
B.build(quiet=True)  # the default objects
objs = []
for src in B.SOURCES:
    obj = os.path.join(B.LIBDIR, src.replace(".cu", ".o"))
    if src in srcs:
        obj = os.path.join(B.LIBDIR, src.replace(".cu", "_%s.o" % variant))
        subprocess.check_call(["/usr/local/cuda/bin/nvcc"] + B.NVCC_FLAGS + B.EXTRA.get(src, []) + flags +
                              ["-c", os.path.join(B.CSRC, src), "-o", obj])
    objs.append(obj)
lib = os.path.join(B.LIBDIR, "libcrocoddyl_b200_%s.so" % variant)
subprocess.check_call(["/usr/local/cuda/bin/nvcc", "-shared", "-o", lib] + objs + ["-gencode", "arch=compute_100a,code=sm_100a"])
print(lib)
