"""Opcode histogram of the kernels in the built objects (cuobjdump -xelf + nvdisasm): which memory / FP64 / tensor
instructions each kernel actually contains (LDGSTS = cp.async, DMMA = FP64 tensor-core MMA, LDG.E.128 ...).
  python tools/sass_histogram.py > profiles/r02_sass_opcodes.txt"""
import collections, os, re, subprocess, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ("DFMA", "DMUL", "DADD", "DMMA", "MUFU", "LDG", "STG", "LDGSTS", "LDS", "STS", "LDL", "STL", "SHFL", "BAR", "ATOMG", "RED",
        "LDC", "UBLKCP", "UTMALDG", "BRA", "CALL")
for obj in ("tpp_backend.o", "cta_backend.o", "cta_arm7.o", "noalloc.o"):
    with tempfile.TemporaryDirectory() as td:
        subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.join(ROOT, "crocoddyl_b200", "lib", obj)], cwd=td,
                              stdout=subprocess.DEVNULL)
        cubin = [os.path.join(td, f) for f in os.listdir(td) if f.endswith(".cubin")][0]
        sass = subprocess.check_output(["nvdisasm", "-c", cubin], text=True)
        res = subprocess.check_output(["cuobjdump", "-res-usage", cubin], text=True)
    regs = dict(re.findall(r"Function (\S+):\n\s+REG:(\d+)", res))
    fn, total, ops = None, collections.Counter(), collections.defaultdict(collections.Counter)
    for line in sass.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", line)
        if m:
            fn = m.group(1)
            continue
        m = re.match(r"\s+(?:/\*[0-9a-f]+\*/)?\s*(?:@!?U?P\d\s+)?([A-Z][A-Z0-9_.]+)", line)
        if m and fn:
            total[fn] += 1
            op = m.group(1)
            base = op.split(".")[0]
            if base in KEEP:
                key = op if base in ("LDG", "STG", "LDGSTS", "DMMA", "LDS", "STS") else base
                ops[fn][key] += 1
    print("== %s" % obj)
    for f, n in total.most_common():
        if n < 400:
            continue
        name = subprocess.check_output(["c++filt", f], text=True).strip().split("(")[0][:100]
        print("%s  [%d instructions, %s registers]" % (name, n, regs.get(f, "?")))
        print("    " + "  ".join("%s:%d" % kv for kv in sorted(ops[f].items(), key=lambda kv: -kv[1])))
