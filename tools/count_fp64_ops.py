"""Static FP64 operation count of the arm node functions from the SASS of the built library.

SURVEY §8(d): the reference gives no closed form for the ABA / ABA-derivative work per node; the figure used for the
FP64 roofline of the arm kernels is the operation count of the implementation. The node functions are fully unrolled
straight-line code (NV is a compile-time constant), so the static count of DFMA (2 flops), DADD and DMUL (1 flop each)
in `arm_calc_node_dev<7>` / `arm_calcdiff_node_dev<7>` (both __noinline__) is the per-node count; backward branches
(loops the compiler kept) are reported so that a non-unrolled loop cannot hide work.

usage: python tools/count_fp64_ops.py [path/to/cta_arm7.o]
"""
import collections, os, re, subprocess, sys, tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    obj = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "crocoddyl_b200", "lib", "cta_arm7.o")
    with tempfile.TemporaryDirectory() as td:
        subprocess.check_call(["cuobjdump", "-xelf", "all", os.path.abspath(obj)], cwd=td, stdout=subprocess.DEVNULL)
        cubin = [os.path.join(td, f) for f in os.listdir(td) if f.endswith(".cubin")][0]
        sass = subprocess.check_output(["nvdisasm", "-c", cubin], text=True)
    fn = None
    kernel = None
    ops = collections.defaultdict(collections.Counter)
    for line in sass.splitlines():
        m = re.match(r"\s*\.text\.(\S+):", line)
        if m:
            kernel, fn = m.group(1), "<kernel body>"
            continue
        m = re.match(r"\s*(\$\S+):", line)
        if m and kernel and "$" in m.group(1)[1:]:
            fn = m.group(1).split("$")[-1]
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if m and kernel:
            op = m.group(1).split(".")[0]
            ops[(kernel, fn)][op] += 1
            if op == "BRA" and re.search(r"`\(\.L_x_\d+\)", line):
                ops[(kernel, fn)]["_bra"] += 1
    for (k, f), c in ops.items():
        if "cta_problem_kernel" in k and "ArmModelILi7" in k:
            flops = 2 * c["DFMA"] + c["DADD"] + c["DMUL"]
            print("%-40s DFMA %6d DADD %5d DMUL %5d MUFU %3d -> %7d flops; instructions %6d, local ld/st %d/%d, branches %d"
                  % (f[:40], c["DFMA"], c["DADD"], c["DMUL"], c["MUFU"], flops, sum(v for kk, v in c.items() if kk != "_bra"),
                     c["LDL"], c["STL"], c["_bra"]))


if __name__ == "__main__":
    main()
