"""Static evidence for the kernel-shape claims of DESIGN.md section 3, taken from the BUILT
library without a GPU: per kernel, registers / spills / shared memory from `ptxas -v`
(mx-deepim_b200/build/ptxas.log) and counts of the SASS mnemonics that show how memory is
moved -- 128-bit global loads / stores, cp.async staging (LDGSTS), shared-memory atomics, warp
shuffles / match / vote, distributed-shared-memory and cluster barriers, FP64 instructions.

    python scripts/sass_evidence.py > profiles/sass_evidence.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "mx-deepim_b200", "librendercompare.so")
PTXAS = os.path.join(ROOT, "mx-deepim_b200", "build", "ptxas.log")

COLS = [
    ("LDG.128", r"\bLDG\.E\.(?:\w+\.)*128"), ("STG.128", r"\bSTG\.E\.(?:\w+\.)*128"),
    ("LDG", r"\bLDG\."), ("STG", r"\bSTG\."), ("LDGSTS", r"\bLDGSTS"), ("LDS", r"\bLDS"),
    ("STS", r"\bSTS"), ("ATOMS", r"\bATOMS"), ("ATOMG/RED", r"\b(?:ATOMG|REDG?)\."),
    ("SHFL", r"\bSHFL"), ("MATCH/VOTE", r"\b(?:MATCH|VOTE)"), ("REDUX", r"\bREDUX"),
    ("cluster", r"\b(?:UCGABAR|CGABAR|MAPA|ST\.E\.\w*CLUSTER|LD\.E\.\w*CLUSTER|MEMBAR\.\w*\.CTA)|ACQBULK"),
    ("FP64", r"\bD(?:FMA|MUL|ADD|SETP|MNMX)\b"), ("MUFU.RCP64", r"MUFU\.RCP64"),
    # sm_100 packed fp32 pipe (fma.rn.f32x2 and friends)
    ("F32x2", r"\b(?:FFMA2|FMUL2|FADD2)\b"),
    # 64-bit min reductions / atomics in global memory (the rasteriser's z-winner blocks)
    ("REDG/ATOMG.MIN.64", r"\b(?:REDG?|ATOMG)\.E\.MIN\.64"),
]


def demangle(names):
    out = subprocess.run(["cu++filt"] + names, capture_output=True, text=True).stdout.split("\n")
    return dict(zip(names, out))


def strip_params(name):
    """Drop the trailing (parameter list) of a demangled name, keep template arguments."""
    name = name.strip()
    if not name.endswith(")"):
        return name
    depth = 0
    for i in range(len(name) - 1, -1, -1):
        depth += name[i] == ")"
        depth -= name[i] == "("
        if depth == 0:
            return name[:i]
    return name


def main():
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = []
        elif cur and re.match(r"\s+/\*[0-9a-f]{4}\*/", line):
            kernels[cur].append(line)
    res = {}
    if os.path.exists(PTXAS):
        txt = open(PTXAS).read()
        for m in re.finditer(r"Compiling entry function '(\S+)' for 'sm_100a'.*?\n.*?Function properties for \S+\n\s+(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads\n.*?Used (\d+) registers(?:, used \d+ barriers)?(?:, (\d+) bytes smem)?", txt):
            res[m.group(1)] = (m.group(5), m.group(2), m.group(3), m.group(6) or "0")
    names = demangle(list(kernels))
    print("# Static kernel evidence (cuobjdump -sass / ptxas -v of librendercompare.so, sm_100a)\n")
    print("Counts are static instruction counts in the SASS of each kernel (not executed counts).")
    print("`LDGSTS` = cp.async global->shared staging; `cluster` = DSMEM / cluster-barrier "
          "instructions.\n")
    print("| kernel | regs | stack B | spill st B | static smem B | instrs | " +
          " | ".join(c for c, _ in COLS) + " |")
    print("|---|---|---|---|---|---|" + "---|" * len(COLS))
    for k, lines in kernels.items():
        body = "\n".join(lines)
        short = strip_params(names.get(k, k)).replace("void ", "")
        short = short.replace("(anonymous namespace)::", "").replace("(int)", "").replace("(bool)", "")
        r = res.get(k, ("?", "?", "?", "?"))
        counts = [len(re.findall(p, body)) for _, p in COLS]
        print("| `%s` | %s | %s | %s | %s | %d | %s |" % (short, r[0], r[1], r[2], r[3], len(lines),
                                                         " | ".join(str(c) for c in counts)))


if __name__ == "__main__":
    sys.exit(main())
