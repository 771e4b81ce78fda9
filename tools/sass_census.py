#!/usr/bin/env python
"""SASS mnemonic census of the built library (no GPU needed): which kernels carry tcgen05 / TMA / TMEM /
mbarrier instructions and which still use legacy mma.sync.  Writes a markdown table.

    python tools/sass_census.py profiles/r01_head_sass_census.md
"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "nano-vllm-candle_b200", "libnvc_b200.so")
MNEMONICS = ["UTCHMMA", "UTCBAR", "UTMALDG", "UTMAPF", "UTCATOMSWS", "LDTM", "STTM", "SYNCS", "ACQBULK", "UCGABAR",
             "HMMA", "LDG", "STG", "LDS", "STS", "BAR", "ATOM", "RED", "SHFL", "MUFU"]
HEADER = """# SASS mnemonic census of libnvc_b200.so (`cuobjdump -sass`, sm_100a; made on the build host, no GPU needed)

Counts of instructions whose opcode starts with the mnemonic: `UTCHMMA` = tcgen05.mma, `UTCBAR` = tcgen05.commit,
`UTMALDG` = TMA tensor load, `UTMAPF` = TMA L2 prefetch, `UTCATOMSWS` = TMEM alloc/dealloc, `LDTM`/`STTM` = tcgen05.ld/st,
`SYNCS` = mbarrier ops, `UCGABAR` = cluster barrier, `ACQBULK` = griddepcontrol.wait, `HMMA` = legacy mma.sync
(none of the GEMM / prefill kernels use it; the paged decode kernel is the one deliberate mma.sync user, DESIGN.md 4.3).
`regs / stack / static smem` is `cuobjdump -res-usage` (dynamic shared memory is not listed).  No kernel spills: the 8-24 B
stack frames of the tcgen05 and peer-memory kernels are the argument buffer of the cold-path `printf` that reports an
mbarrier / peer-flag wait that timed out before trapping (ptx.cuh `mbar_wait`, p2p_allreduce.cu); their only STL
instructions sit in those blocks.
"""


def demangle(name):
    d = subprocess.run(["c++filt", "-p", name], capture_output=True, text=True).stdout.strip()
    return d.replace("(anonymous namespace)::", "").replace("void ", "")


def main():
    out_path = sys.argv[1] if len(sys.argv) > 1 else "/dev/stdout"
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    per, fn = collections.OrderedDict(), None
    for line in sass.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            fn = m.group(1)
            per[fn] = collections.Counter()
            continue
        if fn is None:
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_]+)", line)
        if m:
            per[fn]["_total"] += 1
            for k in MNEMONICS:
                if m.group(1).startswith(k):
                    per[fn][k] += 1
                    break
    # registers / stack (spills) / static shared memory per kernel, from the same cubin
    res, fn = {}, None
    for line in subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout.splitlines():
        m = re.match(r"\s*Function (\S+):", line)
        if m:
            fn = m.group(1)
            continue
        m = re.search(r"REG:(\d+) STACK:(\d+) SHARED:(\d+) LOCAL:(\d+)", line)
        if m and fn:
            res[fn] = "%s / %s / %s" % (m.group(1), m.group(2), m.group(3))
    rows = [HEADER, "| kernel | regs / stack B / static smem B | instrs | " + " | ".join(MNEMONICS) + " |",
            "|---|---|---|" + "---|" * len(MNEMONICS)]
    for f, c in sorted(per.items(), key=lambda kv: demangle(kv[0])):
        rows.append("| `%s` | %s | %d | " % (demangle(f)[:110], res.get(f, "?"), c["_total"])
                    + " | ".join(str(c[k]) if c[k] else "" for k in MNEMONICS) + " |")
    with open(out_path, "w") as fh:
        fh.write("\n".join(rows) + "\n")


if __name__ == "__main__":
    main()
