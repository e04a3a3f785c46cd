#!/usr/bin/env python3
"""Static evidence from the built library (no GPU needed): per-kernel register / stack / shared-memory use
(cuobjdump -res-usage) and counts of the SASS mnemonics that matter for this path (cuobjdump -sass).
    python tools/sass_summary.py r01   ->  profiles/r01_sass_summary.md
"""
import os
import re
import subprocess
import sys
from collections import Counter, OrderedDict

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "lattice-boltzmann_b200", "liblbm_b200.so")
KEYS = ["LDG.E.128", "LDG.E.", "STG.E.128", "STG.E.", "LDGSTS", "LDS", "STS", "SHFL", "FFMA", "FADD", "FMUL", "MUFU",
        "BAR.SYNC", "MEMBAR", "ERRBAR", "CCTL", "NANOSLEEP", "VOTE", "CALL", "LDL", "STL", "ATOM", "RED"]


def demangle(name):
    try:
        return subprocess.run(["cu++filt", name], capture_output=True, text=True, check=True).stdout.strip()
    except Exception:
        return name


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r01"
    res = subprocess.run(["cuobjdump", "-res-usage", LIB], capture_output=True, text=True, check=True).stdout
    usage = {}
    cur = None
    for line in res.splitlines():
        m = re.search(r"Function (\S+):", line)
        if m:
            cur = m.group(1)
        elif cur and "REG:" in line:
            usage[cur] = " ".join(re.findall(r"(?:REG|STACK|SHARED):\d+", line))
            cur = None
    sass = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = OrderedDict()
    cur = None
    for line in sass.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
        if cur and m:
            op = m.group(1)
            kernels[cur]["_total"] += 1
            for k in KEYS:
                if op.startswith(k):
                    kernels[cur][k] += 1
    want = [k for k in kernels if "persistent_kernel" in k or "dataflow_kernel" in k or "lbm_step_kernel" in k]
    lines = [f"# {tag} - static SASS / resource summary of the timestep kernels (tools/sass_summary.py, cuobjdump, sm_100a)", "",
             "Counts are static instructions in the kernel text (not executed counts). LDG.E. / STG.E. include the 128-bit ones.",
             "Strict kernels (first template argument 1) carry no contracted FFMA in the collision itself; their FFMAs are the",
             "exact division / square-root sequences (DESIGN.md section 2).  Template arguments: persistent <strict, ring, profile>,",
             "step <strict>, dataflow <strict, several tasks per warp>.", "",
             "| kernel | resources | instr | " + " | ".join(KEYS) + " |", "|---|---|---|" + "---|" * len(KEYS)]
    for k in want:
        c = kernels[k]
        name = demangle(k)
        name = (name.split(">(")[0] + ">").replace("void lbm::", "").replace("(bool)", "")
        lines.append(f"| `{name}` | {usage.get(k, '?')} | {c['_total']} | " + " | ".join(str(c[x]) for x in KEYS) + " |")
    with open(os.path.join(ROOT, "profiles", f"{tag}_sass_summary.md"), "w") as fh:
        fh.write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
