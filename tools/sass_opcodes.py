#!/usr/bin/env python
"""Lists, per kernel of libigs.so, the SASS opcodes that show which Blackwell units it uses: tcgen05 MMAs
(UTC*MMA), tensor-memory loads / stores (LDTM / STTM), TMA tensor and bulk copies (UTMALDG / UTMASTG / UBLKCP),
mbarrier waits (SYNCS), system-scope atomics (peer-memory windows), warp match (MATCH), 256-bit global moves,
and the `.2CTA` forms of the CTA-pair MLP kernels.

    python tools/sass_opcodes.py [> profiles/r2_sass_opcodes.txt]

Needs only cuobjdump (no GPU).  Regenerate after every kernel change."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "igaming-platform_b200", "libigs.so")
PAT = re.compile(r"\b(UTC[A-Z]*MMA[.A-Z0-9_]*|LDG\.E[.A-Z0-9_]*\.256[.A-Z0-9_]*|STG\.E[.A-Z0-9_]*\.256[.A-Z0-9_]*|UTCBAR[.A-Z0-9_]*|LDTM[.A-Za-z0-9_]*|STTM[.A-Za-z0-9_]*|UTMALDG[.A-Z0-9_]*|UTMASTG[.A-Z0-9_]*|"
                 r"UBLKCP[.A-Z0-9_]*|SYNCS[.A-Z0-9_]*|MATCH[.A-Z0-9_]*|ATOM[A-Z]*\.[.A-Z0-9_]*SYS[.A-Z0-9_]*|"
                 r"RED[.A-Z0-9_]*SYS[.A-Z0-9_]*|HMMA[.A-Z0-9_]*|REDUX[.A-Z0-9_]*)")


def main():
    out = subprocess.run(["cuobjdump", "-sass", LIB], capture_output=True, text=True, check=True).stdout
    kernels = collections.OrderedDict()
    cur = None
    for line in out.splitlines():
        m = re.match(r"\s*Function : (\S+)", line)
        if m:
            cur = m.group(1)
            kernels[cur] = collections.Counter()
            continue
        if cur is None:
            continue
        for op in PAT.findall(line):
            kernels[cur][op.split(".")[0] if op.startswith(("SYNCS", "MATCH", "REDUX")) else op] += 1
    demangled = subprocess.run(["c++filt"], input="\n".join(kernels), capture_output=True, text=True).stdout.splitlines()
    sha = subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True, text=True).stdout.strip()
    print(f"# SASS opcode census of igaming-platform_b200/libigs.so (sm_100a), source tree at {sha}")
    print("# cuobjdump -sass | per kernel: opcode x count   (tools/sass_opcodes.py)")
    total = collections.Counter()
    for name, dm in zip(kernels, demangled):
        c = kernels[name]
        total.update(c)
        short = re.sub(r"\(anonymous namespace\)::", "", dm)
        short = short.split("(")[0] if "<" not in short else short[:short.rfind("(")] if "(" in short else short
        ops = ", ".join(f"{k} x{v}" for k, v in sorted(c.items())) or "-"
        print(f"{short[:110]:110s} {ops}")
    print("# total: " + ", ".join(f"{k} x{v}" for k, v in sorted(total.items())))


if __name__ == "__main__":
    main()
