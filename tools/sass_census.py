#!/usr/bin/env python
"""Static SASS instruction census per kernel of a built library (cuobjdump -sass), the mnemonics that prove what the kernels use:
UBLKCP = bulk (TMA) copy, SYNCS = mbarrier, ELECT = elect.sync, BRA.U.ANY = loop over active threads around a uniform-operand
instruction (none left around the bulk copies, DESIGN.md section 3 rule 9), LDGSTS = cp.async, STL/LDL = local-memory spills.
usage: python tools/sass_census.py mpas-model_b200/libmpas_b200.so > profiles/<round>_sass_census.txt"""
import re
import subprocess
import sys

COLS = ["UBLKCP", "SYNCS", "ELECT", "BRA.U.ANY", "LDGSTS", "LDS", "STS", "LDL", "STL", "DADD", "DMUL", "DFMA", "SHFL", "BAR", "ATOM"]


def main(lib):
    out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True).stdout
    names = subprocess.run(["cu++filt"], input="\n".join(re.findall(r"Function : (\S+)", out)), capture_output=True, text=True).stdout.split("\n")
    print("# cuobjdump -sass %s (sm_100a): static instruction census per kernel (tools/sass_census.py)" % lib)
    print("%-72s %5s " % ("kernel", "instr") + " ".join("%6s" % c[:6] for c in COLS))
    blocks = re.split(r"\n\s*Function : ", out)[1:]
    for blk, name in zip(blocks, names):
        depth, cut = 0, len(name)
        for i in range(len(name) - 1, -1, -1):  # drop the trailing parameter list, keep "(bool)1" inside the template arguments
            if name[i] == ")":
                depth += 1
            elif name[i] == "(":
                depth -= 1
                if depth == 0:
                    cut = i
                    break
        name = name[:cut].replace("<unnamed>::", "").replace("void ", "").replace("(bool)", "").replace("(int)", "")
        ins = re.findall(r"^\s*/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", blk, flags=re.M)
        cnt = {c: 0 for c in COLS}
        for i in ins:
            for c in COLS:
                if i == c or i.startswith(c + ".") or (c == "BRA.U.ANY" and i.startswith("BRA.U") and "ANY" in i):
                    cnt[c] += 1
        # BRA.U.ANY appears as "BRA.U.ANY" mnemonic
        print("%-72s %5d " % (name[:72], len(ins)) + " ".join("%6d" % cnt[c] for c in COLS))


main(sys.argv[1])
