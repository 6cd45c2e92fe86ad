#!/usr/bin/env python
"""Instruction mnemonics per kernel of libtorchkafka_b200.so -> profiles/rNN_sass_evidence.txt.

    python tools/sass_evidence.py > profiles/r02_sass_evidence.txt

Extracts the sm_100a cubins (cuobjdump -xelf), disassembles them (nvdisasm -c), demangles the kernel
names (c++filt) and counts, per kernel, the mnemonics that show what the hardware is asked to do.
Runs without a GPU."""
import collections
import glob
import os
import re
import subprocess
import sys
import tempfile

HERE = os.path.dirname(os.path.abspath(__file__))
LIB = os.path.join(HERE, "..", "torchkafka_b200", "libtorchkafka_b200.so")
SHOW = ("ACQBULK", "ATOMG", "CCTL.E.PF1", "DMUL", "F2F.F32.F64", "I2F.F64", "IDP.4A", "LDG.E.128", "LDS.128",
        "MATCH", "PRMT", "REDUX", "SHFL", "STG.E.128", "SYNCS", "UBLKCP", "VOTE", "LDL", "STL")


def key_of(op: str):
    for k in SHOW:
        if op == k or op.startswith(k + "."):
            return k
    return None


def main():
    lib = os.path.abspath(sys.argv[1] if len(sys.argv) > 1 else LIB)
    print("# SASS evidence: instruction mnemonics per kernel of libtorchkafka_b200.so (sm_100a)")
    print("# python tools/sass_evidence.py  (cuobjdump -xelf all; nvdisasm -c; c++filt)")
    print("# UBLKCP = cp.async.bulk (TMA bulk copy), SYNCS = mbarrier, ACQBULK = griddepcontrol (programmatic"
          " dependent launch), IDP.4A = byte dot product, LDL / STL = local memory")
    print("# A kernel's section holds its out-of-line callees too (the scalar JSON grammar, Eisel-Lemire): their"
          " register saves and local arrays are the LDL / STL counted here; `-Xptxas -v` has the kernels' own spills")
    print()
    with tempfile.TemporaryDirectory() as tmp:
        subprocess.run(["cuobjdump", "-xelf", "all", lib], cwd=tmp, check=True, capture_output=True)
        for cubin in sorted(glob.glob(os.path.join(tmp, "*.cubin"))):
            text = subprocess.run(["nvdisasm", "-c", cubin], check=True, capture_output=True, text=True).stdout
            name, counts, total = None, None, 0
            rows = []
            for line in text.splitlines():
                m = re.match(r"\s*\.text\.(\S+):", line)
                if m:
                    if name:
                        rows.append((name, total, counts))
                    name, counts, total = m.group(1), collections.Counter(), 0
                    continue
                m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P[T\d]+\s+)?([A-Z][A-Z0-9_.]+)\b.*;", line)
                if m and name:
                    total += 1
                    k = key_of(m.group(1))
                    if k:
                        counts[k] += 1
            if name:
                rows.append((name, total, counts))
            names = subprocess.run(["c++filt"], input="\n".join(r[0] for r in rows), capture_output=True,
                                   text=True).stdout.splitlines()
            for (raw, total, counts), nm in zip(rows, names):
                if total == 0:
                    continue
                what = ", ".join("%s %d" % (k, counts[k]) for k in sorted(counts))
                print("%-88s %6d instr  %s" % (nm[:88], total, what))


if __name__ == "__main__":
    main()
