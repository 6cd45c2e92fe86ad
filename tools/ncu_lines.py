#!/usr/bin/env python3
"""Per-source-line executed-instruction counts of one kernel.

    tools/ncu_lines.py REPORT.ncu-rep KERNEL_REGEX MANGLED_PREFIX [divisor]

Joins the SASS page of an ncu report (instructions executed per SASS instruction) with the
line table of the same kernel in the built library (nvdisasm -g), by instruction index, and
prints the totals per source line (divided by `divisor`, e.g. the number of warp-elements).
The library must be the build that was profiled."""
import collections
import csv
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB = os.path.join(ROOT, "torchkafka_b200", "libtorchkafka_b200.so")


def sass_lines(prefix):
    tmp = tempfile.mkdtemp()
    subprocess.run(["cuobjdump", "-xelf", "all", LIB], cwd=tmp, check=True, capture_output=True)
    for f in sorted(os.listdir(tmp)):
        txt = subprocess.run(["nvdisasm", "-g", "-c", f], cwd=tmp, capture_output=True, text=True).stdout
        if ".text." + prefix not in txt:
            continue
        out, cur, on = [], None, False
        for l in txt.splitlines():
            if l.startswith(".text."):
                on = l.startswith(".text." + prefix)
                continue
            if not on:
                continue
            m = re.search(r'//## File "([^"]+)", line (\d+)', l)
            if m:
                cur = (os.path.basename(m.group(1)), int(m.group(2)))
                continue
            m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
            if m:
                out.append((cur, m.group(2).strip()))
        return out
    raise SystemExit("kernel not found in " + LIB)


def main():
    rep, regex, prefix = sys.argv[1:4]
    div = float(sys.argv[4]) if len(sys.argv) > 4 else 1.0
    txt = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-name", "regex:" + regex],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    starts = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"]
    h = rows[starts[0] + 1]
    ci, si = h.index("Instructions Executed"), h.index("Source")
    end = starts[1] if len(starts) > 1 else len(rows)
    body = [r for r in rows[starts[0] + 2:end] if len(r) > ci]
    sass = sass_lines(prefix)
    if len(body) != len(sass):
        raise SystemExit(f"instruction count differs: report {len(body)}, library {len(sass)}")
    by, ops = collections.Counter(), collections.Counter()
    for b, s in zip(body, sass):
        n = int(b[ci])
        by[s[0]] += n
        ops[[o for o in s[1].split() if not o.startswith("@")][0].split(".")[0]] += n
    tot = sum(by.values())
    print("total", tot, round(tot / div, 1))
    print("-- opcodes")
    for k, v in ops.most_common(16):
        print(f"{k:8s} {v / div:10.1f}")
    print("-- lines")
    for k, v in sorted(by.items(), key=lambda kv: -kv[1])[:70]:
        print(f"{k[0]}:{k[1]:<5d} {v / div:10.1f}")


if __name__ == "__main__":
    main()
