#!/usr/bin/env python3
"""Instruction histogram of one kernel of a .so / .cubin (cuobjdump -sass), optionally between two BAR.SYNC lines.
usage: python tools/sass_hist.py <file> <kernel-substring> [--dump out.sass]"""
import collections
import re
import subprocess
import sys


def main():
    path, key = sys.argv[1], sys.argv[2]
    txt = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
    for p in re.split(r"\n\s*Function : ", txt)[1:]:
        name = p.split("\n")[0]
        if key not in name:
            continue
        lines = [l for l in p.split("\n") if re.match(r"\s+/\*[0-9a-f]{4,6}\*/", l)]
        ops = collections.Counter()
        for l in lines:
            m = re.match(r"\s+/\*[0-9a-f]{4,6}\*/\s+(@!?U?P\d\s+)?([A-Z0-9_.]+)", l)
            ops[m.group(2).split(".")[0]] += 1
        print(name, len(lines), "instructions")
        print(", ".join(f"{k}:{v}" for k, v in ops.most_common(45)))
        if "--dump" in sys.argv:
            open(sys.argv[sys.argv.index("--dump") + 1], "w").write("\n".join(lines))


if __name__ == "__main__":
    main()
