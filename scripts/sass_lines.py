#!/usr/bin/env python
"""Static SASS instruction count per source line / inlined function of ONE kernel (nvdisasm -g output of a cubin built
with -lineinfo).  The step kernels are almost straight-line code, so the static count is close to the executed count
and shows where an env step's instructions go (rules, Philox, writer, address arithmetic).

usage: sass_lines.py <cubin> <mangled kernel name substring> [--lines]"""
import collections
import re
import subprocess
import sys


def main():
    cubin, key = sys.argv[1], sys.argv[2]
    text = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.splitlines()
    start = next(i for i, l in enumerate(text) if l.startswith(".text.") and key in l)
    per_line, per_op = collections.Counter(), collections.Counter()
    cur = ("?", 0)
    total = 0
    for l in text[start + 1:]:
        if l.startswith("\t.section") or l.startswith(".text."):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).rsplit("/", 1)[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*[0-9a-f]{4,}\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", l)
        if m:
            op = m.group(2).split(".")[0]
            if op in ("NOP",):
                continue
            per_line[cur] += 1
            per_op[op] += 1
            total += 1
    print("static instructions:", total)
    print("by opcode:", ", ".join("%s %d" % kv for kv in per_op.most_common(24)))
    if "--lines" in sys.argv:
        for (f, n), c in sorted(per_line.items()):
            print("%-20s %5d  %d" % (f, n, c))
    else:                                   # bucket by 'function' = nearest preceding definition line in the source
        by_file = collections.defaultdict(list)
        for (f, n), c in per_line.items():
            by_file[f].append((n, c))
        import os
        root = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "gymothelloenv_b200", "csrc")
        out = collections.Counter()
        for f, items in by_file.items():
            try:
                src = open(os.path.join(root, f)).read().splitlines()
            except OSError:
                out[(f, "?")] += sum(c for _, c in items)
                continue
            for n, c in items:
                name = "?"
                for k in range(min(n, len(src)) - 1, -1, -1):
                    m = re.match(r"^(?:OTH_DEV|OTH_HD|__device__|__global__|template|static|inline)?.*?\b([A-Za-z_][A-Za-z0-9_]*)\s*\([^;]*$", src[k])
                    if m and not src[k].startswith((" ", "\t", "#", "//")) and m.group(1) not in ("if", "for", "while"):
                        name = m.group(1)
                        break
                out[(f, name)] += c
        for (f, name), c in out.most_common():
            print("%5d  %s:%s" % (c, f, name))


if __name__ == "__main__":
    main()
