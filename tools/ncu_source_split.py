#!/usr/bin/env python3
"""Per-function share of a kernel's executed instructions and stall samples from the correlated
source page of an .ncu-rep:
    ncu -i X.ncu-rep --page source --print-source cuda,sass --csv --kernel-name regex:K > page.csv
    tools/ncu_source_split.py page.csv path/to/source.cu [top_lines]
Functions are taken as the spans between top-level `static __device__` / `__global__` / `template`
lines of the source file."""
import csv
import re
import sys


def main():
    page, src_path = sys.argv[1], sys.argv[2]
    top = int(sys.argv[3]) if len(sys.argv) > 3 else 12
    rows = list(csv.reader(open(page)))
    hi = next(i for i, r in enumerate(rows) if r and r[0] == "Line No")
    hdr = rows[hi]
    ix, sx = hdr.index("Instructions Executed"), hdr.index("# Samples")
    data, tot, tots = [], 0, 0
    for r in rows[hi + 1:]:
        if len(r) <= ix or r[2] != "-":  # source-level rows carry '-' in the address column
            continue
        try:
            n, s = int(r[ix]), int(r[sx])
        except ValueError:
            continue
        tot += n
        tots += s
        data.append((n, s, int(r[0]), r[1]))
    src = open(src_path).read().split("\n")
    starts = [i for i, l in enumerate(src, 1)
              if re.match(r"^(static )?(__device__|__global__|template)", l) and ("(" in l or "template" in l)]
    agg = {}
    for n, s, line, _ in data:
        k = [f for f in starts if f <= line]
        f = k[-1] if k else 0
        a = agg.setdefault(f, [0, 0])
        a[0] += n
        a[1] += s
    print("instructions %d, samples %d" % (tot, tots))
    for f, (n, s) in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
        print("%5.1f %% instr %5.1f %% samples  L%-4d %s" % (100.0 * n / tot, 100.0 * s / max(tots, 1), f,
                                                            src[f - 1].strip()[:90] if f else ""))
    print()
    for n, s, line, text in sorted(data, reverse=True)[:top]:
        print("%5.1f %% instr %5.1f %% samples  L%-4d %s" % (100.0 * n / tot, 100.0 * s / max(tots, 1), line,
                                                            text.strip()[:100]))


if __name__ == "__main__":
    main()
