#!/usr/bin/env python3
"""Dynamic SASS opcode histogram (warp-level instructions executed per warp) from an .ncu-rep source page.
Usage: python profiles/ncu_opcodes.py report.ncu-rep [launch-index]"""
import collections
import csv
import io
import re
import subprocess
import sys


def main(path, idx='1', top='22'):
    out = subprocess.run(['ncu', '-i', path, '--page', 'source', '--csv', '--kernel-id', ':::' + idx],
                         capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ci, si, sa = hdr.index('Instructions Executed'), hdr.index('Source'), hdr.index('# Samples')
    by, samp = collections.Counter(), collections.Counter()
    first = None
    seen = set()
    for r in rows[2:]:
        if len(r) <= ci or r[0] in seen:      # the page lists every instruction once per view; keep the first
            continue
        seen.add(r[0])
        try:
            n, s = int(r[ci]), int(r[sa])
        except ValueError:
            continue
        if first is None:
            first = n
        m = re.match(r'\s*(@!?U?P\d+\s+)?([A-Z0-9_]+)', r[si])
        op = m.group(2) if m else r[si]
        by[op] += n
        samp[op] += s
    tot = sum(by.values())
    print(rows[0][1][:90])
    print('warps %d, instructions per warp %.0f' % (first, tot / first))
    for op, n in by.most_common(int(top)):
        print('  %-8s %7.1f /warp   %5.1f%% of stall samples' % (op, n / first, 100.0 * samp[op] / max(1, sum(samp.values()))))


if __name__ == '__main__':
    main(*sys.argv[1:])
