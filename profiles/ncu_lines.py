#!/usr/bin/env python3
"""Dynamic warp-instruction count per SOURCE LINE: joins the per-instruction 'Instructions Executed' of an .ncu-rep
(SASS source page) with the line table of the same kernel disassembled from the library (nvdisasm -g).
Usage: python profiles/ncu_lines.py report.ncu-rep launch-index all.sass mangled-kernel-name [top]
(all.sass: cuobjdump -xelf <tu> libdg1d.so; nvdisasm -g <cubin> > all.sass).  The two instruction streams must be the
same build: the script checks opcode by opcode."""
import collections, csv, io, re, subprocess, sys


def main(rep, idx, sass, name, top='40'):
    out = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv', '--kernel-id', ':::' + idx], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[1]
    ci, si = hdr.index('Instructions Executed'), hdr.index('Source')
    dyn, seen = [], set()
    for r in rows[2:]:
        if len(r) <= ci or r[0] in seen:
            continue
        seen.add(r[0])
        if not r[ci].isdigit():
            continue
        m = re.match(r'\s*(@!?U?P\w+\s+)?([A-Z0-9_.]+)', r[si])
        dyn.append((m.group(2).split('.')[0] if m else '?', int(r[ci])))
    lines = open(sass).read().split('\n')
    start = lines.index('.text.' + name + ':')
    cur, stat = None, []
    for l in lines[start + 1:]:
        if l.startswith('.text.') or l.strip().startswith('.section'):
            break
        m = re.search(r'//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split('/')[-1], int(m.group(2)))
            continue
        m = re.match(r'\s+/\*[0-9a-f]+\*/\s+(@!?U?P\w+\s+)?([A-Z0-9_.]+)', l)
        if m:
            stat.append((m.group(2).split('.')[0], cur))
    n = min(len(dyn), len(stat))
    bad = sum(1 for i in range(n) if dyn[i][0] != stat[i][0])
    print('instructions: report %d, disassembly %d, opcode mismatches %d' % (len(dyn), len(stat), bad))
    by = collections.Counter(); fp = collections.Counter()
    warps = dyn[0][1]
    for i in range(n):
        by[stat[i][1]] += dyn[i][1]
        if dyn[i][0] in ('DFMA', 'DMUL', 'DADD', 'DSETP', 'MUFU'):
            fp[stat[i][1]] += dyn[i][1]
    tot = sum(by.values())
    print('warps %d, instructions per warp %.0f' % (warps, tot / warps))
    src = {}
    for (f, ln), c in by.most_common(int(top)):
        if f not in src:
            try:
                src[f] = open('/root/repo/fem_b200/csrc/' + f).read().split('\n')
            except OSError:
                src[f] = []
        text = src[f][ln - 1].strip()[:90] if ln - 1 < len(src[f]) else ''
        print('%6.1f /warp (fp64 %5.1f)  %s:%d  %s' % (c / warps, fp[(f, ln)] / warps, f, ln, text))


if __name__ == '__main__':
    main(*sys.argv[1:])
