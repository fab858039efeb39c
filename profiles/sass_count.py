#!/usr/bin/env python3
"""Static opcode histogram of one kernel from an nvdisasm listing: python profiles/sass_count.py all.sass <substring of the mangled name>"""
import collections, re, sys
lines = open(sys.argv[1]).read().split('\n')
pat = sys.argv[2]
for i, l in enumerate(lines):
    if l.startswith('.text.') and pat in l:
        c = collections.Counter()
        for m in lines[i + 1:]:
            if m.startswith('.text.') or m.strip().startswith('.section'):
                break
            mm = re.match(r'\s+/\*[0-9a-f]+\*/\s+(@!?U?P\w+\s+)?([A-Z0-9_.]+)', m)
            if mm:
                c[mm.group(2).split('.')[0]] += 1
        tot = sum(c.values())
        fp = sum(c[k] for k in ('DFMA', 'DMUL', 'DADD', 'DSETP', 'MUFU'))
        print(l[:120]); print(' total %d fp64 %d :' % (tot, fp), ' '.join('%s %d' % kv for kv in c.most_common(14)))
