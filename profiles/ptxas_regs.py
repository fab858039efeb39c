"""Registers / spills of the headline stage-kernel instantiations from the ptxas logs (make writes *.ptxas.log)."""
import re, sys, glob, os
here = os.path.join(os.path.dirname(os.path.abspath(__file__)), '..', 'fem_b200', 'csrc')
for f in sorted(glob.glob(os.path.join(here, 'dg1d_euler.ptxas.log')) + glob.glob(os.path.join(here, 'dg1d_burger.ptxas.log'))):
    log = open(f).read()
    for e in re.split(r"ptxas info\s+: Compiling entry function ", log)[1:]:
        name = e.split("'")[1]
        m = re.search(r"stage_kernelILi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELi(\d+)ELb(\d)EL[bi](\d)", name)
        if not m:
            continue
        pde, flux, K, LIM, VAR, BLOCK, MINB, C, PARK, CLOSED = map(int, m.groups())
        want = (pde == 2 and VAR != 2 and CLOSED in (1, 3) and ((K == 2 and LIM == 2 and flux in (2, 4)) or (K == 1 and LIM == 1 and flux == 4))) or \
               (pde == 1 and VAR != 2 and K == 2 and LIM == 1 and flux == 2 and CLOSED in (0, 2))
        if not want:
            continue
        regs = re.search(r"Used (\d+) registers", e).group(1)
        sp = re.search(r"(\d+) bytes stack frame, (\d+) bytes spill stores, (\d+) bytes spill loads", e).groups()
        print("pde %d flux %d K %d LIM %d VAR %d edge %d block %d minb %d C %d: regs %s stack/spill-st/spill-ld %s" %
              (pde, flux, K, LIM, VAR, CLOSED, BLOCK, MINB, C, regs, "/".join(sp)))
