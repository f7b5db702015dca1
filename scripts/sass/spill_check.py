"""Reports local-memory (spill) accesses inside the loops of a kernel in a `cuobjdump -sass` dump, per loop
(backward branches): usage: spill_check.py file.sass <substring of the kernel name>"""
import re, sys
txt = open(sys.argv[1]).read()
for f in re.split(r'\n\s*Function : ', txt)[1:]:
    if sys.argv[2] not in f.split('\n')[0]: continue
    rows = [(int(m.group(1), 16), m.group(2).strip()) for m in re.finditer(r'/\*([0-9a-f]{4,5})\*/\s+(.*?);', f)]
    loops = []
    for a, t in rows:
        m = re.search(r'BRA(?:\.\w+)* (?:\w+, )?0x([0-9a-f]+)', t)
        if m and int(m.group(1), 16) < a and a - int(m.group(1), 16) > 0x800: loops.append((int(m.group(1), 16), a))
    for lo, hi in loops:
        sp = [(a, t) for a, t in rows if lo <= a <= hi and re.search(r'\b(LDL|STL)', t)]
        print('loop %x..%x (%d instrs): %d local-memory accesses' % (lo, hi, (hi - lo) // 16 + 1, len(sp)))
        for a, t in sp: print('   %x %s' % (a, t))
