"""Per-source-line opcode classes of an address range of an `nvdisasm --print-line-info` listing.
usage: line_stat.py file.dis lo hi [source file substring]  (hex addresses of the loop, from loop_stat.py)"""
import re, sys, collections
lo, hi = int(sys.argv[2], 16), int(sys.argv[3], 16)
want = sys.argv[4] if len(sys.argv) > 4 else ""
cur = ("?", 0)
tab = collections.defaultdict(collections.Counter)
for line in open(sys.argv[1]):
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);', line)
    if not m: continue
    a = int(m.group(1), 16)
    if not (lo <= a <= hi): continue
    op = re.sub(r'^@!?U?P\d+\s+', '', m.group(2).strip()).split()[0]
    base = op.split('.')[0]
    cls = 'FP64' if base in ('DADD', 'DFMA', 'DMUL') else 'LDS' if base == 'LDS' else 'TMEM' if base in ('LDTM', 'STTM') else 'MOV' if base in ('MOV', 'IMAD') and 'MOV' in op else base
    tab[cur][cls] += 1
tot = collections.Counter()
for k in sorted(tab):
    if want and want not in k[0]: continue
    c = tab[k]; tot.update(c)
    other = sum(v for kk, v in c.items() if kk not in ('FP64', 'LDS'))
    print('%-20s %4d  n=%4d FP64 %3d LDS %3d other %3d  %s' % (k[0], k[1], sum(c.values()), c['FP64'], c['LDS'], other, ' '.join('%s:%d' % kv for kv in c.most_common() if kv[0] not in ('FP64', 'LDS'))))
print('total', sum(tot.values()), dict(tot.most_common()))
