"""Instruction histogram of the main loop of a kernel in a `cuobjdump -sass` dump.
usage: loop_stat.py file.sass <substring of the mangled kernel name> [lo hi]
The loop is the widest backward branch unless the hex addresses lo hi are given (variants 5-7 run their chunk prologue
inside the first iteration, variant 8's outermost backward branch is an mbarrier retry: pass the range of the z loop,
from its BAR.SYNC/BAR.ARV pair).  Also prints the sum of the stall-count fields of the control words."""
import re,collections,sys
txt=open(sys.argv[1]).read()
funcs=re.split(r'\n\s*Function : ',txt)
for f in funcs[1:]:
    name=f.split('\n')[0]
    if sys.argv[2] not in name: continue
    lines=f.split('\n')
    rows=[]
    i=0
    while i<len(lines):
        m=re.match(r'\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);\s*/\* 0x([0-9a-f]{16}) \*/',lines[i])
        if m and i+1<len(lines):
            m2=re.match(r'\s+/\* 0x([0-9a-f]{16}) \*/',lines[i+1])
            if m2:
                hi=int(m2.group(1),16)
                rows.append((int(m.group(1),16),m.group(2).strip(),(hi>>41)&0xf)); i+=2; continue
        i+=1
    # main loop = last backward BRA with largest span
    best=None
    for a,t,s in rows:
        m=re.search(r'BRA(?:\.\w+)* (?:\w+, )?0x([0-9a-f]+)',t)
        if m:
            tgt=int(m.group(1),16)
            if tgt<a and (best is None or a-tgt>best[1]-best[0]): best=(tgt,a)
    lo,hi=best
    if len(sys.argv)>=5: lo,hi=int(sys.argv[3],16),int(sys.argv[4],16)
    # exclude inner regions: blocks that are skipped by forward branches over >100 instrs whose predicate is j0==0 or tid<32 can't be detected; report whole loop + histogram
    h=collections.Counter(); st=0; n=0
    for a,t,s in rows:
        if lo<=a<=hi:
            op=re.sub(r'^@!?U?P\d+\s+','',t).split()[0]
            h[op]+=1; st+=s; n+=1
    print(name); print(' loop %x..%x static instrs %d, stall sum %d'%(lo,hi,n,st))
    fp=h['DADD']+h['DFMA']+h['DMUL']
    print(' FP64',fp,'(DADD %d DFMA %d DMUL %d)'%(h['DADD'],h['DFMA'],h['DMUL']),'LDS',h['LDS.64'],'other',n-fp-h['LDS.64'])
    print(' ',', '.join('%s %d'%(k,v) for k,v in h.most_common(30) if k not in('DADD','DFMA','DMUL','LDS.64')))
