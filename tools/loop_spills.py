"""Per-loop count of local-memory (spill) instructions in the SASS of one kernel:  python tools/loop_spills.py <lib.so> <mangled kernel name>"""
import re,sys,subprocess
lib,fun=sys.argv[1],sys.argv[2]
out=subprocess.run(['cuobjdump','-sass','-fun',fun,lib],capture_output=True,text=True).stdout
lines=[]
for l in out.splitlines():
    m=re.match(r'^\s+/\*([0-9a-f]{4,5})\*/\s+(.*?);',l)
    if m: lines.append((int(m.group(1),16),m.group(2)))
loops=[]
for a,t in lines:
    m=re.search(r'BRA(?:\.U)? (?:!?U?P\d, )?0x([0-9a-f]+)',t)
    if m and int(m.group(1),16)<a: loops.append((int(m.group(1),16),a))
print('insts',len(lines),'local',sum(1 for a,t in lines if re.search(r'\b(STL|LDL)',t)))
for t0,s0 in sorted(loops):
    body=[(a,t) for a,t in lines if t0<=a<=s0]
    loc=sum(1 for a,t in body if re.search(r'\b(STL|LDL)',t))
    kinds=set()
    for a,t in body:
        for k in ('UBLKCP','SYNCS','SHFL','REDG','DFMA','BAR','ATOMS','LDG','MEMBAR'):
            if k in t: kinds.add(k)
    print(hex(t0),hex(s0),'insts',len(body),'local',loc,sorted(kinds))
